// Device side of the entropy back end (stream scatter, DC delta, RLE, histogram, bit packing) — declarations.
// Replaces the data passes of
//   compressed::encodeImageFast   stream scatter                 CompressionLib/src/CompressedImage.cpp:612-629
//   compressed::writeCompressed   DC delta, RLE choice, ordering  CompressionLib/src/CompressedImage.cpp:487-517
//   huffman::runLengthEncode                                      CompressionLib/src/Huffman.cpp:246-279
//   huffman::huffmanEncode        frequency pass + payload codes  CompressionLib/src/Huffman.cpp:47-64,157-162
//   bitbuffer::writeGolombCode / golombCodeLength                 CompressionLib/src/BitBuffer.cpp:228-269
//   bitbuffer::BitBuffer::WriteBits                               CompressionLib/src/BitBuffer.cpp:78-110
#pragma once

#include <cuda_runtime.h>

#include <cstdint>

#include "mnenc_internal.h"

namespace mnenc {

constexpr int kTileBlocks = 256;   // image blocks per scatter tile
constexpr int kHistBins = 65536;   // u16 symbols
constexpr int kSmemHistBins = 4096;
constexpr uint32_t kNoPos = 0x7FFFFFFFu;  // "symbol not seen": first-occurrence sentinel (positive as int32, so that a signed MIN
                                          // all-reduce over the ranks of a stripe-sharded image is also the unsigned minimum)

// Per coded sequence (0 = `lengths`, 1 + s = symbol stream s), written by decideKernel.
struct SeqDesc {
    uint32_t len;        // symbols to code
    uint32_t rawLen;     // symbols before RLE (stream size)
    uint32_t off;        // offset into the source array (u16 units)
    uint32_t src;        // 0 = counts array, 1 = stream arena, 2 = RLE arena
};

struct SeqTable {                 // device + pinned host mirror, filled by decideKernel
    SeqDesc desc[kMaxSeq];
    uint32_t start[kMaxSeq + 1];  // exclusive prefix of desc[].len; start[nSeq] = M
    uint32_t posBase[kMaxSeq];    // position of the slice's first element in the WHOLE sequence (0 unless stripe-sharded)
    uint32_t nRuns;               // diagnostic
};

// Per sequence coding plan uploaded by the host after the table build.
struct SeqCode {
    uint64_t bitBase;   // absolute bit position of the sequence's first payload code minus the scan value at
                        // its first symbol (so position = bitBase + exclusive scan of code lengths)
    uint32_t golombM;   // 0 = Huffman
    uint32_t pad;
};
struct CodeTable {
    SeqCode seq[kMaxSeq];
};

struct HistEntry {
    uint32_t seq;
    uint32_t symbol;
    uint32_t freq;
    uint32_t firstPos;
};

struct DcPrev {        // DC value preceding the slice's first entry of each colour channel (0 at the image's left edge)
    uint32_t v[3];
};
struct StripeStitch {  // device table of the stripe path: stream offsets + run stitching of this rank's slices
    uint32_t off[6 * kMaxK + 1];
    uint32_t carry[6 * kMaxK];
    uint32_t ext[6 * kMaxK];
};

struct StreamOffsets {
    uint32_t off[6 * kMaxK + 1];  // start of stream s in the arena; off[6K] = N
};

struct BackendBuffers {
    // inputs from the matching-pursuit kernel
    const uint16_t* counts;   // [3*nblocks]
    const uint32_t* pairs;    // [3*nblocks][K]
    int K;
    uint32_t nblocks;
    uint32_t ntiles;
    // scatter
    uint32_t* tileCounts;     // [ntiles][3*K] -> exclusive offsets after tileScan
    uint32_t* totals;         // [3*K]
    uint16_t* arena;          // [N + slack] raw streams, DC streams replaced in place
    uint16_t* dcTmp;          // [3*nblocks]
    // RLE
    uint8_t* flags;           // [N]
    uint32_t* runIdx;         // [N+1] exclusive scan of flags
    uint32_t* startPos;       // [N+1] position of run r; startPos[nRuns] = N
    uint32_t* outOff;         // [N+1] exclusive scan of RLE output sizes per run
    uint16_t* rleArena;       // [N + N/2 + 8]
    uint32_t* blockSums32;    // scan scratch
    unsigned long long* blockSums64;
    // sequences
    SeqTable* seqTable;       // device
    // histogram
    uint32_t* hist;           // [kMaxSeq][65536], all zero between encodes
    uint32_t* firstPos;       // [kMaxSeq][65536], all kNoPos between encodes; holds the entry index after compaction
    uint32_t* maxSym;         // [kMaxSeq]
    HistEntry* entries;       // [entryCap]
    uint32_t* nEntries;
    uint32_t entryCap;
    // packing
    const unsigned long long* entryCodes;  // [nEntries] code | len << 32
    const CodeTable* codeTable;
    uint32_t* outWords;       // zeroed, big-endian bit stream as bytes
};

// All launch wrappers return the number of kernels they launched (for gpu_launches accounting).
int launchTileCounts(const BackendBuffers& b, cudaStream_t st);
int launchScatter(const BackendBuffers& b, const StreamOffsets& so, cudaStream_t st);      // raw scatter + DC delta
int launchScatterRaw(const BackendBuffers& b, const StreamOffsets& so, cudaStream_t st);
int launchDcDelta(const BackendBuffers& b, const StreamOffsets& so, const DcPrev& prev, cudaStream_t st);
int launchRleRuns(const BackendBuffers& b, const StreamOffsets& so, uint32_t N, cudaStream_t st);
int launchStripeRecords(const BackendBuffers& b, const StreamOffsets& so, uint32_t N, StripeBoundary* rec, cudaStream_t st);
int launchRleWrite(const BackendBuffers& b, uint32_t N, const StripeStitch* stitch, cudaStream_t st);
int launchHistogramOnly(const BackendBuffers& b, uint32_t upperBoundM, int nSeq, cudaStream_t st);
int launchCompact(const BackendBuffers& b, int nSeq, cudaStream_t st);
int launchSeqBits(const BackendBuffers& b, uint32_t M, int nSeq, unsigned long long* seqBits, cudaStream_t st);
int launchFillU32(uint32_t* p, size_t n, uint32_t value, cudaStream_t st);
int launchReadIndirect(const uint32_t* arr, const uint32_t* idx, uint32_t* out, cudaStream_t st);  // *out = arr[*idx]
int launchRleAndDecide(const BackendBuffers& b, const StreamOffsets& so, uint32_t N, cudaStream_t st);
int launchDecide(const BackendBuffers& b, const StreamOffsets& so, SeqTable* hostMirror, cudaStream_t st);
int launchHistogram(const BackendBuffers& b, uint32_t upperBoundM, int nSeq, cudaStream_t st);
int launchPack(const BackendBuffers& b, uint32_t M, int nSeq, cudaStream_t st);
int launchResetHistogram(const BackendBuffers& b, uint32_t nEntries, cudaStream_t st);
// counts[r * countStride] = valid entries of rank r (the in-library exchange gathers {count, status} pairs: stride 2)
int launchMergeEntries(const BackendBuffers& b, const HistEntry* gathered, const uint32_t* counts, int countStride, int nranks, uint32_t nMax, cudaStream_t st);

}  // namespace mnenc
