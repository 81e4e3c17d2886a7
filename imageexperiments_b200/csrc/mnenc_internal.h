// Internal declarations shared by the CUDA kernels, the host entropy-table builder and the C-ABI layer.
// Not installed; the public boundary is include/mnenc.h.
#pragma once

#include <cstddef>
#include <cstdint>
#include <functional>
#include <string>
#include <vector>

#ifdef __CUDACC__
#include <cuda_runtime.h>
#endif

namespace mnenc {

constexpr int kN = 64;        // pixels per 8x8 block (only block_size 8 is supported)
constexpr int kMaxK = 32;     // greedy iterations per block-channel (Compression.cpp:28)
constexpr int kRowsPad = 512; // base dictionary rows padded (510 real atoms)
constexpr int kMaxSeq = 6 * kMaxK + 1;  // coded sequences: `lengths` + 6K symbol streams

// host: fp16 operand images + exact row table of the fp16 kernel; returns max ||d - fp16(d)||_2 and max ||d||_2 over all atoms
void buildF16Images(const float* baseRowMajor, int nBase, const int32_t* detailRows, const float* const detail[3], uint16_t* baseUm16,
                    uint16_t* detUm16, float* rows32, double* maxRoundingNorm, double* maxNorm);
size_t f16Rows32Count(int nBase);

// ------------------------------------------------------------------------------------------------
// Level-synchronous matching pursuit (mp_wave.cu): every greedy iteration of ALL active block-channels is one
// round of kernels, so that dictionary tiles are shared by whole tiles of block-channels instead of being streamed
// once per block-channel.
// ------------------------------------------------------------------------------------------------
constexpr int kCandSlots = 8;            // base-atom candidate slots per block-channel (more candidates: overflow queue)
constexpr int kCandSlotsDetail = 4;      // candidate slots for rows of the appended KLT groups
constexpr uint32_t kWaveNone = 0xFFFFFFFFu;
constexpr int kWaveTile = 128;           // block-channels per MMA tile (M of tcgen05.mma)
#ifdef __CUDACC__
struct WaveParams {
    // image
    const uint8_t* rgb;
    size_t stride;
    int width, height, nbx, nby;
    int K, nBase;
    // several frames of one geometry in one pursuit (batch mode): frame f owns blocks [f * frameBlocks, f * frameBlocks + nbx * nby)
    // of the block numbering; frameBlocks >= nbx * nby, the blocks in between do not exist. frames == nullptr: one image (`rgb`).
    const uint8_t* const* frames;
    uint32_t frameBlocks;
    // dictionary (device)
    const uint16_t* baseUm16;  // [512][64] fp16, canonical K-major UMMA image
    const uint16_t* detUm16;   // [3][nBase] fp16 images of the KLT groups, 8 KB each
    const float* rows32;       // exact FP32 rows: base atoms, then 64 per group
    const uint8_t* detRows;    // [nBase] rows per group (62 / 63)
    const float* quant;        // [3][32]
    float epsRel16, epsAbs16;
    // outputs
    uint16_t* counts;          // [nbc]
    uint32_t* pairs;           // [nbc][K]; entries >= counts[bc] are not written
    // per block-channel state
    float* res32;              // [nbc][64] exact residual
    uint16_t* res16;           // [nbc][64] residual rounded to fp16 (MMA operand rows, 128 B each)
    uint4* state;              // {prev | height << 16, reps | ngroups << 8, twoEps bits, 0}
    uint16_t* groups;          // [nbc][32] base atoms picked so far, pick order (dynamicBasisFast, CompressedImage.cpp:252-289)
    unsigned long long* best;  // [nbc] exact arg-max key of candidates that found no free slot
    uint32_t* amax;            // [nbc] max |approximate base score| of the current iteration (float bits; reset by the decide pass)
    uint32_t* ncand;           // [nbc] candidates among the base atoms (rewritten by every base pass)
    uint32_t* cand;            // [nbc][kCandSlots] dictionary index | row id in rows32 << 12
    uint32_t* ncandDetail;     // [nbc] candidates among the rows of the appended groups (reset by the decide pass)
    uint32_t* candDetail;      // [nbc][kCandSlotsDetail]
    uint32_t* ovfCount;        // candidates that found no slot: queue re-scored by waveOverflowKernel (reset by the decide pass)
    uint2* ovf;                // [ovfCap] {block-channel, dictionary index | row id << 12}
    uint32_t ovfCap;
    // work lists
    uint32_t* act[2];          // active block-channels of the current / next iteration
    uint32_t* nAct;            // [K + 1] active block-channels per iteration (zeroed before the launch)
    uint32_t* binCount;        // [3 * nBase] (channel, atom) bins: groups in use by active block-channels
    uint32_t* binCursor;       // [3 * nBase]
    uint32_t* binOffset;       // [3 * nBase + 1] exclusive scan of binCount
    uint32_t* binTileStart;    // [3 * nBase + 1] exclusive scan of ceil(binCount / 128)
    uint32_t* members;         // bin member lists: block-channel | group position << 27
    uint2* memberInfo;         // per member: {dictionary index of the group's row 0 | fp16(2 eps, rounded up) << 16,
                               //              float bits of (max |base score| - 2 eps)}
    unsigned long long* stats; // [2]: sum of dictionary heights over all iterations, sum of counts
    uint32_t nbc;
};
size_t waveMembersCapacity(uint32_t nbc, int K);
cudaError_t launchWaveInit(const WaveParams& p, cudaStream_t st);
// one greedy iteration of all active block-channels; nActHint >= the number of active block-channels (sizes the grids).
// Returns the number of kernels launched through *launches.
cudaError_t launchWaveIteration(const WaveParams& p, int iter, uint32_t nActHint, int numSMs, cudaStream_t st, int* launches);
#endif

// ------------------------------------------------------------------------------------------------
// Host side of the entropy back end (entropy_host.cpp). Everything here is computed from per-sequence
// histograms + first-occurrence positions only.
// ------------------------------------------------------------------------------------------------
struct SymbolStat {
    uint32_t symbol;
    uint32_t freq;
    uint32_t firstPos;
};

// MSB-first bit string builder for the (small) header parts written on the host.
struct BitString {
    std::vector<uint8_t> bytes;
    uint64_t nbits = 0;
    void put(uint64_t value, int width);   // width <= 64
    void append(const BitString& other);
};

struct SequencePlan {
    // decision
    bool golomb = false;
    uint32_t golombM = 0;
    // Huffman table in the order of the `stats` vector passed in: code/len per entry
    std::vector<uint32_t> code;
    std::vector<uint8_t> len;
    uint32_t eofCode = 0;
    uint8_t eofLen = 0;
    uint32_t maxCodeLen = 0; // longest Huffman code (the device packer handles <= 32 bits)
    BitString header;        // everything before the first payload code: mode flag, table / m
    uint64_t payloadBits = 0;  // sum of code lengths of the data symbols (without EOF)
    uint64_t totalBits() const { return header.nbits + payloadBits + (golomb ? 0 : eofLen); }
};

// Builds the plan of one coded sequence (compressed::writeHuffmanOrGolomb, CompressedImage.cpp:359-390).
// `stats` may be in any order; it is sorted by firstPos internally and code/len are returned in the
// caller's original order.
void planSequence(const std::vector<SymbolStat>& stats, SequencePlan& plan);

// ------------------------------------------------------------------------------------------------
// Multi-GPU stripe path, host logic (stripe_host.cpp)
// ------------------------------------------------------------------------------------------------
struct StripeBoundary {   // one per (rank, symbol stream), all-gathered
    uint32_t n;           // entries of the stream held by the rank
    uint32_t first, firstRun;  // first value and length of the run it starts (n if the slice is one run)
    uint32_t last, lastRun;    // last value and length of the run it ends
    uint32_t interior;         // RLE symbols of the runs strictly between the first and the last run of the slice
};
// For `rank`: carryIn[s] = length of the equal-valued run that ends right before its slice of stream s,
// extendOut[s] = length by which its last run continues on later ranks.
void stitchRuns(const StripeBoundary* rec, int nranks, int nStreams, int rank, uint32_t* carryIn, uint32_t* extendOut);
// RLE symbols a slice emits for `len` elements of a run that has `carry` elements before it and `total` overall.
uint32_t rleChunkedSize(uint32_t carry, uint32_t len, uint64_t total);

struct StripePlan {
    int nranks = 0, nStreams = 0;
    std::vector<uint32_t> emitted, carry, ext;  // [nranks][nStreams]: RLE symbols emitted by a rank; its run stitching
    std::vector<uint64_t> rawLen, rleLen;       // [nStreams] totals over the ranks
    std::vector<uint8_t> useRle;                // [nStreams] decision on the WHOLE stream
};
void planStripes(const StripeBoundary* rec, int nranks, int nStreams, StripePlan& plan);

// compressed::createQuantizationTables, float overload (CompressedImage.cpp:168-210).
int quantTables(int K, int blockSize, double bpp, float* out3K);

// ------------------------------------------------------------------------------------------------
// Decoder, host side (decode_host.cpp): readCompressedFast (CompressedImage.cpp:709-782) in three steps, so that the
// caller can size its (pinned) buffers between them and start uploading `lengths` while the streams are still being read.
// Each step returns 0, or -1 with a message in `err`.
// ------------------------------------------------------------------------------------------------
class HostPool;
class MnParser {
public:
    // pool: large Huffman payloads are decoded by its threads in parallel (speculative chunk starts, decode_host.cpp);
    // nullptr or a pool of one thread = the serial reader. Same result either way.
    MnParser(const uint8_t* bytes, size_t n, HostPool* pool = nullptr);
    ~MnParser();
    MnParser(const MnParser&) = delete;
    MnParser& operator=(const MnParser&) = delete;
    int readHeader(std::string& err, bool headerOnly = false);  // magic, dimensions, K, block size, 3K quantisers -> fields below
    int readLengths(uint16_t* counts, std::string& err);  // counts[3 * nblocks]; fills streamOffset
    int readStreams(uint16_t* arena, std::string& err);   // arena[streamOffset[6K]]: run lengths expanded, DC chains undone
    // Runs body() (calls of readLengths / readStreams) with the pool's threads standing by as spinning helpers for the
    // parallel reader; without a pool it just calls body().
    int withTeam(const std::function<int()>& body);

    uint32_t width = 0, height = 0;
    int K = 0, blockSize = 0;
    float quant[3 * kMaxK] = {};
    size_t nblocks = 0;
    size_t streamOffset[6 * kMaxK + 1] = {};  // stream s = arena[streamOffset[s], streamOffset[s + 1])
    uint64_t bitsRead = 0;

private:
    struct Impl;
    Impl* impl_;
    size_t nbytes_;
};

}  // namespace mnenc
