// Host logic of the multi-GPU stripe path (one image sharded by block columns over N ranks).
// Streams are column-major, so rank r holds the r-th contiguous slice of every symbol stream. What has to be
// stitched across ranks before the per-rank kernels can produce the reference's bytes:
//   * the DC delta chain (CompressedImage.cpp:487-505): the value preceding a rank's first DC entry,
//   * run-length coding (huffman::runLengthEncode, Huffman.cpp:246-279): a run may start on an earlier rank and end
//     on a later one; its output symbols [v, v, len-2] are emitted in chunks of 32 769 counted from the run's GLOBAL
//     start, each chunk by the rank that holds the chunk's first element,
//   * positions: first-occurrence indices and bit offsets are global (exclusive sums over ranks).
// Everything here is computed identically on every rank from all-gathered boundary records; no symbol data moves.
#include <algorithm>

#include "mnenc_internal.h"

namespace mnenc {

void stitchRuns(const StripeBoundary* rec, int nranks, int nStreams, int rank, uint32_t* carryIn, uint32_t* extendOut) {
    for (int s = 0; s < nStreams; ++s) {
        const StripeBoundary& me = rec[static_cast<size_t>(rank) * nStreams + s];
        uint32_t carry = 0, ext = 0;
        if (me.n) {
            for (int q = rank - 1; q >= 0; --q) {  // same-valued run ending right before this slice
                const StripeBoundary& o = rec[static_cast<size_t>(q) * nStreams + s];
                if (o.n == 0) continue;
                if (o.last != me.first) break;
                carry += o.lastRun;
                if (o.lastRun < o.n) break;  // the run started inside rank q
            }
            for (int q = rank + 1; q < nranks; ++q) {  // continuation of this slice's last run
                const StripeBoundary& o = rec[static_cast<size_t>(q) * nStreams + s];
                if (o.n == 0) continue;
                if (o.first != me.last) break;
                ext += o.firstRun;
                if (o.firstRun < o.n) break;
            }
        }
        carryIn[s] = carry;
        extendOut[s] = ext;
    }
}

uint32_t rleChunkedSize(uint32_t carry, uint32_t len, uint64_t total) {
    // chunks of the global run start at multiples of 32 769; this slice emits those that start inside [carry, carry+len)
    const uint64_t kChunk = 32769;
    uint32_t out = 0;
    for (uint64_t k = (static_cast<uint64_t>(carry) + kChunk - 1) / kChunk; k * kChunk < static_cast<uint64_t>(carry) + len; ++k) {
        const uint64_t l = std::min<uint64_t>(kChunk, total - k * kChunk);
        out += l == 1 ? 1u : 3u;
    }
    return out;
}

}  // namespace mnenc
