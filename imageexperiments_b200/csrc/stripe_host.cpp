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

namespace mnenc {

// Everything every rank derives — identically — from the all-gathered boundary records of one image: how many RLE
// symbols each rank emits for each stream, the RLE-or-raw decision of compressed::writeCompressed
// (CompressedImage.cpp:508-516: RLE iff rle.size() + 4 < stream.size(), on the WHOLE stream), and the lengths.
void planStripes(const StripeBoundary* rec, int nranks, int nStreams, StripePlan& plan) {
    plan.nranks = nranks;
    plan.nStreams = nStreams;
    plan.emitted.assign(static_cast<size_t>(nranks) * nStreams, 0);
    plan.carry.assign(static_cast<size_t>(nranks) * nStreams, 0);
    plan.ext.assign(static_cast<size_t>(nranks) * nStreams, 0);
    plan.rawLen.assign(nStreams, 0);
    plan.rleLen.assign(nStreams, 0);
    plan.useRle.assign(nStreams, 0);
    std::vector<uint32_t> carry(nStreams), ext(nStreams);
    for (int q = 0; q < nranks; ++q) {
        stitchRuns(rec, nranks, nStreams, q, carry.data(), ext.data());
        for (int s = 0; s < nStreams; ++s) {
            const StripeBoundary& r = rec[static_cast<size_t>(q) * nStreams + s];
            uint32_t e = 0;
            if (r.n) {
                if (r.firstRun >= r.n) {  // the slice is one run
                    e = rleChunkedSize(carry[s], r.n, static_cast<uint64_t>(carry[s]) + r.n + ext[s]);
                } else {
                    e = rleChunkedSize(carry[s], r.firstRun, static_cast<uint64_t>(carry[s]) + r.firstRun) + r.interior +
                        rleChunkedSize(0, r.lastRun, static_cast<uint64_t>(r.lastRun) + ext[s]);
                }
            }
            const size_t i = static_cast<size_t>(q) * nStreams + s;
            plan.emitted[i] = e;
            plan.carry[i] = carry[s];
            plan.ext[i] = ext[s];
            plan.rawLen[s] += r.n;
            plan.rleLen[s] += e;
        }
    }
    for (int s = 0; s < nStreams; ++s) plan.useRle[s] = plan.rleLen[s] + 4 < plan.rawLen[s] ? 1 : 0;
}

}  // namespace mnenc
