// Device side of the entropy back end. See backend_kernels.cuh for the reference functions replaced.
// All kernels are HBM/L2-bound integer work: coalesced accesses, warp-shuffle prefix sums, grids sized by
// the data (a few MB per image), no tensor-core shaping.
#include "backend_kernels.cuh"
#include "word_acc.h"

#include <algorithm>
#include <cstdlib>

namespace mnenc {

namespace {

constexpr unsigned kFull = 0xFFFFFFFFu;
constexpr int kScanThreads = 256;
constexpr int kScanItems = 8;
constexpr int kScanTile = kScanThreads * kScanItems;

// ------------------------------------------------------------------------------------------------
// Generic three-phase device-wide exclusive scan with fused input / output functors.
//   phase 1: blockSums[b] = sum of in(i) over tile b
//   phase 2: one CTA turns blockSums into exclusive prefixes, total at blockSums[nTiles]
//   phase 3: out(i, exclusive prefix of i, in(i)) for every i, and out.total(sum) once
// `nPtr` (optional) is a device-resident element count that overrides `n` (tiles past it exit at once).
// ------------------------------------------------------------------------------------------------
template <class T>
__device__ __forceinline__ T warpInclusive(T v, int lane) {
#pragma unroll
    for (int off = 1; off < 32; off <<= 1) {
        T t = __shfl_up_sync(kFull, v, off);
        if (lane >= off) v += t;
    }
    return v;
}

// Exclusive prefix of `v` across the CTA (kScanThreads threads); returns the prefix, sets total.
template <class T>
__device__ __forceinline__ T blockExclusive(T v, T* total) {
    __shared__ T warpSums[kScanThreads / 32];
    __shared__ T blockTotal;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    T incl = warpInclusive(v, lane);
    if (lane == 31) warpSums[warp] = incl;
    __syncthreads();
    if (warp == 0) {
        T w = lane < kScanThreads / 32 ? warpSums[lane] : T(0);
        T wi = warpInclusive(w, lane);
        if (lane < kScanThreads / 32) warpSums[lane] = wi - w;
        if (lane == kScanThreads / 32 - 1) blockTotal = wi;
    }
    __syncthreads();
    T res = warpSums[warp] + incl - v;
    *total = blockTotal;
    __syncthreads();
    return res;
}

template <class T, class In>
__global__ void __launch_bounds__(kScanThreads) scanReduceKernel(In in, uint32_t n, const uint32_t* nPtr, T* blockSums) {
    if (nPtr) n = *nPtr;
    const uint32_t base = blockIdx.x * kScanTile;
    if (base >= n) {
        if (threadIdx.x == 0) blockSums[blockIdx.x] = T(0);
        return;
    }
    T sum = T(0);
#pragma unroll
    for (int k = 0; k < kScanItems; ++k) {
        const uint32_t i = base + k * kScanThreads + threadIdx.x;  // strided: coalesced, order irrelevant for a sum
        if (i < n) sum += static_cast<T>(in(i));
    }
    T total;
    blockExclusive(sum, &total);
    if (threadIdx.x == 0) blockSums[blockIdx.x] = total;
}

template <class T>
__global__ void __launch_bounds__(1024) scanBlockSumsKernel(T* blockSums, uint32_t nTiles) {
    __shared__ T warpSums[32];
    __shared__ T carryS, chunkTotal;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) carryS = T(0);
    __syncthreads();
    for (uint32_t base = 0; base < nTiles; base += 1024) {
        const uint32_t i = base + threadIdx.x;
        const T v = i < nTiles ? blockSums[i] : T(0);
        const T incl = warpInclusive(v, lane);
        if (lane == 31) warpSums[warp] = incl;
        __syncthreads();
        if (warp == 0) {
            const T w = warpSums[lane];
            const T wi = warpInclusive(w, lane);
            warpSums[lane] = wi - w;
            if (lane == 31) chunkTotal = wi;
        }
        __syncthreads();
        if (i < nTiles) blockSums[i] = carryS + warpSums[warp] + incl - v;
        __syncthreads();
        if (threadIdx.x == 0) carryS += chunkTotal;
        __syncthreads();
    }
    if (threadIdx.x == 0) blockSums[nTiles] = carryS;
}

template <class T, class In, class Out>
__global__ void __launch_bounds__(kScanThreads) scanApplyKernel(In in, Out out, uint32_t n, const uint32_t* nPtr, const T* blockSums,
                                                                uint32_t nTiles) {
    if (nPtr) n = *nPtr;
    if (blockIdx.x == 0 && threadIdx.x == 0) out.total(blockSums[nTiles], n);
    const uint32_t base = blockIdx.x * kScanTile;
    if (base >= n) return;
    // blocked arrangement: thread t owns items [base + t*8, base + t*8 + 8) so prefixes stay in registers
    T vals[kScanItems];
    T sum = T(0);
    const uint32_t first = base + threadIdx.x * kScanItems;
#pragma unroll
    for (int k = 0; k < kScanItems; ++k) {
        vals[k] = (first + k < n) ? static_cast<T>(in(first + k)) : T(0);
        sum += vals[k];
    }
    T total;
    T prefix = blockExclusive(sum, &total) + blockSums[blockIdx.x];
#pragma unroll
    for (int k = 0; k < kScanItems; ++k) {
        if (first + k < n) out(first + k, prefix, vals[k]);
        prefix += vals[k];
    }
}

template <class T, class In, class Out>
int deviceScan(In in, Out out, uint32_t nUpper, const uint32_t* nPtr, T* blockSums, cudaStream_t st) {
    const uint32_t nTiles = (nUpper + kScanTile - 1) / kScanTile;
    if (nTiles == 0) {
        // still report the total (0) so downstream sentinels are written
        scanBlockSumsKernel<T><<<1, 1024, 0, st>>>(blockSums, 0);
        scanApplyKernel<T, In, Out><<<1, kScanThreads, 0, st>>>(in, out, 0, nPtr, blockSums, 0);
        return 2;
    }
    scanReduceKernel<T, In><<<nTiles, kScanThreads, 0, st>>>(in, nUpper, nPtr, blockSums);
    scanBlockSumsKernel<T><<<1, 1024, 0, st>>>(blockSums, nTiles);
    scanApplyKernel<T, In, Out><<<nTiles, kScanThreads, 0, st>>>(in, out, nUpper, nPtr, blockSums, nTiles);
    return 3;
}

// ------------------------------------------------------------------------------------------------
// Stream scatter (CompressedImage.cpp:612-629): stream (layer, depth) receives one entry from every block
// whose count in that layer exceeds depth, in column-major block order == ascending block index.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kTileBlocks) tileCountsKernel(const uint16_t* counts, uint32_t nblocks, int K, uint32_t* tileCounts) {
    __shared__ uint32_t hist[3][kMaxK + 1];
    for (int i = threadIdx.x; i < 3 * (kMaxK + 1); i += blockDim.x) (&hist[0][0])[i] = 0;
    __syncthreads();
    const uint32_t blk = blockIdx.x * kTileBlocks + threadIdx.x;
    if (blk < nblocks) {
#pragma unroll
        for (int ch = 0; ch < 3; ++ch) atomicAdd(&hist[ch][counts[3 * blk + ch]], 1u);
    }
    __syncthreads();
    // #blocks with count > d  ==  sum_{c > d} hist[c]
    for (int i = threadIdx.x; i < 3 * K; i += blockDim.x) {
        const int ch = i / K, d = i - ch * K;
        uint32_t s = 0;
        for (int c = d + 1; c <= K; ++c) s += hist[ch][c];
        tileCounts[static_cast<size_t>(blockIdx.x) * 3 * K + i] = s;
    }
}

// One warp per (layer, depth) column: exclusive scan over tiles in place, column total -> totals[col].
__global__ void tileScanKernel(uint32_t* tileCounts, uint32_t ntiles, int cols, uint32_t* totals) {
    const int lane = threadIdx.x & 31;
    const int col = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (col >= cols) return;
    uint32_t carry = 0;
    for (uint32_t base = 0; base < ntiles; base += 32) {
        const uint32_t t = base + lane;
        const uint32_t v = t < ntiles ? tileCounts[static_cast<size_t>(t) * cols + col] : 0;
        const uint32_t incl = warpInclusive(v, lane);
        if (t < ntiles) tileCounts[static_cast<size_t>(t) * cols + col] = carry + incl - v;
        carry += __shfl_sync(kFull, incl, 31);
    }
    if (lane == 0) totals[col] = carry;
}

__global__ void __launch_bounds__(kTileBlocks) scatterKernel(const uint16_t* counts, const uint32_t* pairs, uint32_t nblocks, int K,
                                                             const uint32_t* tileOffsets, const __grid_constant__ StreamOffsets so,
                                                             uint16_t* arena) {
    __shared__ uint32_t warpCnt[kTileBlocks / 32][kMaxK];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t blk = blockIdx.x * kTileBlocks + threadIdx.x;
    for (int ch = 0; ch < 3; ++ch) {
        const int count = blk < nblocks ? counts[3 * blk + ch] : 0;
        // rank of this block among the tile's blocks with count > d, for every d < count
        int maxCount = count;
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) maxCount = max(maxCount, __shfl_xor_sync(kFull, maxCount, off));
        for (int d = lane; d < K; d += 32) warpCnt[warp][d] = 0;
        __syncwarp();
        uint32_t myRank[kMaxK];  // only entries < count are meaningful; kept in local memory only when K large
        for (int d = 0; d < maxCount; ++d) {
            const unsigned m = __ballot_sync(kFull, count > d);
            if (count > d) myRank[d] = __popc(m & ((1u << lane) - 1u));
            if (lane == 0) warpCnt[warp][d] = __popc(m);
        }
        __syncthreads();
        if (count > 0) {
            const uint32_t* tileOff = tileOffsets + static_cast<size_t>(blockIdx.x) * 3 * K + ch * K;
            const uint32_t* pp = pairs + (static_cast<size_t>(3) * blk + ch) * K;
            for (int d = 0; d < count; ++d) {
                uint32_t before = 0;
                for (int w = 0; w < warp; ++w) before += warpCnt[w][d];
                const uint32_t pos = tileOff[d] + before + myRank[d];
                const uint32_t pr = pp[d];
                const int sid = 2 * (ch * K + d);
                arena[so.off[sid] + pos] = static_cast<uint16_t>(pr & 0xFFFFu);
                arena[so.off[sid + 1] + pos] = static_cast<uint16_t>(pr >> 16);
            }
        }
        __syncthreads();
    }
}

// DC delta + zig-zag of the three depth-0 coefficient streams (CompressedImage.cpp:487-505), two steps so the
// update can be written back in place.
__global__ void dcDeltaKernel(const uint16_t* arena, const __grid_constant__ StreamOffsets so, int K, uint16_t* tmp, uint32_t tmpStride,
                              DcPrev dcPrev) {
    const int ch = blockIdx.y;
    const uint32_t off = so.off[2 * ch * K + 1], n = so.off[2 * ch * K + 2] - off;
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int32_t cur = arena[off + i];
    const int32_t prev = i ? arena[off + i - 1] : static_cast<int32_t>(dcPrev.v[ch]);
    const int32_t d = cur - prev;
    tmp[static_cast<size_t>(ch) * tmpStride + i] = static_cast<uint16_t>((static_cast<uint32_t>(d) << 1) ^ static_cast<uint32_t>(d >> 31));
}
__global__ void dcWriteBackKernel(uint16_t* arena, const __grid_constant__ StreamOffsets so, int K, const uint16_t* tmp, uint32_t tmpStride) {
    const int ch = blockIdx.y;
    const uint32_t off = so.off[2 * ch * K + 1], n = so.off[2 * ch * K + 2] - off;
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) arena[off + i] = tmp[static_cast<size_t>(ch) * tmpStride + i];
}

}  // namespace

namespace {

// ------------------------------------------------------------------------------------------------
// Run-length coding of all 6K streams at once (Huffman.cpp:246-279). Runs never cross a stream boundary.
//   run of v, length L  ->  floor(L / 32769) x [v, v, 32767]  then  L % 32769 == 1 ? [v] : >= 2 ? [v, v, rem-2]
// ------------------------------------------------------------------------------------------------
__global__ void rleFlagsKernel(const uint16_t* arena, uint32_t n, uint8_t* flags) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) flags[i] = (i == 0 || arena[i] != arena[i - 1]) ? 1 : 0;
}
__global__ void rleMarkStreamStartsKernel(const __grid_constant__ StreamOffsets so, int nStreams, uint32_t n, uint8_t* flags) {
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s < nStreams && so.off[s] < n) flags[so.off[s]] = 1;
}

struct FlagIn {
    const uint8_t* flags;
    __device__ uint32_t operator()(uint32_t i) const { return flags[i]; }
};
struct RunStartOut {
    uint32_t* runIdx;    // [n+1]
    uint32_t* startPos;  // [n+1]
    __device__ void operator()(uint32_t i, uint32_t prefix, uint32_t v) const {
        runIdx[i] = prefix;
        if (v) startPos[prefix] = i;
    }
    __device__ void total(uint32_t nRuns, uint32_t n) const {
        runIdx[n] = nRuns;
        startPos[nRuns] = n;
    }
};

// A maximal run of v is coded in chunks of 32 769 elements counted from the run's start: [v, v, len-2], or [v] for a
// chunk of one element. When the image is stripe-sharded a run may begin on an earlier rank (`carry` elements there)
// and continue on later ranks (`ext` elements); a slice emits the chunks whose first element it holds.
constexpr uint32_t kRleChunk = 32769u;
__device__ __forceinline__ uint32_t rleOutSize(uint32_t L) {
    const uint32_t full = L / kRleChunk, rem = L - full * kRleChunk;
    return 3u * full + (rem == 0 ? 0u : (rem == 1 ? 1u : 3u));
}
__device__ __forceinline__ uint32_t rleOutSizeStitched(uint32_t carry, uint32_t L, uint32_t ext) {
    const unsigned long long total = static_cast<unsigned long long>(carry) + L + ext;
    uint32_t out = 0;
    for (unsigned long long k = (carry + static_cast<unsigned long long>(kRleChunk) - 1) / kRleChunk; k * kRleChunk < static_cast<unsigned long long>(carry) + L; ++k)
        out += (total - k * kRleChunk == 1) ? 1u : 3u;
    return out;
}
// carry / ext of run [s0, s1): non-zero only for the first / last run of a stream
__device__ __forceinline__ void stitchOf(const StripeStitch* st, int nStreams, uint32_t s0, uint32_t s1, uint32_t* carry, uint32_t* ext) {
    int lo = 0, hi = nStreams;  // largest s with off[s] <= s0 (the non-empty stream that contains s0)
    while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (st->off[mid] <= s0) lo = mid;
        else hi = mid;
    }
    *carry = st->off[lo] == s0 ? st->carry[lo] : 0u;
    *ext = st->off[lo + 1] == s1 ? st->ext[lo] : 0u;
}
struct RunSizeIn {
    const uint32_t* startPos;
    const StripeStitch* stitch;  // nullptr: the slice is the whole stream
    int nStreams;
    __device__ uint32_t operator()(uint32_t r) const {
        const uint32_t s0 = startPos[r], s1 = startPos[r + 1];
        if (!stitch) return rleOutSize(s1 - s0);
        uint32_t carry, ext;
        stitchOf(stitch, nStreams, s0, s1, &carry, &ext);
        return (carry | ext) ? rleOutSizeStitched(carry, s1 - s0, ext) : rleOutSize(s1 - s0);
    }
};
struct RunOffsetOut {  // offsets only (stripe path, first pass: sizes of the interior runs for the boundary records)
    uint32_t* outOff;
    __device__ void operator()(uint32_t r, uint32_t prefix, uint32_t /*size*/) const { outOff[r] = prefix; }
    __device__ void total(uint32_t sum, uint32_t nRuns) const { outOff[nRuns] = sum; }
};
struct RleWriteOut {
    const uint32_t* startPos;
    const uint16_t* arena;
    uint16_t* rle;
    uint32_t* outOff;  // [nRuns+1]
    const StripeStitch* stitch;
    int nStreams;
    __device__ void operator()(uint32_t r, uint32_t prefix, uint32_t /*size*/) const {
        outOff[r] = prefix;
        const uint32_t s0 = startPos[r], s1 = startPos[r + 1];
        uint32_t L = s1 - s0;
        const uint16_t v = arena[s0];
        uint32_t o = prefix;
        uint32_t carry = 0, ext = 0;
        if (stitch) stitchOf(stitch, nStreams, s0, s1, &carry, &ext);
        if (carry | ext) {
            const unsigned long long total = static_cast<unsigned long long>(carry) + L + ext;
            for (unsigned long long k = (carry + static_cast<unsigned long long>(kRleChunk) - 1) / kRleChunk; k * kRleChunk < static_cast<unsigned long long>(carry) + L; ++k) {
                const unsigned long long len = min(static_cast<unsigned long long>(kRleChunk), total - k * kRleChunk);
                if (len == 1) rle[o++] = v;
                else rle[o] = v, rle[o + 1] = v, rle[o + 2] = static_cast<uint16_t>(len - 2), o += 3;
            }
            return;
        }
        while (L >= kRleChunk) {
            rle[o] = v, rle[o + 1] = v, rle[o + 2] = 32767;
            o += 3;
            L -= kRleChunk;
        }
        if (L == 1) rle[o] = v;
        else if (L >= 2) rle[o] = v, rle[o + 1] = v, rle[o + 2] = static_cast<uint16_t>(L - 2);
    }
    __device__ void total(uint32_t sum, uint32_t nRuns) const { outOff[nRuns] = sum; }
};

// Stripe path: boundary record of every stream slice (what the other ranks need to stitch runs).
__global__ void stripeRecordsKernel(const __grid_constant__ StreamOffsets so, int nStreams, const uint16_t* arena, const uint32_t* runIdx,
                                    const uint32_t* startPos, const uint32_t* outOff, StripeBoundary* rec) {
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= nStreams) return;
    const uint32_t a = so.off[s], b = so.off[s + 1];
    StripeBoundary r{};
    if (b > a) {
        const uint32_t r0 = runIdx[a], r1 = runIdx[b];
        r.n = b - a;
        r.first = arena[a];
        r.firstRun = startPos[r0 + 1] - a;
        r.last = arena[b - 1];
        r.lastRun = b - startPos[r1 - 1];
        r.interior = (r1 - r0 >= 3) ? outOff[r1 - 1] - outOff[r0 + 1] : 0u;
    }
    rec[s] = r;
}

// One CTA: RLE-or-raw decision per stream (CompressedImage.cpp:508-516) and the table of coded sequences.
__global__ void decideKernel(const __grid_constant__ StreamOffsets so, int nStreams, uint32_t nCounts, const uint32_t* runIdx,
                             const uint32_t* outOff, SeqTable* table, SeqTable* hostMirror) {
    __shared__ SeqTable t;
    const int s = threadIdx.x;
    if (s == 0) {
        t.desc[0] = SeqDesc{nCounts, nCounts, 0u, 0u};
        t.nRuns = runIdx[so.off[nStreams]];
    }
    if (s < nStreams) {
        const uint32_t a = so.off[s], b = so.off[s + 1];
        const uint32_t raw = b - a;
        const uint32_t rleOff = outOff[runIdx[a]];
        const uint32_t rleLen = outOff[runIdx[b]] - rleOff;
        SeqDesc d;
        d.rawLen = raw;
        if (rleLen + 4 < raw) d.len = rleLen, d.off = rleOff, d.src = 2;
        else d.len = raw, d.off = a, d.src = 1;
        t.desc[1 + s] = d;
    }
    for (int q = threadIdx.x; q < kMaxSeq; q += blockDim.x) t.posBase[q] = 0;
    __syncthreads();
    if (s == 0) {
        uint32_t acc = 0;
        for (int q = 0; q <= nStreams; ++q) {
            t.start[q] = acc;
            acc += t.desc[q].len;
        }
        t.start[nStreams + 1] = acc;
    }
    __syncthreads();
    for (int i = threadIdx.x; i < static_cast<int>(sizeof(SeqTable) / 4); i += blockDim.x) {
        reinterpret_cast<uint32_t*>(table)[i] = reinterpret_cast<const uint32_t*>(&t)[i];
        if (hostMirror) reinterpret_cast<uint32_t*>(hostMirror)[i] = reinterpret_cast<const uint32_t*>(&t)[i];
    }
}

// ------------------------------------------------------------------------------------------------
// Access to the concatenation of the coded sequences: element g in [0, M).
// ------------------------------------------------------------------------------------------------
struct SeqView {
    const SeqTable* table;
    const uint16_t* counts;
    const uint16_t* arena;
    const uint16_t* rle;
};
__device__ __forceinline__ int findSeq(const uint32_t* start, int nSeq, uint32_t g) {
    int lo = 0, hi = nSeq;  // largest q with start[q] <= g, skipping empty sequences automatically
    while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (start[mid] <= g) lo = mid;
        else hi = mid;
    }
    return lo;
}
__device__ __forceinline__ uint16_t loadSym(const SeqView& v, const SeqDesc& d, uint32_t k) {
    const uint16_t* base = d.src == 0 ? v.counts : (d.src == 1 ? v.arena : v.rle);
    return base[d.off + k];
}

// ------------------------------------------------------------------------------------------------
// Histogram + first-occurrence position per sequence (Huffman.cpp:47-64).
// Each CTA owns 4096 consecutive elements of the concatenation; symbols < 4096 of the CTA's first
// sequence are privatised in shared memory, everything else goes straight to L2 atomics.
// ------------------------------------------------------------------------------------------------
constexpr int kHistChunk = 16384;  // elements per CTA: amortises the 32 KB bin initialisation and the flush of the private bins
__global__ void __launch_bounds__(256) histogramKernel(SeqView v, int nSeq, uint32_t* hist, uint32_t* firstPos, uint32_t* maxSym) {
    __shared__ uint32_t sStart[kMaxSeq + 1];
    __shared__ uint32_t sBase[kMaxSeq];
    __shared__ uint32_t sHist[kSmemHistBins];
    __shared__ uint32_t sFirst[kSmemHistBins];
    for (int i = threadIdx.x; i <= nSeq; i += blockDim.x) sStart[i] = v.table->start[i];
    for (int i = threadIdx.x; i < nSeq; i += blockDim.x) sBase[i] = v.table->posBase[i];
    for (int i = threadIdx.x; i < kSmemHistBins; i += blockDim.x) sHist[i] = 0, sFirst[i] = kNoPos;
    __syncthreads();
    const uint32_t M = sStart[nSeq];
    const uint32_t base = blockIdx.x * kHistChunk;
    if (base >= M) return;
    const int q0 = findSeq(sStart, nSeq, base);
    uint32_t localMax = 0;
    int qPrev = -1;
    SeqDesc d;
    const uint32_t end = min(base + kHistChunk, M);
    const int lane = threadIdx.x & 31;
    for (uint32_t g0 = base; g0 < end; g0 += blockDim.x) {  // uniform trip count: __match_any_sync below needs the whole warp
        const uint32_t g = g0 + threadIdx.x;
        const bool ok = g < end;
        uint32_t key = 0xFFFFFFFFu, pos = 0, sym = 0;
        int q = -1;
        if (ok) {
            q = (g < sStart[q0 + 1]) ? q0 : findSeq(sStart, nSeq, g);
            if (q != qPrev) {
                if (qPrev >= 0) atomicMax(&maxSym[qPrev], localMax);
                localMax = 0;
                d = v.table->desc[q];
                qPrev = q;
            }
            const uint32_t k = g - sStart[q];
            sym = loadSym(v, d, k);
            pos = k + sBase[q];  // position in the WHOLE sequence (posBase = 0 unless the image is stripe-sharded)
            localMax = max(localMax, sym);
            key = (static_cast<uint32_t>(q) << 16) | sym;
        }
        // Symbol statistics are very skewed (0, 1, 2 dominate): lanes holding the same (sequence, symbol) elect their lowest
        // lane — it also holds the smallest position — to do ONE add of the group size and ONE min, instead of up to 32
        // serialised same-address atomics each.
        const unsigned peers = __match_any_sync(kFull, key);
        if (ok && lane == __ffs(peers) - 1) {
            const uint32_t cnt = __popc(peers);
            if (q == q0 && sym < kSmemHistBins) {
                atomicAdd(&sHist[sym], cnt);
                atomicMin(&sFirst[sym], pos);
            } else {
                atomicAdd(&hist[static_cast<size_t>(q) * kHistBins + sym], cnt);
                atomicMin(&firstPos[static_cast<size_t>(q) * kHistBins + sym], pos);
            }
        }
    }
    if (qPrev >= 0) atomicMax(&maxSym[qPrev], localMax);
    __syncthreads();
    for (int i = threadIdx.x; i < kSmemHistBins; i += blockDim.x) {
        const uint32_t c = sHist[i];
        if (c) {
            atomicAdd(&hist[static_cast<size_t>(q0) * kHistBins + i], c);
            atomicMin(&firstPos[static_cast<size_t>(q0) * kHistBins + i], sFirst[i]);
        }
    }
}

// Non-empty bins -> compact entry list; the bin's firstPos slot is re-used as symbol -> entry index map,
// the hist bin is cleared for the next encode.
__global__ void __launch_bounds__(256) compactKernel(uint32_t* hist, uint32_t* firstPos, const uint32_t* maxSym, HistEntry* entries,
                                                     uint32_t* nEntries, uint32_t cap) {
    const int q = blockIdx.y;
    const uint32_t sym = blockIdx.x * blockDim.x + threadIdx.x;
    if (blockIdx.x * blockDim.x > maxSym[q]) return;
    const size_t bin = static_cast<size_t>(q) * kHistBins + sym;
    const uint32_t c = hist[bin];
    if (c) {
        const uint32_t e = atomicAdd(nEntries, 1u);
        if (e < cap) {
            entries[e] = HistEntry{static_cast<uint32_t>(q), sym, c, firstPos[bin]};
            firstPos[bin] = e;
        }
        hist[bin] = 0;
    }
}
// Stripe path, sparse histogram exchange: the all-gathered entry lists of every rank ([nranks][nMax], counts[r] valid each) are
// merged into the dense tables: SUM of the counts, MIN of the (global) first-occurrence positions, MAX of the largest symbol.
__global__ void mergeEntriesKernel(const HistEntry* gathered, const uint32_t* counts, int countStride, int nranks, uint32_t nMax, uint32_t* hist,
                                   uint32_t* firstPos, uint32_t* maxSym) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    const int r = blockIdx.y;
    if (r >= nranks || i >= counts[r * countStride]) return;
    const HistEntry e = gathered[static_cast<size_t>(r) * nMax + i];
    if (e.seq >= static_cast<uint32_t>(kMaxSeq) || e.symbol >= static_cast<uint32_t>(kHistBins)) return;
    const size_t bin = static_cast<size_t>(e.seq) * kHistBins + e.symbol;
    atomicAdd(&hist[bin], e.freq);
    atomicMin(&firstPos[bin], e.firstPos);
    atomicMax(&maxSym[e.seq], e.symbol);
}
__global__ void resetHistogramKernel(const HistEntry* entries, uint32_t n, uint32_t* firstPos, uint32_t* maxSym, uint32_t* nEntries) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) firstPos[static_cast<size_t>(entries[i].seq) * kHistBins + entries[i].symbol] = kNoPos;
    if (i < kMaxSeq) maxSym[i] = 0;
    if (i == 0) *nEntries = 0;
}

// ------------------------------------------------------------------------------------------------
// Bit packing (Huffman.cpp:157-161, BitBuffer.cpp:78-110,228-241): code length per symbol -> exclusive
// scan = bit offset -> MSB-first OR into the zeroed output. The output is a byte stream; a logical
// big-endian 32-bit word w covers bytes 4w..4w+3, so OR-ing the byte-swapped value needs no final pass.
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t golombLength(uint32_t v, uint32_t M) {
    const uint32_t q = v / M, rem = v - q * M;
    const uint32_t b = 32 - __clz(M);
    const uint32_t limit = (1u << (b + 1)) - M;
    return rem < limit ? b + q + 1 : b + q + 2;
}
struct CodeLenIn {
    SeqView v;
    int nSeq;
    const CodeTable* codes;
    const uint32_t* symToEntry;  // == firstPos after compaction
    const unsigned long long* entryCodes;
    __device__ uint32_t operator()(uint32_t g) const {
        const int q = findSeq(v.table->start, nSeq, g);
        const SeqDesc d = v.table->desc[q];
        const uint32_t sym = loadSym(v, d, g - v.table->start[q]);
        const uint32_t m = codes->seq[q].golombM;
        if (m) return golombLength(sym, m);
        return static_cast<uint32_t>(entryCodes[symToEntry[static_cast<size_t>(q) * kHistBins + sym]] >> 32);
    }
};
struct PackOut {
    SeqView v;
    int nSeq;
    const CodeTable* codes;
    const uint32_t* symToEntry;
    const unsigned long long* entryCodes;
    uint32_t* out;
};

// Pack scan. A thread owns eight consecutive codes:
//  * they are merged in a register (WordAcc, csrc/word_acc.h — the host build of that header is checked against a bit-by-bit
//    writer by tests/test_native.py) and leave as one atomicOr per output word touched instead of one or two per code
//    (mean code ≈ 4.6 bits: ≈2.2 atomics per thread instead of ≈9);
//  * the sequence of an item is found once per thread (binary search) and then advanced, its descriptor and code plan are
//    reloaded only when it changes, and the table entry fetched for the length is kept for the emission.
// Measured on the 16 MP bench image (round 2, B200): 0.122 ms for the generic per-element scan + per-code atomicOr,
// 0.088 ms with this apply phase alone, 0.072 ms with the reduce phase in the same arrangement — the only variant kept.
struct OrSink {
    uint32_t* out;
    __device__ __forceinline__ void operator()(unsigned long long w, uint32_t v) const { atomicOr(&out[w], v); }
};
// Codes of the eight consecutive elements a thread owns (Huffman: length << 32 | bits, i.e. the table entry; Golomb:
// length << 32 | symbol) and their sequences; returns the sum of the lengths.
__device__ __forceinline__ unsigned long long packThreadCodes(const PackOut& p, uint32_t first, uint32_t n, unsigned long long (&code)[kScanItems],
                                                              int (&seqOf)[kScanItems]) {
    unsigned long long sum = 0;
    SeqWalk walk(p.v.table->start, p.nSeq, first < n ? first : 0u);
    int qLoaded = -1;
    SeqDesc d{};
    uint32_t golombM = 0;
#pragma unroll
    for (int k = 0; k < kScanItems; ++k) {
        const uint32_t g = first + k;
        code[k] = 0;
        seqOf[k] = 0;
        if (g >= n) continue;
        walk.advance(g);
        const int q = walk.q;
        if (q != qLoaded) {
            d = p.v.table->desc[q];
            golombM = p.codes->seq[q].golombM;
            qLoaded = q;
        }
        const uint32_t sym = loadSym(p.v, d, g - walk.seqStart);
        seqOf[k] = q;
        code[k] = golombM ? (static_cast<unsigned long long>(golombLength(sym, golombM)) << 32) | sym
                          : p.entryCodes[p.symToEntry[static_cast<size_t>(q) * kHistBins + sym]];
        sum += code[k] >> 32;
    }
    return sum;
}
// reduce phase in the same blocked arrangement (one sequence search per thread instead of one per element)
__global__ void __launch_bounds__(kScanThreads) packReduceKernel(PackOut p, uint32_t n, unsigned long long* blockSums) {
    const uint32_t base = blockIdx.x * kScanTile;
    if (base >= n) {
        if (threadIdx.x == 0) blockSums[blockIdx.x] = 0ull;
        return;
    }
    unsigned long long code[kScanItems];
    int seqOf[kScanItems];
    const unsigned long long sum = packThreadCodes(p, base + threadIdx.x * kScanItems, n, code, seqOf);
    unsigned long long total;
    blockExclusive(sum, &total);
    if (threadIdx.x == 0) blockSums[blockIdx.x] = total;
}
__global__ void __launch_bounds__(kScanThreads) packApplyKernel(PackOut p, uint32_t n, const unsigned long long* blockSums) {
    const uint32_t base = blockIdx.x * kScanTile;
    if (base >= n) return;
    const uint32_t first = base + threadIdx.x * kScanItems;
    unsigned long long code[kScanItems];
    int seqOf[kScanItems];
    const unsigned long long sum = packThreadCodes(p, first, n, code, seqOf);
    int qLoaded = -1;
    uint32_t golombM = 0;
    unsigned long long total;
    unsigned long long prefix = blockExclusive(sum, &total) + blockSums[blockIdx.x];
    WordAcc<OrSink> acc(OrSink{p.out});
    unsigned long long bitBase = 0;
#pragma unroll
    for (int k = 0; k < kScanItems; ++k) {
        if (first + k < n) {
            if (seqOf[k] != qLoaded) {
                const SeqCode c = p.codes->seq[seqOf[k]];
                bitBase = c.bitBase;
                golombM = c.golombM;
                qLoaded = seqOf[k];
            }
            const unsigned long long pos = bitBase + prefix;
            if (golombM == 0) acc.put(pos, static_cast<uint32_t>(code[k]), static_cast<int>(code[k] >> 32));
            else acc.putGolomb(pos, static_cast<uint32_t>(code[k]), golombM, 32 - __clz(golombM));
        }
        prefix += code[k] >> 32;
    }
    acc.flush();
}


// Stripe path: payload bits this rank contributes to every sequence (sum of the code lengths of its slice).
__global__ void __launch_bounds__(256) seqBitsKernel(CodeLenIn in, uint32_t M, unsigned long long* seqBits) {
    __shared__ unsigned long long sBits[kMaxSeq];
    for (int i = threadIdx.x; i < kMaxSeq; i += blockDim.x) sBits[i] = 0;
    __syncthreads();
    const uint32_t base = blockIdx.x * kHistChunk;
    const uint32_t* start = in.v.table->start;
    const uint32_t end = min(base + kHistChunk, M);
    for (uint32_t g0 = base; g0 < end; g0 += blockDim.x) {  // uniform trip count: the warp reductions below need every lane
        const uint32_t g = g0 + threadIdx.x;
        const bool ok = g < end;
        const int q = ok ? findSeq(start, in.nSeq, g) : -1;
        const uint32_t len = ok ? in(g) : 0u;
        // a warp's 32 consecutive symbols nearly always lie in one sequence: one shared-memory atomic per warp, not per symbol
        const int q0 = __shfl_sync(kFull, q, 0);
        if (__all_sync(kFull, q == q0 || q < 0)) {
            const uint32_t sum = __reduce_add_sync(kFull, len);
            if ((threadIdx.x & 31) == 0 && q0 >= 0) atomicAdd(&sBits[q0], static_cast<unsigned long long>(sum));
        } else if (ok) {
            atomicAdd(&sBits[q], static_cast<unsigned long long>(len));
        }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < in.nSeq; i += blockDim.x)
        if (sBits[i]) atomicAdd(&seqBits[i], sBits[i]);
}

__global__ void readIndirectKernel(const uint32_t* arr, const uint32_t* idx, uint32_t* out) { *out = arr[*idx]; }

__global__ void fillU32Kernel(uint32_t* p, size_t n, uint32_t v) {
    for (size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < n; i += static_cast<size_t>(gridDim.x) * blockDim.x) p[i] = v;
}

}  // namespace

// ------------------------------------------------------------------------------------------------
// Launch wrappers
// ------------------------------------------------------------------------------------------------
int launchTileCounts(const BackendBuffers& b, cudaStream_t st) {
    const int cols = 3 * b.K;
    tileCountsKernel<<<b.ntiles, kTileBlocks, 0, st>>>(b.counts, b.nblocks, b.K, b.tileCounts);
    const int warpsPerCta = 8;
    tileScanKernel<<<(cols + warpsPerCta - 1) / warpsPerCta, warpsPerCta * 32, 0, st>>>(b.tileCounts, b.ntiles, cols, b.totals);
    return 2;
}

int launchScatterRaw(const BackendBuffers& b, const StreamOffsets& so, cudaStream_t st) {
    scatterKernel<<<b.ntiles, kTileBlocks, 0, st>>>(b.counts, b.pairs, b.nblocks, b.K, b.tileCounts, so, b.arena);
    return 1;
}

int launchDcDelta(const BackendBuffers& b, const StreamOffsets& so, const DcPrev& prev, cudaStream_t st) {
    const dim3 grid((b.nblocks + 255) / 256, 3);
    dcDeltaKernel<<<grid, 256, 0, st>>>(b.arena, so, b.K, b.dcTmp, b.nblocks, prev);
    dcWriteBackKernel<<<grid, 256, 0, st>>>(b.arena, so, b.K, b.dcTmp, b.nblocks);
    return 2;
}

int launchScatter(const BackendBuffers& b, const StreamOffsets& so, cudaStream_t st) {
    return launchScatterRaw(b, so, st) + launchDcDelta(b, so, DcPrev{}, st);
}

int launchRleRuns(const BackendBuffers& b, const StreamOffsets& so, uint32_t N, cudaStream_t st) {
    const int nStreams = 6 * b.K;
    int launches = 0;
    if (N) {
        rleFlagsKernel<<<(N + 255) / 256, 256, 0, st>>>(b.arena, N, b.flags);
        rleMarkStreamStartsKernel<<<(nStreams + 255) / 256, 256, 0, st>>>(so, nStreams, N, b.flags);
        launches += 2;
    }
    return launches + deviceScan<uint32_t>(FlagIn{b.flags}, RunStartOut{b.runIdx, b.startPos}, N, nullptr, b.blockSums32, st);
}

int launchRleWrite(const BackendBuffers& b, uint32_t N, const StripeStitch* stitch, cudaStream_t st) {
    // runs: count lives in device memory (runIdx[N]); launch over the upper bound N, extra tiles exit at once
    const int nStreams = 6 * b.K;
    return deviceScan<uint32_t>(RunSizeIn{b.startPos, stitch, nStreams}, RleWriteOut{b.startPos, b.arena, b.rleArena, b.outOff, stitch, nStreams},
                                N, b.runIdx + N, b.blockSums32, st);
}

int launchRleAndDecide(const BackendBuffers& b, const StreamOffsets& so, uint32_t N, cudaStream_t st) {
    return launchRleRuns(b, so, N, st) + launchRleWrite(b, N, nullptr, st);
}

int launchStripeRecords(const BackendBuffers& b, const StreamOffsets& so, uint32_t N, StripeBoundary* rec, cudaStream_t st) {
    const int nStreams = 6 * b.K;
    int launches = deviceScan<uint32_t>(RunSizeIn{b.startPos, nullptr, nStreams}, RunOffsetOut{b.outOff}, N, b.runIdx + N, b.blockSums32, st);
    stripeRecordsKernel<<<(nStreams + 255) / 256, 256, 0, st>>>(so, nStreams, b.arena, b.runIdx, b.startPos, b.outOff, rec);
    return launches + 1;
}

int launchDecide(const BackendBuffers& b, const StreamOffsets& so, SeqTable* hostMirror, cudaStream_t st) {
    decideKernel<<<1, 256, 0, st>>>(so, 6 * b.K, 3 * b.nblocks, b.runIdx, b.outOff, b.seqTable, hostMirror);
    return 1;
}

int launchHistogramOnly(const BackendBuffers& b, uint32_t upperBoundM, int nSeq, cudaStream_t st) {
    SeqView v{b.seqTable, b.counts, b.arena, b.rleArena};
    const uint32_t grid = (upperBoundM + kHistChunk - 1) / kHistChunk;
    if (grid) histogramKernel<<<grid, 256, 0, st>>>(v, nSeq, b.hist, b.firstPos, b.maxSym);
    return grid ? 1 : 0;
}

int launchCompact(const BackendBuffers& b, int nSeq, cudaStream_t st) {
    compactKernel<<<dim3(kHistBins / 256, nSeq), 256, 0, st>>>(b.hist, b.firstPos, b.maxSym, b.entries, b.nEntries, b.entryCap);
    return 1;
}

int launchHistogram(const BackendBuffers& b, uint32_t upperBoundM, int nSeq, cudaStream_t st) {
    return launchHistogramOnly(b, upperBoundM, nSeq, st) + launchCompact(b, nSeq, st);
}

int launchSeqBits(const BackendBuffers& b, uint32_t M, int nSeq, unsigned long long* seqBits, cudaStream_t st) {
    SeqView v{b.seqTable, b.counts, b.arena, b.rleArena};
    cudaMemsetAsync(seqBits, 0, sizeof(unsigned long long) * kMaxSeq, st);
    const uint32_t grid = (M + kHistChunk - 1) / kHistChunk;
    if (grid) seqBitsKernel<<<grid, 256, 0, st>>>(CodeLenIn{v, nSeq, b.codeTable, b.firstPos, b.entryCodes}, M, seqBits);
    return grid ? 1 : 0;
}

int launchReadIndirect(const uint32_t* arr, const uint32_t* idx, uint32_t* out, cudaStream_t st) {
    readIndirectKernel<<<1, 1, 0, st>>>(arr, idx, out);
    return 1;
}

int launchFillU32(uint32_t* p, size_t n, uint32_t value, cudaStream_t st) {
    if (!n) return 0;
    fillU32Kernel<<<static_cast<unsigned>(std::min<size_t>((n + 255) / 256, 4096)), 256, 0, st>>>(p, n, value);
    return 1;
}

int launchPack(const BackendBuffers& b, uint32_t M, int nSeq, cudaStream_t st) {
    if (!M) return 0;
    SeqView v{b.seqTable, b.counts, b.arena, b.rleArena};
    const uint32_t nTiles = (M + kScanTile - 1) / kScanTile;
    const PackOut out{v, nSeq, b.codeTable, b.firstPos, b.entryCodes, b.outWords};
    packReduceKernel<<<nTiles, kScanThreads, 0, st>>>(out, M, b.blockSums64);
    scanBlockSumsKernel<unsigned long long><<<1, 1024, 0, st>>>(b.blockSums64, nTiles);
    packApplyKernel<<<nTiles, kScanThreads, 0, st>>>(out, M, b.blockSums64);
    return 3;
}

int launchMergeEntries(const BackendBuffers& b, const HistEntry* gathered, const uint32_t* counts, int countStride, int nranks, uint32_t nMax, cudaStream_t st) {
    if (!nMax) return 0;
    mergeEntriesKernel<<<dim3((nMax + 255) / 256, static_cast<unsigned>(nranks)), 256, 0, st>>>(gathered, counts, countStride, nranks, nMax, b.hist, b.firstPos, b.maxSym);
    return 1;
}

int launchResetHistogram(const BackendBuffers& b, uint32_t nEntries, cudaStream_t st) {
    const uint32_t n = nEntries > kMaxSeq ? nEntries : kMaxSeq;
    resetHistogramKernel<<<(n + 255) / 256, 256, 0, st>>>(b.entries, nEntries, b.firstPos, b.maxSym, b.nEntries);
    return 1;
}

}  // namespace mnenc
