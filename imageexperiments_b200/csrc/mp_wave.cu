// Level-synchronous matching pursuit (sm_100a). One greedy iteration of ALL active block-channels is one round of
// kernels; between rounds the residuals live in HBM / L2. What the reference does per block-channel
// (CalcMPDynamicFast, MatchingPursuit.cpp:76-107, with dynamicBasisFast, CompressedImage.cpp:252-289, materialising the
// grown dictionary every iteration) becomes, per iteration r:
//
//   waveBaseKernel     scores of the 510 base atoms for tiles of 128 active block-channels: tcgen05.mma kind::f16,
//                      A = 128 residual rows (fp16), B = the base dictionary image (N = 256 per pass), D in TMEM with the
//                      block-channel on the TMEM lane, so a block-channel's 512 scores belong to ONE thread: maximum and
//                      the 2-eps candidate filter are thread-local loops, no cross-lane reduction
//   waveScanKernel     the KLT groups appended so far, as (block-channel, group) items sorted by (channel, atom): counting
//   waveBinFillKernel  sort = count (by the previous decide pass), scan, fill
//   waveDetailKernel   per (channel, atom) bin, tiles of 128 member block-channels against the bin's 64-row group image:
//                      the group is loaded ONCE per tile instead of once per block-channel (the streamed-group traffic of
//                      the slot-persistent kernel was 18.7 GB per 16 MP image)
//   waveDecideKernel   exact re-score of the candidates in the oracle's operation order (SelectFast,
//                      MatchingPursuit.cpp:27-37; Eigen GEMV without FMA, GeneralMatrixVector.h:108-215), "first strictly
//                      greater wins" by a 64-bit key (Visitor.h:227), quantise, update (MatchingPursuit.cpp:86-104),
//                      append the survivors to the next iteration's list
//
// Pre-filter bound and exactness argument are those of the fp16 tensor-core pre-filter (mnenc_create): every row whose
// approximate |score| is within 2 eps of the largest approximate |score| seen for the block-channel is a candidate; thresholds
// taken from PARTIAL maxima (one 256-atom half, the base atoms only) are lower, i.e. they only add candidates.
#include "mnenc_internal.h"
#include "mp_tc_common.cuh"

#include <algorithm>
#include <atomic>
#include <cstdio>

namespace mnenc {

namespace {

using namespace tc;

constexpr uint32_t kImgBytes = 8192;  // one fp16 group image: 64 rows x 64
constexpr uint32_t kSbo = 1024, kLbo = 128, kKStep = 256;
constexpr int kRStr = 68;             // floats per residual in shared memory (bank spread for 8 threads per block-channel)
constexpr uint32_t kBcMask = 0x07FFFFFFu;

__host__ __device__ constexpr uint32_t idescF16(uint32_t M, uint32_t N) { return (1u << 4) | ((N >> 3) << 17) | ((M >> 4) << 24); }

__device__ __forceinline__ float twoEpsOf(const WaveParams& p, float ss) { return 2.0f * (p.epsRel16 * sqrtf(ss) * 1.0001f + p.epsAbs16); }

// A candidate that found no free slot (near-tie floods: ~1 % of the block-channel iterations) goes to a queue that
// waveOverflowKernel re-scores before the decide pass: the scanning warps never wait for a dependent chain of loads.
__device__ __noinline__ void queueOverflow(const WaveParams& p, uint32_t bc, uint32_t code) {
    const uint32_t pos = atomicAdd(p.ovfCount, 1u);
    if (pos < p.ovfCap) {
        p.ovf[pos] = make_uint2(bc, code);
    } else {  // queue full: re-score on the spot
        const float a = exactDot(p.rows32 + static_cast<size_t>(code >> 12) * kN, p.res32 + static_cast<size_t>(bc) * kN);
        atomicMax(&p.best[bc], makeKey(a, static_cast<int>(code & 0xFFFu)));
    }
}
// Thread-local candidate collector: the scan loops only fill registers (no memory round trip, no divergent atomics); the
// candidates leave the thread in one convergent step after the loop.
struct CandLocal {
    uint32_t c[kCandSlotsDetail];
    int n = 0;
    __device__ __forceinline__ void add(const WaveParams& p, uint32_t bc, uint32_t idx, uint32_t rowId) {
        const uint32_t code = idx | (rowId << 12);
        if (n == 0) c[0] = code;
        else if (n == 1) c[1] = code;
        else if (n == 2) c[2] = code;
        else if (n == 3) c[3] = code;
        else queueOverflow(p, bc, code);
        ++n;
    }
};

// ------------------------------------------------------------------------------------------------
// Block gather + YUVFromRGB (CompressedImage.cpp:595-611, misc.cpp:12-21): 8 threads per image block, thread l8 owns
// pixel row l8; all three channels from one read of the pixels.
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t rawByte(const uint32_t (&w)[6], int k) { return (w[k >> 2] >> (8 * (k & 3))) & 0xFFu; }

__global__ void __launch_bounds__(256) waveInitKernel(const __grid_constant__ WaveParams p) {
    __shared__ uint32_t sCount, sBase;
    const uint32_t frameBlocksReal = static_cast<uint32_t>(p.nbx) * static_cast<uint32_t>(p.nby);
    const int tid = threadIdx.x, lane = tid & 31, l8 = tid & 7;
    const unsigned groupMask = 0xFFu << (lane & 24);
    const uint32_t blk = blockIdx.x * 32u + static_cast<uint32_t>(tid >> 3);
    if (tid == 0) sCount = 0;
    __syncthreads();
    uint32_t w[6] = {0u, 0u, 0u, 0u, 0u, 0u};
    uint32_t local = blk;
    const uint8_t* image = p.rgb;
    bool valid = blk < p.nbc / 3u;
    if (p.frames && valid) {
        const uint32_t f = blk / p.frameBlocks;
        local = blk - f * p.frameBlocks;
        image = p.frames[f];
    }
    valid = valid && local < frameBlocksReal;
    if (valid) {
        const int bx = static_cast<int>(local / static_cast<uint32_t>(p.nby)), by = static_cast<int>(local - static_cast<uint32_t>(bx) * p.nby);
        const int x0 = bx * 8, y = by * 8 + l8;
        if (y < p.height) {
            const uint8_t* row = image + static_cast<size_t>(y) * p.stride + 3 * static_cast<size_t>(x0);
            if (x0 + 8 <= p.width && (reinterpret_cast<uintptr_t>(row) & 7u) == 0) {
                const uint2* r2 = reinterpret_cast<const uint2*>(row);
                const uint2 a = __ldg(r2), b = __ldg(r2 + 1), c = __ldg(r2 + 2);
                w[0] = a.x, w[1] = a.y, w[2] = b.x, w[3] = b.y, w[4] = c.x, w[5] = c.y;
            } else {
#pragma unroll
                for (int k = 0; k < 24; ++k)
                    if (x0 + k / 3 < p.width) w[k >> 2] |= static_cast<uint32_t>(__ldg(row + k)) << (8 * (k & 3));
            }
        }
    }
    uint32_t rank[3] = {kWaveNone, kWaveNone, kWaveNone};
    unsigned long long zeroRows = 0;
#pragma unroll
    for (int ch = 0; ch < 3; ++ch) {
        float r[8];
        float ss = 0.0f;
        bool nz = false;
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            r[i] = yuvChannel(ch, static_cast<double>(rawByte(w, 3 * i)), static_cast<double>(rawByte(w, 3 * i + 1)),
                              static_cast<double>(rawByte(w, 3 * i + 2)));
            ss += r[i] * r[i];
            nz |= r[i] != 0.0f;
        }
        ss += __shfl_xor_sync(groupMask, ss, 4);
        ss += __shfl_xor_sync(groupMask, ss, 2);
        ss += __shfl_xor_sync(groupMask, ss, 1);
        const bool nonZero = (__ballot_sync(groupMask, nz) & groupMask) != 0u;
        if (!valid) continue;
        const uint32_t bc = 3u * blk + static_cast<uint32_t>(ch);
        if (nonZero) {
            float4* d32 = reinterpret_cast<float4*>(p.res32 + static_cast<size_t>(bc) * kN + 8 * l8);
            d32[0] = make_float4(r[0], r[1], r[2], r[3]);
            d32[1] = make_float4(r[4], r[5], r[6], r[7]);
            uint32_t h[4];
#pragma unroll
            for (int i = 0; i < 4; ++i)
                h[i] = static_cast<uint32_t>(__half_as_ushort(__float2half_rn(r[2 * i]))) |
                       (static_cast<uint32_t>(__half_as_ushort(__float2half_rn(r[2 * i + 1]))) << 16);
            *reinterpret_cast<uint4*>(p.res16 + static_cast<size_t>(bc) * kN + 8 * l8) = make_uint4(h[0], h[1], h[2], h[3]);
        }
        if (l8 == 0) {
            p.best[bc] = 0ULL;
            p.ncandDetail[bc] = 0u;
            p.amax[bc] = 0u;
            if (nonZero) {
                p.state[bc] = make_uint4(static_cast<uint32_t>(p.nBase) << 16, 0u, __float_as_uint(twoEpsOf(p, ss)), 0u);
                rank[ch] = atomicAdd(&sCount, 1u);
            } else {  // r == 0: every score is +0, atom 0 wins, its coefficient quantises to 0 (MatchingPursuit.cpp:98-99)
                p.counts[bc] = 0;
                zeroRows += static_cast<unsigned long long>(p.nBase);
            }
        }
    }
    __syncthreads();
    if (tid == 0) sBase = sCount ? atomicAdd(&p.nAct[0], sCount) : 0u;
    __syncthreads();
    if (valid && l8 == 0) {
#pragma unroll
        for (int ch = 0; ch < 3; ++ch)
            if (rank[ch] != kWaveNone) p.act[0][sBase + rank[ch]] = 3u * blk + static_cast<uint32_t>(ch);
    }
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) zeroRows += __shfl_xor_sync(kFullMask, zeroRows, off);
    if (lane == 0 && zeroRows) atomicAdd(&p.stats[0], zeroRows);
}

// ------------------------------------------------------------------------------------------------
// Base scores. 256 threads = two independent groups of four warps, 2 CTAs per SM (256 TMEM columns each, 128 per group).
// A group owns a tile of 128 active block-channels: thread t of the group stages residual row t (cp.async straight into the
// canonical operand image), and later scans TMEM lane t, i.e. the scores of ITS block-channel, so block-channel index and
// eps never leave the thread's registers. The 512 dictionary rows are scored in eight parts of 64 atoms (N = 64 per
// tcgen05.mma) that ping-pong between the group's two 64-column TMEM buffers: while the group scans part e, the tensor core
// fills the other buffer with part e + 1. One thread sees the eight parts of its block-channel in ascending atom order, so
// its threshold (running maximum - 2 eps) is that of a sequential scan, and nothing in the loop is a CTA-wide barrier:
// mbarriers between the group's issuing warp and its four warps, one 128-thread named barrier per tile. The next tile's rows
// are fetched during the scan of the last part.
// Candidates go straight to the block-channel's slots (plain stores, the thread is their only writer). A part whose maximum
// exceeds the running maximum by more than 2 eps makes every earlier candidate stale (their scores are below the new
// threshold): the slot counter restarts at 0, which keeps the list at the candidates of the final threshold plus near-ties.
// ------------------------------------------------------------------------------------------------
constexpr int kBaseThreads = 256;
constexpr uint32_t kPartAtoms = 64, kParts = 8;
struct __align__(1024) BaseSmem {
    __half D16[kRowsPad * kN];                    // 64 KB base dictionary image (B operand)
    alignas(1024) __half A[2][kWaveTile * kN];    // one 16 KB residual tile (A operand) per group, canonical K-major image
    unsigned long long mbarDict, full[2][2], free_[2][2];
    uint32_t tmemBase;
};

__device__ __forceinline__ void groupBarrier(int g) {  // immediate barrier ids: a register id makes ptxas reserve all 16
    if (g == 0) asm volatile("bar.sync 1, 128;" ::: "memory");
    else asm volatile("bar.sync 2, 128;" ::: "memory");
}

// One 32-column chunk held in registers: running maximum, stale-list reset, and - only if the chunk reaches the threshold - its
// candidates. The hit mask is built branch-free (the lanes of a warp hit different columns: predicated per-column blocks would
// all be issued), then the few set bits are walked.
__device__ __forceinline__ void scanChunk(const WaveParams& p, const uint32_t (&raw)[32], float eps, bool valid, int atom0, uint32_t bc,
                                          uint32_t* slots, float& mrun, int& n) {
    float m[4] = {0.0f, 0.0f, 0.0f, 0.0f};
#pragma unroll
    for (int j = 0; j < 32; j += 8)
#pragma unroll
        for (int q = 0; q < 4; ++q) m[q] = fmaxf(m[q], fmaxf(fabsf(__uint_as_float(raw[j + 2 * q])), fabsf(__uint_as_float(raw[j + 2 * q + 1]))));
    const float mc = fmaxf(fmaxf(m[0], m[1]), fmaxf(m[2], m[3]));
    if (mc - eps > mrun) n = 0;  // every earlier candidate scored <= mrun < the new threshold: stale
    mrun = fmaxf(mrun, mc);
    const float thr = mrun - eps;
    if (valid && mc >= thr) {
        uint32_t mask = 0u;
#pragma unroll
        for (int j = 0; j < 32; ++j)
            if (fabsf(__uint_as_float(raw[j])) >= thr) mask |= 1u << j;
        while (mask) {
            const int atom = atom0 + __ffs(mask) - 1;
            mask &= mask - 1u;
            if (atom < p.nBase) {
                const uint32_t code = static_cast<uint32_t>(atom) | (static_cast<uint32_t>(atom) << 12);
                if (n < kCandSlots) slots[n] = code;
                else queueOverflow(p, bc, code);
                ++n;
            }
        }
    }
}

__global__ void __launch_bounds__(kBaseThreads, 2) waveBaseKernel(const __grid_constant__ WaveParams p, const int iter) {
    extern __shared__ __align__(1024) unsigned char smemRaw[];
    BaseSmem& s = *reinterpret_cast<BaseSmem*>(smemRaw);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int g = warp >> 2, wq = warp & 3, tg = tid & 127;
    const uint32_t nA = p.nAct[iter];
    const uint32_t ntiles = (nA + kWaveTile - 1) / kWaveTile;
    if (blockIdx.x * 2u >= ntiles) return;
    const uint32_t* act = p.act[iter & 1];
    const uint32_t firstTile = blockIdx.x * 2u + static_cast<uint32_t>(g), tileStride = gridDim.x * 2u;

    if (tid == 0) {
        mbarInit(&s.mbarDict, 1);
        for (int i = 0; i < 2; ++i)
            for (int b = 0; b < 2; ++b) {
                mbarInit(&s.full[i][b], 1);   // tcgen05.commit
                mbarInit(&s.free_[i][b], 4);  // one arrival per warp of the group
            }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        constexpr uint32_t total = kRowsPad * kN * sizeof(__half);
        mbarExpectTx(&s.mbarDict, total);
        for (uint32_t o = 0; o < total; o += 32768)
            bulkCopy(reinterpret_cast<unsigned char*>(s.D16) + o, reinterpret_cast<const unsigned char*>(p.baseUm16) + o, 32768, &s.mbarDict);
    }
    __syncwarp();
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smemAddr(&s.tmemBase)), "r"(256u) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    fenceBefore();
    __syncthreads();
    fenceAfter();
    const uint32_t tmem = s.tmemBase;
    const uint32_t tcolBase = tmem + (static_cast<uint32_t>(wq * 32) << 16) + static_cast<uint32_t>(g) * 128u;
    const uint32_t aAddr = smemAddr(s.A[g]);
    const uint32_t myDst = aAddr + static_cast<uint32_t>(tg >> 3) * kSbo + static_cast<uint32_t>(tg & 7) * 16u;
    const uint32_t dAddr = smemAddr(s.D16);

    if (firstTile < ntiles) {
        uint32_t bcN = kWaveNone;
        float epsN = 0.0f;
        auto loadIndex = [&](uint32_t tile) {
            const uint32_t e = tile * kWaveTile + static_cast<uint32_t>(tg);
            bcN = e < nA ? __ldg(act + e) : kWaveNone;
        };
        auto fetchRow = [&]() {  // residual row -> canonical image: chunk c of row r at (r / 8) * SBO + c * LBO + (r % 8) * 16
            const bool has = bcN != kWaveNone;
            const unsigned char* src = reinterpret_cast<const unsigned char*>(p.res16 + static_cast<size_t>(has ? bcN : 0u) * kN);
            const uint32_t bytes = has ? 16u : 0u;  // src-size 0: the 16 bytes are zero-filled
#pragma unroll
            for (uint32_t c = 0; c < 8; ++c)
                asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(myDst + c * kLbo), "l"(src + 16u * c), "r"(bytes) : "memory");
            epsN = has ? __uint_as_float(p.state[bcN].z) : 0.0f;
        };
        // issue number j of buffer b (j = 0, 1, ...): waits for the group's j-th release of b, commits to full[g][b]
        uint32_t issued0 = 0u, issued1 = 0u;
        auto issuePart = [&](uint32_t e) {  // warp wq == 0 only, converged
            const uint32_t b = e & 1u;
            const uint32_t issued = b ? issued1 : issued0;
            if (issued) mbarWait(&s.free_[g][b], (issued - 1u) & 1u);
            if (b) ++issued1;
            else ++issued0;
            fenceAfter();
            const unsigned long long descA = umDesc(warpUniform(aAddr), kLbo, kSbo);
            const unsigned long long descB = umDesc(warpUniform(dAddr + e * kPartAtoms * 128u), kLbo, kSbo);
            const uint32_t tm = warpUniform(tmem + static_cast<uint32_t>(g) * 128u + b * kPartAtoms);
#pragma unroll
            for (int ks = 0; ks < 4; ++ks)
                ummaF16Elect(tm, descA + ((ks * kKStep) >> 4), descB + ((ks * kKStep) >> 4), idescF16(128, kPartAtoms), ks ? 1u : 0u);
            ummaCommitElect(&s.full[g][b]);
        };
        auto publishTile = [&]() {  // rows landed -> visible to the tensor core -> first two parts in flight
            asm volatile("cp.async.wait_all;" ::: "memory");
            fenceProxyAsync();
            groupBarrier(g);
            if (wq == 0) {
                issuePart(0);
                issuePart(1);
            }
        };
        loadIndex(firstTile);
        fetchRow();
        if (wq == 0) mbarWait(&s.mbarDict, 0);
        publishTile();
        uint32_t scanned0 = 0u, scanned1 = 0u;
        unsigned nCandStat = 0;
        for (uint32_t tile = firstTile; tile < ntiles; tile += tileStride) {
            const uint32_t myBc = bcN;
            const float eps = epsN;
            const bool more = tile + tileStride < ntiles;
            uint32_t* const slots = p.cand + static_cast<size_t>(myBc != kWaveNone ? myBc : 0u) * kCandSlots;
            float mrun = 0.0f;
            int n = 0;
#pragma unroll 1
            for (uint32_t e = 0; e < kParts; ++e) {
                const uint32_t b = e & 1u;
                mbarWait(&s.full[g][b], (b ? scanned1 : scanned0) & 1u);
                if (b) ++scanned1;
                else ++scanned0;
                fenceAfter();
                if (e == kParts - 2 && more) loadIndex(tile + tileStride);
                if (e == kParts - 1 && more) fetchRow();  // every MMA that reads A[g] has completed: refill it behind this scan
                const uint32_t tcol = tcolBase + b * kPartAtoms;
                uint32_t raw0[32], raw1[32];
                tmemLoad32NoWait(tcol, raw0);
                tmemLoad32NoWait(tcol + 32u, raw1);
                tmemWaitLd();
                const int atom0 = static_cast<int>(e * kPartAtoms);
                scanChunk(p, raw0, eps, myBc != kWaveNone, atom0, myBc, slots, mrun, n);
                scanChunk(p, raw1, eps, myBc != kWaveNone, atom0 + 32, myBc, slots, mrun, n);
                fenceBefore();
                __syncwarp();
                if (lane == 0) mbarArrive(&s.free_[g][b]);
                __syncwarp();
                if (wq == 0 && e + 2 < kParts) issuePart(e + 2);
            }
            if (more) publishTile();
            if (myBc != kWaveNone) {
                p.ncand[myBc] = static_cast<uint32_t>(min(n, kCandSlots));
                p.amax[myBc] = __float_as_uint(mrun);
                nCandStat += static_cast<unsigned>(n);
            }
        }
        nCandStat = __reduce_add_sync(kFullMask, nCandStat);
        if (lane == 0 && nCandStat) atomicAdd(&p.stats[2], static_cast<unsigned long long>(nCandStat));
    }
    fenceBefore();
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(256u) : "memory");
}

// ------------------------------------------------------------------------------------------------
// (channel, atom) bins of the appended groups (counted by the decide pass, offsets by the scan): member lists, one thread per
// active block-channel. A member record carries everything the scoring pass needs about it - block-channel, group position,
// dictionary index of the group's first row, 2 eps, and the threshold the base scores imply - so that pass reads its members
// with one coalesced load.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) waveBinFillKernel(const __grid_constant__ WaveParams p, const int iter) {
    // The groups of atoms 0 and nBase-1 have most members (nearly every block-channel picks them): their positions are drawn from
    // per-CTA counters in shared memory and one global reservation per CTA; all other bins straight from the global cursors, four
    // independent atomics in flight per thread.
    __shared__ uint32_t sCnt[6], sBase[6];
    __shared__ uint8_t sDetRows[kRowsPad];
    const uint32_t nA = p.nAct[iter];
    const uint32_t* act = p.act[iter & 1];
    const int tid = threadIdx.x;
    for (int i = tid; i < kRowsPad; i += 256) sDetRows[i] = p.detRows[i];
    for (uint32_t e0 = blockIdx.x * blockDim.x; e0 < nA; e0 += gridDim.x * blockDim.x) {
        if (tid < 6) sCnt[tid] = 0u;
        __syncthreads();
        const uint32_t e = e0 + tid;
        uint32_t bc = 0, ng = 0, epsHalf = 0, thrBase = 0;
        if (e < nA) {
            bc = act[e];
            const uint4 st = p.state[bc];
            ng = (st.y >> 8) & 0xFFu;
            const float twoEps = __uint_as_float(st.z);
            epsHalf = static_cast<uint32_t>(__half_as_ushort(__float2half_ru(twoEps)));  // rounded UP: never a tighter filter
            thrBase = __float_as_uint(__uint_as_float(p.amax[bc]) - twoEps);
        }
        const uint32_t ch = bc % 3u;
        const uint32_t lastAtom = static_cast<uint32_t>(p.nBase - 1);
        const uint16_t* g = p.groups + static_cast<size_t>(bc) * kMaxK;
        uint32_t first = static_cast<uint32_t>(p.nBase);
        uint32_t hotRank[2] = {kWaveNone, kWaveNone}, hotRec[2] = {0u, 0u}, hotFirst[2] = {0u, 0u};
        for (uint32_t g0 = 0; g0 < ng; g0 += 4) {
            const uint2 gw = *reinterpret_cast<const uint2*>(g + g0);
            uint32_t pos[4], bin[4], fst[4];
#pragma unroll
            for (uint32_t j = 0; j < 4; ++j) {
                pos[j] = kWaveNone;
                if (g0 + j < ng) {
                    const uint32_t atom = ((j < 2 ? gw.x : gw.y) >> (16 * (j & 1))) & 0xFFFFu;
                    bin[j] = ch * static_cast<uint32_t>(p.nBase) + atom;
                    fst[j] = first;
                    first += sDetRows[atom];
                    const int hot = atom == 0u ? 0 : (atom == lastAtom ? 1 : -1);
                    if (hot >= 0 && hotRank[hot] == kWaveNone) {
                        hotRank[hot] = atomicAdd(&sCnt[2 * ch + hot], 1u);
                        hotRec[hot] = bc | ((g0 + j) << 27);
                        hotFirst[hot] = fst[j];
                    } else {
                        pos[j] = atomicAdd(&p.binCursor[bin[j]], 1u);
                    }
                }
            }
#pragma unroll
            for (uint32_t j = 0; j < 4; ++j)
                if (pos[j] != kWaveNone) {
                    const uint32_t at = p.binOffset[bin[j]] + pos[j];
                    p.members[at] = bc | ((g0 + j) << 27);
                    p.memberInfo[at] = make_uint2(fst[j] | (epsHalf << 16), thrBase);
                }
        }
        __syncthreads();
        if (tid < 6) {
            const uint32_t hb = static_cast<uint32_t>(tid >> 1) * static_cast<uint32_t>(p.nBase) + ((tid & 1) ? lastAtom : 0u);
            sBase[tid] = p.binOffset[hb] + (sCnt[tid] ? atomicAdd(&p.binCursor[hb], sCnt[tid]) : 0u);
        }
        __syncthreads();
#pragma unroll
        for (int hot = 0; hot < 2; ++hot)
            if (hotRank[hot] != kWaveNone) {
                const uint32_t at = sBase[2 * ch + hot] + hotRank[hot];
                p.members[at] = hotRec[hot];
                p.memberInfo[at] = make_uint2(hotFirst[hot] | (epsHalf << 16), thrBase);
            }
        __syncthreads();
    }
}

// exclusive scans over the bins (one CTA): member offsets and tile starts; counters cleared for the next iteration
__global__ void __launch_bounds__(1024) waveScanKernel(const __grid_constant__ WaveParams p, const int iter) {
    __shared__ uint32_t sWarp[2][32];
    if (p.nAct[iter] == 0) return;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int nbins = 3 * p.nBase;
    const int b0 = 2 * tid, b1 = 2 * tid + 1;
    const uint32_t c0 = b0 < nbins ? p.binCount[b0] : 0u, c1 = b1 < nbins ? p.binCount[b1] : 0u;
    const uint32_t t0 = (c0 + kWaveTile - 1) / kWaveTile, t1 = (c1 + kWaveTile - 1) / kWaveTile;
    uint32_t sc = c0 + c1, stl = t0 + t1;
#pragma unroll
    for (int off = 1; off < 32; off <<= 1) {
        const uint32_t a = __shfl_up_sync(kFullMask, sc, off), b = __shfl_up_sync(kFullMask, stl, off);
        if (lane >= off) sc += a, stl += b;
    }
    if (lane == 31) sWarp[0][warp] = sc, sWarp[1][warp] = stl;
    __syncthreads();
    if (warp == 0) {
        uint32_t a = sWarp[0][lane], b = sWarp[1][lane];
#pragma unroll
        for (int off = 1; off < 32; off <<= 1) {
            const uint32_t x = __shfl_up_sync(kFullMask, a, off), y = __shfl_up_sync(kFullMask, b, off);
            if (lane >= off) a += x, b += y;
        }
        sWarp[0][lane] = a, sWarp[1][lane] = b;
    }
    __syncthreads();
    const uint32_t exC = sc - (c0 + c1) + (warp ? sWarp[0][warp - 1] : 0u);
    const uint32_t exT = stl - (t0 + t1) + (warp ? sWarp[1][warp - 1] : 0u);
    if (b0 < nbins) {
        p.binOffset[b0] = exC;
        p.binTileStart[b0] = exT;
        p.binCount[b0] = 0u;
        p.binCursor[b0] = 0u;
    }
    if (b1 < nbins) {
        p.binOffset[b1] = exC + c0;
        p.binTileStart[b1] = exT + t0;
        p.binCount[b1] = 0u;
        p.binCursor[b1] = 0u;
    }
    if (tid == 1023) {
        p.binOffset[nbins] = sWarp[0][31];
        p.binTileStart[nbins] = sWarp[1][31];
    }
}

// ------------------------------------------------------------------------------------------------
// Scores of the appended KLT groups. 128 threads, 64 TMEM columns, several CTAs per SM.
// ------------------------------------------------------------------------------------------------
struct __align__(1024) DetailSmem {
    __half A[kWaveTile * kN];                  // 16 KB residual rows of the tile's members
    alignas(1024) unsigned char B[kImgBytes];  // 8 KB group image
    uint32_t tileStart[3 * kRowsPad + 1];
    uint32_t binOffset[3 * kRowsPad + 1];
    unsigned long long mbarB, mbarMma;
    uint32_t tmemBase;
};

// Thread t stages the residual row of member t and later scans TMEM lane t, the scores of that member: the member record stays
// in registers. The next tile's records are fetched while the tensor core works on this one, and its rows (cp.async) and group
// image (TMA) start as soon as this tile's MMAs have completed, i.e. behind the scan; one CTA-wide barrier per tile.
__global__ void __launch_bounds__(128, 5) waveDetailKernel(const __grid_constant__ WaveParams p, const int iter) {
    extern __shared__ __align__(1024) unsigned char smemRaw[];
    DetailSmem& s = *reinterpret_cast<DetailSmem*>(smemRaw);
    if (p.nAct[iter] == 0) return;
    const int tid = threadIdx.x, warp = tid >> 5;
    const int nbins = 3 * p.nBase;
    const uint32_t nT = p.binTileStart[nbins];
    if (blockIdx.x >= nT) return;
    for (int i = tid; i <= nbins; i += 128) s.tileStart[i] = p.binTileStart[i], s.binOffset[i] = p.binOffset[i];
    if (tid == 0) {
        mbarInit(&s.mbarB, 1);
        mbarInit(&s.mbarMma, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smemAddr(&s.tmemBase)), "r"(64u) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    fenceBefore();
    __syncthreads();
    fenceAfter();
    const uint32_t tmem = s.tmemBase;
    const uint32_t laneBase = static_cast<uint32_t>(warp * 32) << 16;
    const uint32_t myDst = smemAddr(s.A) + static_cast<uint32_t>(tid >> 3) * kSbo + static_cast<uint32_t>(tid & 7) * 16u;
    uint32_t mmaPhase = 0, bPhase = 0;
    unsigned nDetStat = 0;

    struct Tile {
        int bin;
        uint32_t mem;   // block-channel | group position << 27, kWaveNone beyond the bin's last member
        uint2 info;     // first dictionary index of the group | fp16(2 eps) << 16, threshold implied by the base scores
    };
    auto describe = [&](uint32_t tile) {  // bin of the tile: the last bin whose first tile is <= tile (empty bins have equal starts)
        int lo = 0, hi = nbins;
        while (hi - lo > 1) {
            const int mid = (lo + hi) >> 1;
            if (s.tileStart[mid] <= tile) lo = mid;
            else hi = mid;
        }
        Tile t;
        t.bin = lo;
        const uint32_t chunk = tile - s.tileStart[lo];
        const uint32_t off = s.binOffset[lo], cnt = s.binOffset[lo + 1] - off;
        const uint32_t n = min(static_cast<uint32_t>(kWaveTile), cnt - chunk * kWaveTile);
        const bool valid = static_cast<uint32_t>(tid) < n;
        t.mem = valid ? __ldg(p.members + off + chunk * kWaveTile + tid) : kWaveNone;
        t.info = valid ? __ldg(p.memberInfo + off + chunk * kWaveTile + tid) : make_uint2(0u, 0u);
        return t;
    };
    auto fetchRow = [&](uint32_t mem) {  // residual row -> canonical image: chunk c of row r at (r / 8) * SBO + c * LBO + (r % 8) * 16
        const bool has = mem != kWaveNone;
        const unsigned char* src = reinterpret_cast<const unsigned char*>(p.res16 + static_cast<size_t>(has ? (mem & kBcMask) : 0u) * kN);
        const uint32_t bytes = has ? 16u : 0u;  // src-size 0: the 16 bytes are zero-filled
#pragma unroll
        for (uint32_t c = 0; c < 8; ++c)
            asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(myDst + c * kLbo), "l"(src + 16u * c), "r"(bytes) : "memory");
    };
    auto fetchGroup = [&](int bin) {
        if (tid == 0) {
            mbarExpectTx(&s.mbarB, kImgBytes);
            bulkCopy(s.B, reinterpret_cast<const unsigned char*>(p.detUm16) + static_cast<size_t>(bin) * kImgBytes, kImgBytes, &s.mbarB);
        }
    };

    Tile cur = describe(blockIdx.x);
    fetchRow(cur.mem);
    fetchGroup(cur.bin);
    bool bPending = true;
    for (uint32_t tile = blockIdx.x; tile < nT; tile += gridDim.x) {
        const bool more = tile + gridDim.x < nT;
        asm volatile("cp.async.wait_all;" ::: "memory");
        fenceProxyAsync();
        fenceBefore();      // this thread's TMEM reads of the previous tile precede the barrier, the next MMAs follow it
        __syncthreads();
        if (warp == 0) {
            if (bPending) mbarWait(&s.mbarB, bPhase);
            fenceAfter();
            const unsigned long long descA = umDesc(warpUniform(smemAddr(s.A)), kLbo, kSbo);
            const unsigned long long descB = umDesc(warpUniform(smemAddr(s.B)), kLbo, kSbo);
            const uint32_t tm = warpUniform(tmem);
#pragma unroll
            for (int ks = 0; ks < 4; ++ks)
                ummaF16Elect(tm, descA + ((ks * kKStep) >> 4), descB + ((ks * kKStep) >> 4), idescF16(128, 64), ks ? 1u : 0u);
            ummaCommitElect(&s.mbarMma);
        }
        if (bPending) bPhase ^= 1u, bPending = false;
        Tile nxt = cur;
        if (more) nxt = describe(tile + gridDim.x);  // in flight while the tensor core works
        mbarWait(&s.mbarMma, mmaPhase);
        mmaPhase ^= 1u;
        fenceAfter();
        if (more) {  // A and B have been consumed: refill them behind the scan
            fetchRow(nxt.mem);
            if (nxt.bin != cur.bin) {
                fetchGroup(nxt.bin);
                bPending = true;
            }
        }
        const bool valid = cur.mem != kWaveNone;
        const uint32_t bc = cur.mem & kBcMask;
        const int ch = cur.bin / p.nBase, atom = cur.bin - ch * p.nBase;
        const int rowsValid = p.detRows[atom];
        const uint32_t first = cur.info.x & 0xFFFFu;
        const float twoEps = __half2float(__ushort_as_half(static_cast<unsigned short>(cur.info.x >> 16)));
        uint32_t raw0[32], raw1[32];
        tmemLoad32NoWait(tmem + laneBase, raw0);
        tmemLoad32NoWait(tmem + laneBase + 32u, raw1);
        tmemWaitLd();
        // rows past the group's height are zero in the image: they score exactly 0 and cannot raise the maximum
        float m[4] = {0.0f, 0.0f, 0.0f, 0.0f};
#pragma unroll
        for (int j = 0; j < 32; j += 4)
#pragma unroll
            for (int q = 0; q < 2; ++q) {
                m[q] = fmaxf(m[q], fmaxf(fabsf(__uint_as_float(raw0[j + 2 * q])), fabsf(__uint_as_float(raw0[j + 2 * q + 1]))));
                m[2 + q] = fmaxf(m[2 + q], fmaxf(fabsf(__uint_as_float(raw1[j + 2 * q])), fabsf(__uint_as_float(raw1[j + 2 * q + 1]))));
            }
        const float m0 = fmaxf(m[0], m[1]), m1 = fmaxf(m[2], m[3]), mAll = fmaxf(m0, m1);
        // within 2 eps of the larger of this group's maximum and the base atoms' maximum
        const float thr = valid ? fmaxf(mAll - twoEps, __uint_as_float(cur.info.y)) : __int_as_float(0x7F800000);
        if (mAll >= thr) {  // most groups stay below the base atoms' maximum
            CandLocal cl;
            uint32_t mask0 = 0u, mask1 = 0u;
            if (m0 >= thr) {
#pragma unroll
                for (int j = 0; j < 32; ++j)
                    if (fabsf(__uint_as_float(raw0[j])) >= thr) mask0 |= 1u << j;
            }
            if (m1 >= thr) {
#pragma unroll
                for (int j = 0; j < 32; ++j)
                    if (fabsf(__uint_as_float(raw1[j])) >= thr) mask1 |= 1u << j;
            }
            const uint32_t rowBase = static_cast<uint32_t>(kRowsPad + (ch * p.nBase + atom) * 64);
            while (mask0 | mask1) {
                int row;
                if (mask0) row = __ffs(mask0) - 1, mask0 &= mask0 - 1u;
                else row = 32 + __ffs(mask1) - 1, mask1 &= mask1 - 1u;
                if (row < rowsValid) cl.add(p, bc, first + static_cast<uint32_t>(row), rowBase + static_cast<uint32_t>(row));
            }
            if (cl.n) {  // one atomic per thread with candidates
                nDetStat += static_cast<unsigned>(cl.n);
                const uint32_t take = static_cast<uint32_t>(min(cl.n, kCandSlotsDetail));
                const uint32_t pos = atomicAdd(&p.ncandDetail[bc], take);
                uint32_t* slots = p.candDetail + static_cast<size_t>(bc) * kCandSlotsDetail;
                for (uint32_t j = 0; j < take; ++j) {
                    if (pos + j < static_cast<uint32_t>(kCandSlotsDetail)) slots[pos + j] = cl.c[j];
                    else queueOverflow(p, bc, cl.c[j]);
                }
            }
        }
        cur = nxt;
    }
    nDetStat = __reduce_add_sync(kFullMask, nDetStat);
    if ((tid & 31) == 0 && nDetStat) atomicAdd(&p.stats[4], static_cast<unsigned long long>(nDetStat));  // diagnostics
    fenceBefore();
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(64u) : "memory");
}

// candidates that found no slot: exact re-score, one per thread
__global__ void __launch_bounds__(256) waveOverflowKernel(const __grid_constant__ WaveParams p) {
    const uint32_t n = min(*p.ovfCount, p.ovfCap);
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const uint2 e = p.ovf[i];
        const float a = exactDot(p.rows32 + static_cast<size_t>(e.y >> 12) * kN, p.res32 + static_cast<size_t>(e.x) * kN);
        atomicMax(&p.best[e.x], makeKey(a, static_cast<int>(e.y & 0xFFFu)));
    }
    if (blockIdx.x == 0 && threadIdx.x == 0 && n) atomicAdd(&p.stats[3], static_cast<unsigned long long>(n));  // diagnostics
}

// ------------------------------------------------------------------------------------------------
// Exact decision + step C (MatchingPursuit.cpp:86-104), 8 threads per block-channel, 32 block-channels per CTA pass.
// The candidate rows of a block-channel are fetched together, 32 bytes per thread and row (one L2 round trip for all of
// them), then re-scored from shared memory in the oracle's operation order, one candidate per thread.
// Survivors go to the next iteration's list (batched in shared memory, one reservation per CTA and flush) and are counted
// into the (channel, atom) bins of their groups.
// ------------------------------------------------------------------------------------------------
constexpr int kRowsPerRound = 4;   // candidate rows of a block-channel staged per round (about 3 candidates on average)
__global__ void __launch_bounds__(256, 4) waveDecideKernel(const __grid_constant__ WaveParams p, const int iter) {
    __shared__ __align__(16) float sR[32 * kRStr];                       // residuals
    __shared__ __align__(16) float sRow[32 * kRowsPerRound * kRStr];     // candidate rows
    __shared__ uint32_t sCand[32][kCandSlots + kCandSlotsDetail];
    __shared__ uint32_t sList[8][32];  // survivors, one list per warp: flushed by the warp itself, no CTA barrier in the loop
    __shared__ uint32_t sDc[6];        // bins of the groups of atoms 0 and nBase-1: most members, counted per CTA
    __shared__ uint8_t sDetRows[kRowsPad];
    const uint32_t nA = p.nAct[iter];
    if (nA == 0) return;
    const uint32_t* act = p.act[iter & 1];
    uint32_t* actNext = p.act[(iter + 1) & 1];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, l8 = tid & 7, grp = tid >> 3;
    const unsigned groupMask = 0xFFu << (lane & 24);
    const int leaderLane = lane & 24;
    float* const myR = sR + grp * kRStr;
    float* const myRows = sRow + grp * kRowsPerRound * kRStr;
    unsigned long long statRows = 0, statPicks = 0;
    if (tid < 6) sDc[tid] = 0;
    for (int i = tid; i < kRowsPad; i += 256) sDetRows[i] = p.detRows[i];
    if (blockIdx.x == 0 && tid == 0) *p.ovfCount = 0u;  // the overflow queue of this iteration has been consumed
    __syncthreads();

    uint32_t nList = 0;  // warp-uniform
    auto flushWarp = [&]() {
        uint32_t base = 0;
        if (lane == 0) base = atomicAdd(&p.nAct[iter + 1], nList);
        base = __shfl_sync(kFullMask, base, 0);
        if (static_cast<uint32_t>(lane) < nList) actNext[base + lane] = sList[warp][lane];
        nList = 0;
        __syncwarp();
    };

    const uint32_t stride = gridDim.x * 32u;
    uint32_t e = blockIdx.x * 32u + static_cast<uint32_t>(grp);
    uint32_t bcNext = e < nA ? act[e] : 0u;
    for (uint32_t base = blockIdx.x * 32u; base < nA; base += stride, e += stride) {
        const bool valid = e < nA;
        const uint32_t bc = bcNext;
        if (e + stride < nA) bcNext = act[e + stride];  // next pass's block-channel: its loads start without this round trip
        bool survivor = false;
        if (valid) {
            const int ch = static_cast<int>(bc % 3u);
            const uint4 st = p.state[bc];
            const uint32_t nBaseCand = min(p.ncand[bc], static_cast<uint32_t>(kCandSlots));
            const uint32_t nDet = min(p.ncandDetail[bc], static_cast<uint32_t>(kCandSlotsDetail));
            const uint32_t c0 = p.cand[static_cast<size_t>(bc) * kCandSlots + l8];
            const uint32_t cd = l8 < kCandSlotsDetail ? p.candDetail[static_cast<size_t>(bc) * kCandSlotsDetail + l8] : 0u;
            uint16_t* const g = p.groups + static_cast<size_t>(bc) * kMaxK;
            const uint2 gw = *reinterpret_cast<const uint2*>(g + 4 * l8);  // groups 4 l8 .. 4 l8 + 3 of the block-channel
            unsigned long long key = l8 == 0 ? p.best[bc] : 0ULL;          // best of the overflow queue, if any
            uint32_t keyCode = kWaveNone;
            const int prev = static_cast<int>(st.x & 0xFFFFu), height = static_cast<int>(st.x >> 16);
            const int reps = static_cast<int>(st.y & 0xFFu), ng = static_cast<int>((st.y >> 8) & 0xFFu);
            const float4* src = reinterpret_cast<const float4*>(p.res32 + static_cast<size_t>(bc) * kN + 8 * l8);
            const float4 r0 = src[0], r1 = src[1];
            reinterpret_cast<float4*>(myR + 8 * l8)[0] = r0;
            reinterpret_cast<float4*>(myR + 8 * l8)[1] = r1;
            // the block-channel's candidates: base atoms first, then rows of its appended groups
            const int nCand = static_cast<int>(nBaseCand + nDet);
            if (static_cast<uint32_t>(l8) < nBaseCand) sCand[grp][l8] = c0;
            if (static_cast<uint32_t>(l8) < nDet) sCand[grp][nBaseCand + l8] = cd;
            __syncwarp(groupMask);
            for (int r = 0; r < nCand; r += kRowsPerRound) {
                const int nr = min(kRowsPerRound, nCand - r);
                // 32 bytes per thread and row straight into shared memory (cp.async: no staging registers, all rows in flight at once)
#pragma unroll
                for (int j = 0; j < kRowsPerRound; ++j)
                    if (j < nr) {
                        const float* rowp = p.rows32 + static_cast<size_t>(sCand[grp][r + j] >> 12) * kN + 8 * l8;
                        const uint32_t dst = smemAddr(myRows + j * kRStr + 8 * l8);
                        asm volatile("cp.async.ca.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(rowp) : "memory");
                        asm volatile("cp.async.ca.shared.global [%0], [%1], 16;" ::"r"(dst + 16u), "l"(rowp + 4) : "memory");
                    }
                asm volatile("cp.async.wait_all;" ::: "memory");
                __syncwarp(groupMask);
                if (l8 < nr) {  // acc = fl(acc + fl(d * r)), j ascending, never fused (tc::exactDot)
                    const float4* dp = reinterpret_cast<const float4*>(myRows + l8 * kRStr);
                    const float4* rp = reinterpret_cast<const float4*>(myR);
                    float acc = 0.0f;
#pragma unroll
                    for (int c = 0; c < 16; ++c) {
                        const float4 d4 = dp[c], r4 = rp[c];
                        acc = __fadd_rn(acc, __fmul_rn(d4.x, r4.x));
                        acc = __fadd_rn(acc, __fmul_rn(d4.y, r4.y));
                        acc = __fadd_rn(acc, __fmul_rn(d4.z, r4.z));
                        acc = __fadd_rn(acc, __fmul_rn(d4.w, r4.w));
                    }
                    const uint32_t code = sCand[grp][r + l8];
                    const unsigned long long k2 = makeKey(acc, static_cast<int>(code & 0xFFFu));
                    if (k2 > key) key = k2, keyCode = code;
                }
                __syncwarp(groupMask);
            }
            unsigned long long kmax = key;
#pragma unroll
            for (int off = 4; off > 0; off >>= 1) kmax = max(kmax, __shfl_xor_sync(groupMask, kmax, off));
            const int idx = kmax ? keyIndex(kmax) : 0;
            const float proj = kmax ? keyValue(kmax) : 0.0f;
            // row of the winner: the thread that scored it knows it (a winner from the overflow queue is looked up below)
            const unsigned owners = (__ballot_sync(groupMask, kmax != 0ULL && key == kmax && keyCode != kWaveNone) >> leaderLane) & 0xFFu;
            const uint32_t winCode = __shfl_sync(groupMask, keyCode, leaderLane + (owners ? __ffs(owners) - 1 : 0));
            const float qf = p.quant[ch * kMaxK + reps];
            const int qi = static_cast<int>(roundf(__fdiv_rn(proj, qf)));
            const uint32_t deltaId = (reps ? zigzag(idx - prev) : static_cast<uint32_t>(idx)) & 0xFFFFu;
            statRows += (l8 == 0) ? height : 0;
            int count = -1;
            if (qi == 0) {
                count = reps;
            } else {
                const float c = __fmul_rn(qf, static_cast<float>(qi));
                uint32_t rowId = static_cast<uint32_t>(idx);
                if (owners) {
                    rowId = winCode >> 12;
                } else if (idx >= p.nBase) {
                    int hh = p.nBase;
                    for (int k = 0; k < ng; ++k) {
                        const int atom = g[k];
                        const int rows = sDetRows[atom];
                        if (idx < hh + rows) {
                            rowId = static_cast<uint32_t>(kRowsPad + (ch * p.nBase + atom) * 64 + (idx - hh));
                            break;
                        }
                        hh += rows;
                    }
                }
                const float4* rowp = reinterpret_cast<const float4*>(p.rows32 + static_cast<size_t>(rowId) * kN) + 2 * l8;
                const float4 d0 = __ldg(rowp), d1 = __ldg(rowp + 1);  // fetched a moment ago as a candidate row: L1 / L2 hit
                float r[8];
                r[0] = __fsub_rn(r0.x, __fmul_rn(c, d0.x));
                r[1] = __fsub_rn(r0.y, __fmul_rn(c, d0.y));
                r[2] = __fsub_rn(r0.z, __fmul_rn(c, d0.z));
                r[3] = __fsub_rn(r0.w, __fmul_rn(c, d0.w));
                r[4] = __fsub_rn(r1.x, __fmul_rn(c, d1.x));
                r[5] = __fsub_rn(r1.y, __fmul_rn(c, d1.y));
                r[6] = __fsub_rn(r1.z, __fmul_rn(c, d1.z));
                r[7] = __fsub_rn(r1.w, __fmul_rn(c, d1.w));
                float ss = 0.0f;
                bool nz = false;
#pragma unroll
                for (int i = 0; i < 8; ++i) ss += r[i] * r[i], nz |= r[i] != 0.0f;
                ss += __shfl_xor_sync(groupMask, ss, 4);
                ss += __shfl_xor_sync(groupMask, ss, 2);
                ss += __shfl_xor_sync(groupMask, ss, 1);
                const bool nonZero = (__ballot_sync(groupMask, nz) & groupMask) != 0u;
                const bool isBase = idx < p.nBase;
                const int ngNew = ng + (isBase ? 1 : 0);
                const int heightNew = height + (isBase ? static_cast<int>(sDetRows[idx]) : 0);
                if (l8 == 0) p.pairs[static_cast<size_t>(bc) * p.K + reps] = deltaId | ((zigzag(qi) & 0xFFFFu) << 16);
                if (reps + 1 == p.K) {
                    count = p.K;
                } else if (!nonZero) {  // r == 0 now: the next iteration picks atom 0 with a coefficient that quantises to 0
                    count = reps + 1;
                    statRows += (l8 == 0) ? heightNew : 0;
                } else {
                    float4* d32 = reinterpret_cast<float4*>(p.res32 + static_cast<size_t>(bc) * kN + 8 * l8);
                    d32[0] = make_float4(r[0], r[1], r[2], r[3]);
                    d32[1] = make_float4(r[4], r[5], r[6], r[7]);
                    uint32_t hw[4];
#pragma unroll
                    for (int i = 0; i < 4; ++i)
                        hw[i] = static_cast<uint32_t>(__half_as_ushort(__float2half_rn(r[2 * i]))) |
                                (static_cast<uint32_t>(__half_as_ushort(__float2half_rn(r[2 * i + 1]))) << 16);
                    *reinterpret_cast<uint4*>(p.res16 + static_cast<size_t>(bc) * kN + 8 * l8) = make_uint4(hw[0], hw[1], hw[2], hw[3]);
                    // the survivor's groups (those it had + the one just appended) count into next iteration's bins
                    auto countBin = [&](int atom) {
                        if (atom == 0) atomicAdd(&sDc[2 * ch], 1u);
                        else if (atom == p.nBase - 1) atomicAdd(&sDc[2 * ch + 1], 1u);
                        else atomicAdd(&p.binCount[ch * p.nBase + atom], 1u);
                    };
#pragma unroll
                    for (int j = 0; j < 4; ++j)
                        if (4 * l8 + j < ng) countBin(static_cast<int>(((j < 2 ? gw.x : gw.y) >> (16 * (j & 1))) & 0xFFFFu));
                    if (l8 == 0) {
                        if (isBase) {
                            countBin(idx);
                            g[ng] = static_cast<uint16_t>(idx);
                        }
                        p.state[bc] = make_uint4(static_cast<uint32_t>(idx) | (static_cast<uint32_t>(heightNew) << 16),
                                                 static_cast<uint32_t>(reps + 1) | (static_cast<uint32_t>(ngNew) << 8),
                                                 __float_as_uint(twoEpsOf(p, ss)), 0u);
                        p.amax[bc] = 0u;
                        survivor = true;
                    }
                }
            }
            if (l8 == 0) {
                p.best[bc] = 0ULL;
                p.ncandDetail[bc] = 0u;
                if (count >= 0) {
                    p.counts[bc] = static_cast<uint16_t>(count);
                    statPicks += static_cast<unsigned long long>(count);
                }
            }
        }
        // survivors -> the warp's list -> next iteration's active list
        __syncwarp();
        const unsigned surv = __ballot_sync(kFullMask, survivor);
        if (surv) {
            if (survivor) sList[warp][nList + __popc(surv & ((1u << lane) - 1u))] = bc;
            nList += static_cast<uint32_t>(__popc(surv));
            __syncwarp();
            if (nList > 28u) flushWarp();
        }
    }
    if (nList) flushWarp();
    __syncthreads();
    if (tid < 6 && sDc[tid]) atomicAdd(&p.binCount[(tid >> 1) * p.nBase + ((tid & 1) ? p.nBase - 1 : 0)], sDc[tid]);
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) {
        statRows += __shfl_xor_sync(kFullMask, statRows, off);
        statPicks += __shfl_xor_sync(kFullMask, statPicks, off);
    }
    if (lane == 0) {
        if (statRows) atomicAdd(&p.stats[0], statRows);
        if (statPicks) atomicAdd(&p.stats[1], statPicks);
    }
}

}  // namespace

size_t waveMembersCapacity(uint32_t nbc, int K) { return static_cast<size_t>(nbc) * static_cast<size_t>(std::max(K - 1, 1)); }

cudaError_t launchWaveInit(const WaveParams& p, cudaStream_t st) {
    const uint32_t nblocks = p.nbc / 3;
    if (nblocks == 0) return cudaSuccess;
    waveInitKernel<<<(nblocks + 31) / 32, 256, 0, st>>>(p);
    return cudaGetLastError();
}

cudaError_t launchWaveIteration(const WaveParams& p, int iter, uint32_t nActHint, int numSMs, cudaStream_t st, int* launches) {
    static std::atomic<bool> configuredOn[64];  // the attribute is per device (one process may drive several GPUs)
    int device = 0;
    if (cudaError_t e = cudaGetDevice(&device); e != cudaSuccess) return e;
    std::atomic<bool>& configured = configuredOn[device & 63];
    const size_t smemBase = sizeof(BaseSmem) + 1024, smemDetail = sizeof(DetailSmem) + 1024;
    if (!configured.load(std::memory_order_acquire)) {
        cudaError_t e = cudaFuncSetAttribute(waveBaseKernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smemBase));
        if (e != cudaSuccess) return e;
        e = cudaFuncSetAttribute(waveDetailKernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smemDetail));
        if (e != cudaSuccess) return e;
        configured.store(true, std::memory_order_release);
    }
    if (nActHint == 0) return cudaSuccess;
    const unsigned sms = static_cast<unsigned>(numSMs);
    const unsigned tiles = (nActHint + kWaveTile - 1) / kWaveTile;
    waveBaseKernel<<<std::min(tiles, 2u * sms), kBaseThreads, smemBase, st>>>(p, iter);
    int n = 1;
    if (iter > 0) {
        const unsigned binGrid = (nActHint + 255u) / 256u;  // one pass per thread: the chain of dependent loads is the cost, not the work
        waveScanKernel<<<1, 1024, 0, st>>>(p, iter);
        waveBinFillKernel<<<binGrid, 256, 0, st>>>(p, iter);
        // at most one partial tile per bin on top of the full ones
        const unsigned long long items = static_cast<unsigned long long>(nActHint) * static_cast<unsigned long long>(std::min(iter, kMaxK - 1));
        const unsigned long long maxTiles = items / kWaveTile + static_cast<unsigned long long>(3 * p.nBase);
        waveDetailKernel<<<static_cast<unsigned>(std::min<unsigned long long>(maxTiles, 5ull * sms)), 128, smemDetail, st>>>(p, iter);
        n += 3;
    }
    waveOverflowKernel<<<std::min((nActHint + 1023u) / 1024u + 1u, 2u * sms), 256, 0, st>>>(p);
    waveDecideKernel<<<std::min((nActHint + 31u) / 32u, 16u * sms), 256, 0, st>>>(p, iter);
    n += 2;
    if (launches) *launches += n;
    return cudaGetLastError();
}

}  // namespace mnenc
