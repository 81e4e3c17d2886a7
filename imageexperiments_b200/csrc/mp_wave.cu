// Level-synchronous matching pursuit (sm_100a). One greedy iteration of ALL active block-channels is one round of
// kernels; between rounds the residuals live in HBM / L2. What the reference does per block-channel
// (CalcMPDynamicFast, MatchingPursuit.cpp:76-107, with dynamicBasisFast, CompressedImage.cpp:252-289, materialising the
// grown dictionary every iteration) becomes, per iteration r:
//
//   waveBaseKernel     scores of the 510 base atoms for tiles of 128 active block-channels: tcgen05.mma kind::f16,
//                      A = 128 residual rows (fp16), B = the base dictionary image (N = 256 per pass), D in TMEM with the
//                      block-channel on the TMEM lane, so a block-channel's 512 scores belong to ONE thread: maximum and
//                      the 2-eps candidate filter are thread-local loops, no cross-lane reduction
//   waveBinKernel x2   the KLT groups appended so far, as (block-channel, group) items sorted by (channel, atom): counting
//   waveScanKernel     sort = count, scan, fill
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

// Candidate -> a free slot of its block-channel, or (rare: near-tie floods) re-scored on the spot.
__device__ __noinline__ void pushCandidate(const WaveParams& p, uint32_t bc, uint32_t idx, uint32_t rowId) {
    const uint32_t pos = atomicAdd(&p.ncand[bc], 1u);
    if (pos < static_cast<uint32_t>(kCandSlots)) {
        p.cand[static_cast<size_t>(bc) * kCandSlots + pos] = idx | (rowId << 12);
    } else {
        const float a = exactDot(p.rows32 + static_cast<size_t>(rowId) * kN, p.res32 + static_cast<size_t>(bc) * kN);
        atomicMax(&p.best[bc], makeKey(a, static_cast<int>(idx)));
    }
}

// ------------------------------------------------------------------------------------------------
// Block gather + YUVFromRGB (CompressedImage.cpp:595-611, misc.cpp:12-21): 8 threads per image block, thread l8 owns
// pixel row l8; all three channels from one read of the pixels.
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t rawByte(const uint32_t (&w)[6], int k) { return (w[k >> 2] >> (8 * (k & 3))) & 0xFFu; }

__global__ void __launch_bounds__(256) waveInitKernel(const __grid_constant__ WaveParams p) {
    __shared__ uint32_t sCount, sBase;
    const uint32_t nblocks = static_cast<uint32_t>(p.nbx) * static_cast<uint32_t>(p.nby);
    const int tid = threadIdx.x, lane = tid & 31, l8 = tid & 7;
    const unsigned groupMask = 0xFFu << (lane & 24);
    const uint32_t blk = blockIdx.x * 32u + static_cast<uint32_t>(tid >> 3);
    if (tid == 0) sCount = 0;
    __syncthreads();
    uint32_t w[6] = {0u, 0u, 0u, 0u, 0u, 0u};
    const bool valid = blk < nblocks;
    if (valid) {
        const int bx = static_cast<int>(blk / static_cast<uint32_t>(p.nby)), by = static_cast<int>(blk - static_cast<uint32_t>(bx) * p.nby);
        const int x0 = bx * 8, y = by * 8 + l8;
        if (y < p.height) {
            const uint8_t* row = p.rgb + static_cast<size_t>(y) * p.stride + 3 * static_cast<size_t>(x0);
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
            p.ncand[bc] = 0u;
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
// Base scores. 256 threads, 2 CTAs per SM (256 TMEM columns each).
// ------------------------------------------------------------------------------------------------
struct __align__(1024) BaseSmem {
    __half D16[kRowsPad * kN];                 // 64 KB base dictionary image (B operand)
    alignas(1024) __half A[kWaveTile * kN];    // 16 KB residual tile (A operand), canonical K-major image
    uint32_t bc[kWaveTile];
    float eps[kWaveTile];
    float halfMax[2][kWaveTile];
    unsigned long long mbarDict, mbarMma;
    uint32_t tmemBase;
};

// rows of res16 (128 B each) -> canonical K-major operand image; item = (row, pair of 16-byte chunks)
__device__ __forceinline__ void gatherRows(const uint16_t* res16, const uint32_t* sBc, __half* A, int tid, int nthreads) {
    unsigned char* dst = reinterpret_cast<unsigned char*>(A);
    for (int it = tid; it < kWaveTile * 4; it += nthreads) {
        const int row = it & (kWaveTile - 1), pair = it >> 7;
        const uint32_t bc = sBc[row];
        uint4 v0 = make_uint4(0u, 0u, 0u, 0u), v1 = v0;
        if (bc != kWaveNone) {
            const uint4* src = reinterpret_cast<const uint4*>(res16 + static_cast<size_t>(bc) * kN) + 2 * pair;
            v0 = __ldg(src);
            v1 = __ldg(src + 1);
        }
        unsigned char* d = dst + (row >> 3) * kSbo + (2 * pair) * kLbo + (row & 7) * 16;
        *reinterpret_cast<uint4*>(d) = v0;
        *reinterpret_cast<uint4*>(d + kLbo) = v1;
    }
}

__global__ void __launch_bounds__(256, 2) waveBaseKernel(const __grid_constant__ WaveParams p, const int iter) {
    extern __shared__ __align__(1024) unsigned char smemRaw[];
    BaseSmem& s = *reinterpret_cast<BaseSmem*>(smemRaw);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t nA = p.nAct[iter];
    const uint32_t ntiles = (nA + kWaveTile - 1) / kWaveTile;
    if (blockIdx.x >= ntiles) return;
    const uint32_t* act = p.act[iter & 1];

    if (tid == 0) {
        mbarInit(&s.mbarDict, 1);
        mbarInit(&s.mbarMma, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smemAddr(&s.tmemBase)), "r"(256u) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    fenceBefore();
    __syncthreads();
    fenceAfter();
    const uint32_t tmem = s.tmemBase;
    if (tid == 0) {
        constexpr uint32_t total = kRowsPad * kN * sizeof(__half);
        mbarExpectTx(&s.mbarDict, total);
        for (uint32_t o = 0; o < total; o += 32768)
            bulkCopy(reinterpret_cast<unsigned char*>(s.D16) + o, reinterpret_cast<const unsigned char*>(p.baseUm16) + o, 32768, &s.mbarDict);
    }
    bool dictReady = false;
    uint32_t mmaPhase = 0;
    const int q = warp & 3, h = warp >> 2;
    const int slot = q * 32 + lane;
    const uint32_t laneBase = static_cast<uint32_t>(q * 32) << 16;

    for (uint32_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        if (tid < kWaveTile) {
            const uint32_t e = tile * kWaveTile + tid;
            const uint32_t bc = e < nA ? act[e] : kWaveNone;
            s.bc[tid] = bc;
            s.eps[tid] = bc != kWaveNone ? __uint_as_float(p.state[bc].z) : 0.0f;
        }
        __syncthreads();
        gatherRows(p.res16, s.bc, s.A, tid, 256);
        fenceProxyAsync();
        __syncthreads();
        const uint32_t myBc = s.bc[slot];
        float mrun = 0.0f;
#pragma unroll 1
        for (int half = 0; half < 2; ++half) {
            if (warp == 0) {  // whole warp on the issue path, one elected lane issues (see ummaF16Elect)
                if (!dictReady) mbarWait(&s.mbarDict, 0), dictReady = true;
                fenceAfter();
                const unsigned long long descA = umDesc(warpUniform(smemAddr(s.A)), kLbo, kSbo);
                const unsigned long long descB = umDesc(warpUniform(smemAddr(s.D16) + static_cast<uint32_t>(half) * 256u * 128u), kLbo, kSbo);
                const uint32_t tm = warpUniform(tmem);
#pragma unroll
                for (int ks = 0; ks < 4; ++ks)
                    ummaF16Elect(tm, descA + ((ks * kKStep) >> 4), descB + ((ks * kKStep) >> 4), idescF16(128, 256), ks ? 1u : 0u);
                ummaCommitElect(&s.mbarMma);
            }
            mbarWait(&s.mbarMma, mmaPhase);
            mmaPhase ^= 1u;
            fenceAfter();
            // pass 1: this thread's maximum over its 128 columns (warps 0-3: columns 0-127, warps 4-7: 128-255)
            float m = 0.0f;
#pragma unroll 1
            for (int cc = 0; cc < 4; ++cc) {
                uint32_t raw[32];
                tmemLoad32(tmem + laneBase + static_cast<uint32_t>(h * 128 + cc * 32), raw);
#pragma unroll
                for (int i = 0; i < 32; ++i) m = fmaxf(m, fabsf(__uint_as_float(raw[i])));
            }
            s.halfMax[h][slot] = m;
            __syncthreads();
            mrun = fmaxf(mrun, fmaxf(s.halfMax[0][slot], s.halfMax[1][slot]));
            const float thr = myBc != kWaveNone ? mrun - s.eps[slot] : __int_as_float(0x7F800000);
            // pass 2: rows that can still be the exact maximum
#pragma unroll 1
            for (int cc = 0; cc < 4; ++cc) {
                uint32_t raw[32];
                tmemLoad32(tmem + laneBase + static_cast<uint32_t>(h * 128 + cc * 32), raw);
                const int atom0 = half * 256 + h * 128 + cc * 32;
#pragma unroll
                for (int g = 0; g < 4; ++g) {
                    bool any = false;
#pragma unroll
                    for (int i = 0; i < 8; ++i) any |= fabsf(__uint_as_float(raw[8 * g + i])) >= thr;
                    if (any) {
#pragma unroll
                        for (int i = 0; i < 8; ++i) {
                            const int atom = atom0 + 8 * g + i;
                            if (fabsf(__uint_as_float(raw[8 * g + i])) >= thr && atom < p.nBase)
                                pushCandidate(p, myBc, static_cast<uint32_t>(atom), static_cast<uint32_t>(atom));
                        }
                    }
                }
            }
            fenceBefore();
            __syncthreads();  // TMEM columns and halfMax are reused by the next pass
        }
        if (h == 0 && myBc != kWaveNone) p.amax[myBc] = __float_as_uint(mrun);
    }
    fenceBefore();
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(256u) : "memory");
}

// ------------------------------------------------------------------------------------------------
// (channel, atom) bins of the appended groups: count / fill, one thread per active block-channel.
// ------------------------------------------------------------------------------------------------
template <bool kFill>
__global__ void __launch_bounds__(256) waveBinKernel(const __grid_constant__ WaveParams p, const int iter) {
    const uint32_t nA = p.nAct[iter];
    const uint32_t* act = p.act[iter & 1];
    const int lane = threadIdx.x & 31;
    for (uint32_t e0 = blockIdx.x * blockDim.x; e0 < nA; e0 += gridDim.x * blockDim.x) {
        const uint32_t e = e0 + threadIdx.x;
        uint32_t bc = 0, ng = 0;
        if (e < nA) {
            bc = act[e];
            ng = (p.state[bc].y >> 8) & 0xFFu;
        }
        uint32_t maxNg = ng;
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) maxNg = max(maxNg, __shfl_xor_sync(kFullMask, maxNg, off));
        const uint32_t ch = bc % 3u;
        for (uint32_t gi = 0; gi < maxNg; ++gi) {
            const bool has = gi < ng;
            const unsigned active = __ballot_sync(kFullMask, has);
            if (has) {
                const uint32_t bin = ch * static_cast<uint32_t>(p.nBase) + p.groups[static_cast<size_t>(bc) * kMaxK + gi];
                const unsigned peers = __match_any_sync(active, bin);
                const int leader = __ffs(peers) - 1;
                uint32_t base = 0;
                if (lane == leader) base = atomicAdd(kFill ? &p.binCursor[bin] : &p.binCount[bin], static_cast<uint32_t>(__popc(peers)));
                if (kFill) {
                    base = __shfl_sync(peers, base, leader);
                    const uint32_t rank = __popc(peers & ((1u << lane) - 1u));
                    p.members[p.binOffset[bin] + base + rank] = bc | (gi << 27);
                }
            }
        }
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
    uint32_t member[kWaveTile];
    uint32_t bc[kWaveTile];
    uint32_t tileStart[3 * kRowsPad + 1];
    unsigned long long mbarB, mbarMma;
    uint32_t tmemBase;
};

__global__ void __launch_bounds__(128, 6) waveDetailKernel(const __grid_constant__ WaveParams p, const int iter) {
    extern __shared__ __align__(1024) unsigned char smemRaw[];
    DetailSmem& s = *reinterpret_cast<DetailSmem*>(smemRaw);
    if (p.nAct[iter] == 0) return;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int nbins = 3 * p.nBase;
    const uint32_t nT = p.binTileStart[nbins];
    if (blockIdx.x >= nT) return;
    for (int i = tid; i <= nbins; i += 128) s.tileStart[i] = p.binTileStart[i];
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
    uint32_t mmaPhase = 0, bPhase = 0;
    int curBin = -1;
    bool bPending = false;

    for (uint32_t tile = blockIdx.x; tile < nT; tile += gridDim.x) {
        // bin of this tile: the last bin whose first tile is <= tile (empty bins have equal starts and are skipped)
        int lo = 0, hi = nbins;
        while (hi - lo > 1) {
            const int mid = (lo + hi) >> 1;
            if (s.tileStart[mid] <= tile) lo = mid;
            else hi = mid;
        }
        const int bin = lo;
        const uint32_t chunk = tile - s.tileStart[bin];
        const uint32_t off = p.binOffset[bin], cnt = p.binOffset[bin + 1] - off;
        const uint32_t n = min(static_cast<uint32_t>(kWaveTile), cnt - chunk * kWaveTile);
        const int ch = bin / p.nBase, atom = bin - ch * p.nBase;
        if (bin != curBin) {  // the previous tile's MMAs have completed (mbarMma was waited for), B may be overwritten
            curBin = bin;
            bPending = true;
            if (tid == 0) {
                mbarExpectTx(&s.mbarB, kImgBytes);
                bulkCopy(s.B, reinterpret_cast<const unsigned char*>(p.detUm16) + static_cast<size_t>(bin) * kImgBytes, kImgBytes, &s.mbarB);
            }
        }
        {
            const uint32_t mem = static_cast<uint32_t>(tid) < n ? p.members[off + chunk * kWaveTile + tid] : kWaveNone;
            s.member[tid] = mem;
            s.bc[tid] = mem != kWaveNone ? (mem & kBcMask) : kWaveNone;
        }
        __syncthreads();
        gatherRows(p.res16, s.bc, s.A, tid, 128);
        fenceProxyAsync();
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
        mbarWait(&s.mbarMma, mmaPhase);
        mmaPhase ^= 1u;
        fenceAfter();
        const uint32_t mem = s.member[tid];
        const int rowsValid = p.detRows[atom];
        float m = 0.0f, thr = __int_as_float(0x7F800000);
        uint32_t bc = 0, first = 0;
        if (mem != kWaveNone) {
            bc = mem & kBcMask;
            const uint32_t gi = mem >> 27;
            first = static_cast<uint32_t>(p.nBase);
            const uint16_t* g = p.groups + static_cast<size_t>(bc) * kMaxK;
            for (uint32_t k = 0; k < gi; ++k) first += p.detRows[g[k]];
        }
#pragma unroll 1
        for (int cc = 0; cc < 2; ++cc) {
            uint32_t raw[32];
            tmemLoad32(tmem + laneBase + static_cast<uint32_t>(cc * 32), raw);
#pragma unroll
            for (int i = 0; i < 32; ++i)
                if (cc * 32 + i < rowsValid) m = fmaxf(m, fabsf(__uint_as_float(raw[i])));
        }
        if (mem != kWaveNone) thr = fmaxf(m, __uint_as_float(p.amax[bc])) - __uint_as_float(p.state[bc].z);
#pragma unroll 1
        for (int cc = 0; cc < 2; ++cc) {
            uint32_t raw[32];
            tmemLoad32(tmem + laneBase + static_cast<uint32_t>(cc * 32), raw);
#pragma unroll
            for (int g = 0; g < 4; ++g) {
                bool any = false;
#pragma unroll
                for (int i = 0; i < 8; ++i) any |= fabsf(__uint_as_float(raw[8 * g + i])) >= thr;
                if (any) {
#pragma unroll
                    for (int i = 0; i < 8; ++i) {
                        const int row = cc * 32 + 8 * g + i;
                        if (fabsf(__uint_as_float(raw[8 * g + i])) >= thr && row < rowsValid)
                            pushCandidate(p, bc, first + static_cast<uint32_t>(row),
                                          static_cast<uint32_t>(kRowsPad + (ch * p.nBase + atom) * 64 + row));
                    }
                }
            }
        }
        fenceBefore();
        __syncthreads();
    }
    fenceBefore();
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(64u) : "memory");
}

// ------------------------------------------------------------------------------------------------
// Exact decision + step C (MatchingPursuit.cpp:86-104), 8 threads per block-channel, 32 block-channels per CTA pass.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) waveDecideKernel(const __grid_constant__ WaveParams p, const int iter) {
    __shared__ __align__(16) float sR[32 * kRStr];
    __shared__ uint32_t sSurv, sBase;
    const uint32_t nA = p.nAct[iter];
    if (nA == 0) return;
    const uint32_t* act = p.act[iter & 1];
    uint32_t* actNext = p.act[(iter + 1) & 1];
    const int tid = threadIdx.x, lane = tid & 31, l8 = tid & 7, grp = tid >> 3;
    const unsigned groupMask = 0xFFu << (lane & 24);
    const int leaderLane = lane & 24;
    float* const myR = sR + grp * kRStr;
    unsigned long long statRows = 0, statPicks = 0;
    if (tid == 0) sSurv = 0;
    __syncthreads();

    for (uint32_t base = blockIdx.x * 32u; base < nA; base += gridDim.x * 32u) {
        const uint32_t e = base + static_cast<uint32_t>(grp);
        const bool valid = e < nA;
        uint32_t survRank = kWaveNone, bc = 0;
        if (valid) {
            bc = act[e];
            const int ch = static_cast<int>(bc % 3u);
            const uint4 st = p.state[bc];
            const int prev = static_cast<int>(st.x & 0xFFFFu), height = static_cast<int>(st.x >> 16);
            const int reps = static_cast<int>(st.y & 0xFFu), ng = static_cast<int>((st.y >> 8) & 0xFFu);
            const float4* src = reinterpret_cast<const float4*>(p.res32 + static_cast<size_t>(bc) * kN + 8 * l8);
            const float4 r0 = src[0], r1 = src[1];
            reinterpret_cast<float4*>(myR + 8 * l8)[0] = r0;
            reinterpret_cast<float4*>(myR + 8 * l8)[1] = r1;
            __syncwarp(groupMask);
            // exact re-score of the listed candidates, one per thread
            const uint32_t nc = min(p.ncand[bc], static_cast<uint32_t>(kCandSlots));
            unsigned long long key = 0ULL;
            if (static_cast<uint32_t>(l8) < nc) {
                const uint32_t c = p.cand[static_cast<size_t>(bc) * kCandSlots + l8];
                const float a = exactDot(p.rows32 + static_cast<size_t>(c >> 12) * kN, myR);
                key = makeKey(a, static_cast<int>(c & 0xFFFu));
            }
            if (l8 == 0) key = max(key, p.best[bc]);
#pragma unroll
            for (int off = 4; off > 0; off >>= 1) key = max(key, __shfl_xor_sync(groupMask, key, off));
            const int idx = keyIndex(key);
            const float proj = keyValue(key);
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
                const uint16_t* g = p.groups + static_cast<size_t>(bc) * kMaxK;
                if (idx >= p.nBase) {
                    int hh = p.nBase;
                    for (int k = 0; k < ng; ++k) {
                        const int atom = g[k];
                        const int rows = p.detRows[atom];
                        if (idx < hh + rows) {
                            rowId = static_cast<uint32_t>(kRowsPad + (ch * p.nBase + atom) * 64 + (idx - hh));
                            break;
                        }
                        hh += rows;
                    }
                }
                const float4* rowp = reinterpret_cast<const float4*>(p.rows32 + static_cast<size_t>(rowId) * kN) + 2 * l8;
                const float4 d0 = __ldg(rowp), d1 = __ldg(rowp + 1);
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
                const int heightNew = height + (isBase ? static_cast<int>(p.detRows[idx]) : 0);
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
                    if (l8 == 0) {
                        if (isBase) p.groups[static_cast<size_t>(bc) * kMaxK + ng] = static_cast<uint16_t>(idx);
                        p.state[bc] = make_uint4(static_cast<uint32_t>(idx) | (static_cast<uint32_t>(heightNew) << 16),
                                                 static_cast<uint32_t>(reps + 1) | (static_cast<uint32_t>(ngNew) << 8),
                                                 __float_as_uint(twoEpsOf(p, ss)), 0u);
                        survRank = atomicAdd(&sSurv, 1u);
                    }
                }
            }
            if (l8 == 0) {
                p.best[bc] = 0ULL;
                p.ncand[bc] = 0u;
                if (count >= 0) {
                    p.counts[bc] = static_cast<uint16_t>(count);
                    statPicks += static_cast<unsigned long long>(count);
                }
            }
        }
        __syncthreads();
        if (tid == 0) {
            sBase = sSurv ? atomicAdd(&p.nAct[iter + 1], sSurv) : 0u;
            sSurv = 0;
        }
        __syncthreads();
        if (survRank != kWaveNone) actNext[sBase + survRank] = bc;
    }
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
    static bool configured = false;
    const size_t smemBase = sizeof(BaseSmem) + 1024, smemDetail = sizeof(DetailSmem) + 1024;
    if (!configured) {
        cudaError_t e = cudaFuncSetAttribute(waveBaseKernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smemBase));
        if (e != cudaSuccess) return e;
        e = cudaFuncSetAttribute(waveDetailKernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smemDetail));
        if (e != cudaSuccess) return e;
        configured = true;
    }
    if (nActHint == 0) return cudaSuccess;
    const unsigned sms = static_cast<unsigned>(numSMs);
    const unsigned tiles = (nActHint + kWaveTile - 1) / kWaveTile;
    waveBaseKernel<<<std::min(tiles, 2u * sms), 256, smemBase, st>>>(p, iter);
    int n = 1;
    if (iter > 0) {
        const unsigned binGrid = std::min((nActHint + 255u) / 256u, 8u * sms);
        waveBinKernel<false><<<binGrid, 256, 0, st>>>(p, iter);
        waveScanKernel<<<1, 1024, 0, st>>>(p, iter);
        waveBinKernel<true><<<binGrid, 256, 0, st>>>(p, iter);
        // at most one partial tile per bin on top of the full ones
        const unsigned long long items = static_cast<unsigned long long>(nActHint) * static_cast<unsigned long long>(std::min(iter, kMaxK - 1));
        const unsigned long long maxTiles = items / kWaveTile + static_cast<unsigned long long>(3 * p.nBase);
        waveDetailKernel<<<static_cast<unsigned>(std::min<unsigned long long>(maxTiles, 6ull * sms)), 128, smemDetail, st>>>(p, iter);
        n += 4;
    }
    waveDecideKernel<<<std::min((nActHint + 31u) / 32u, 16u * sms), 256, 0, st>>>(p, iter);
    n += 1;
    if (launches) *launches += n;
    return cudaGetLastError();
}

}  // namespace mnenc
