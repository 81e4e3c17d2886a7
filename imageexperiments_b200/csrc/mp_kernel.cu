// Matching-pursuit kernel for sm_100a: block gather + RGB->YUV, shared-dictionary correlation, per-block
// KLT-detail correlation, arg-max / quantize / residual update — all of
//   compressed::encodeImageFast's block loop   CompressionLib/src/CompressedImage.cpp:592-631
//   img::YUVFromRGB                            ImageHelper/src/misc.cpp:12-21
//   matching::CalcMPDynamicFast / SelectFast   CompressionLib/src/MatchingPursuit.cpp:27-37,76-107
//   compressed::dynamicBasisFast               CompressionLib/src/CompressedImage.cpp:252-289 (as index arithmetic)
// in one persistent launch.
//
// Shape of the kernel
//   * one CTA per SM, 512 threads, the 510x64 base dictionary resident in shared memory (transposed, 128 KB,
//     brought in by one TMA bulk copy), 32 "slots" = block-channels in flight per CTA.
//   * every trip of the CTA loop advances each slot by one greedy iteration; a slot that finishes
//     (quantised coefficient 0, or K atoms) writes its symbols and immediately claims the next block-channel
//     from a global counter, so all 32 columns of the correlation stay busy whatever the per-block atom counts.
//   * step A (base atoms): a [512 x 64] x [64 x 32] product on the FP32 pipe with a 4x8 register tile per
//     thread. The oracle's arithmetic is reproduced exactly: per atom acc = fl(acc + fl(d*r)), j ascending,
//     never fused (Eigen GEMV without FMA, GeneralMatrixVector.h:108-215).
//   * step B (detail atoms): one warp per (slot, appended group); atoms of the group are read coalesced from
//     the transposed table in L2, residual broadcast from shared memory, warp-shuffle max reduction.
//   * step C: per slot a half-warp decodes the winner (64-bit key: |p| bits, then lowest index), quantises,
//     updates the residual in shared memory with unfused mul/sub, appends the detail group.
#include "mnenc_internal.h"

#include <cuda_runtime.h>

namespace mnenc {

namespace {

constexpr unsigned kFull = 0xFFFFFFFFu;

struct __align__(16) MpSmem {
    float D[kN * kRowsPad];          // base atoms, D[j*512 + atom]
    float R[kN * kSlots];            // residuals, R[j*32 + slot]
    unsigned long long best[kSlots];  // arg-max key of the current iteration
    unsigned long long mbar;
    uint32_t pairs[kSlots * kMaxK];
    uint16_t groupAtom[kSlots * kMaxK];   // appended detail groups in pick order
    uint16_t groupFirst[kSlots * kMaxK];  // dictionary index of the first row of that group
    float quant[3 * kMaxK];
    int bc[kSlots];       // block-channel owned by the slot, -1 = idle
    int reps[kSlots];
    int prev[kSlots];
    int ngroups[kSlots];
    int height[kSlots];   // current dictionary height: nBase + rows of the appended groups
    uint8_t detRows[kRowsPad];
};

// Key such that max(key) == the oracle's arg-max: larger |p| first, then the LOWER index (Visitor.h:227,
// strict '>' while scanning upwards). Bit 0 carries the sign of p so the signed projection can be recovered.
__device__ __forceinline__ unsigned long long makeKey(float p, int idx) {
    unsigned int b = __float_as_uint(p);
    return (static_cast<unsigned long long>(b & 0x7FFFFFFFu) << 32) | (static_cast<unsigned long long>(0x7FFFFFFFu - idx) << 1) |
           (b >> 31);
}
__device__ __forceinline__ unsigned long long umax64(unsigned long long a, unsigned long long b) { return a > b ? a : b; }
__device__ __forceinline__ unsigned long long warpMax64(unsigned long long k) {
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) k = umax64(k, __shfl_xor_sync(kFull, k, off));
    return k;
}
__device__ __forceinline__ uint32_t zigzag(int x) { return (static_cast<uint32_t>(x) << 1) ^ static_cast<uint32_t>(x >> 31); }

__device__ __forceinline__ uint32_t smemAddr(const void* p) { return static_cast<uint32_t>(__cvta_generic_to_shared(p)); }

// RGB -> one channel of YUV in double, then float (misc.cpp:12-21; left-associative sums, no contraction).
__device__ __forceinline__ float yuvChannel(int ch, double R, double G, double B) {
    double y = __dadd_rn(__dadd_rn(__dmul_rn(0.299, R), __dmul_rn(0.587, G)), __dmul_rn(0.114, B));
    if (ch == 0) return __double2float_rn(y);
    if (ch == 1) return __double2float_rn(__dmul_rn(0.436 / (1.0 - 0.114), __dsub_rn(B, y)));
    return __double2float_rn(__dmul_rn(0.615 / (1.0 - 0.299), __dsub_rn(R, y)));
}

__global__ void __launch_bounds__(kMpThreads, 1) matchingPursuitKernel(const MpParams p) {
    extern __shared__ __align__(128) unsigned char smemRaw[];
    MpSmem& s = *reinterpret_cast<MpSmem*>(smemRaw);
    const int tid = threadIdx.x;
    const int lane = tid & 31;
    const int warp = tid >> 5;

    // ---- base dictionary -> shared memory with one TMA bulk copy per 32 KB ----
    if (tid == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smemAddr(&s.mbar)));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    for (int i = tid; i < 3 * kMaxK; i += kMpThreads) s.quant[i] = p.quant[i];
    for (int i = tid; i < kRowsPad; i += kMpThreads) s.detRows[i] = i < p.nBase ? p.detRows[i] : 0;
    __syncthreads();
    if (tid == 0) {
        constexpr uint32_t total = kN * kRowsPad * sizeof(float);
        constexpr uint32_t chunk = 32768;
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smemAddr(&s.mbar)), "r"(total) : "memory");
        for (uint32_t o = 0; o < total; o += chunk) {
            asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                             smemAddr(reinterpret_cast<unsigned char*>(s.D) + o)),
                         "l"(reinterpret_cast<const unsigned char*>(p.baseT) + o), "r"(chunk), "r"(smemAddr(&s.mbar))
                         : "memory");
        }
    }

    // Slot bookkeeping: half-warp `slot` owns slot `slot` in step C.
    const int slot = tid >> 4;
    const int l16 = tid & 15;
    const unsigned halfMask = 0xFFFFu << (lane & 16);
    unsigned long long statRows = 0, statPicks = 0;

    // Claim the next block-channel for this slot and gather its pixels (CompressedImage.cpp:595-611).
    auto refill = [&]() {
        unsigned int next = 0;
        if (l16 == 0) next = atomicAdd(p.workCounter, 1u);
        next = __shfl_sync(halfMask, next, lane & 16);
        if (next < p.nbc) {
            const int blk = next / 3, ch = next - 3 * blk;
            const int bx = blk / p.nby, by = blk - bx * p.nby;
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                const int px = l16 + 16 * e;  // block vector index dx + 8*dy
                const int x = bx * 8 + (px & 7), y = by * 8 + (px >> 3);
                float v = 0.0f;
                if (x < p.width && y < p.height) {
                    const uint8_t* q = p.rgb + static_cast<size_t>(y) * p.stride + 3 * static_cast<size_t>(x);
                    v = yuvChannel(ch, static_cast<double>(q[0]), static_cast<double>(q[1]), static_cast<double>(q[2]));
                }
                s.R[px * kSlots + slot] = v;
            }
            if (l16 == 0) {
                s.bc[slot] = static_cast<int>(next);
                s.reps[slot] = 0;
                s.prev[slot] = 0;
                s.ngroups[slot] = 0;
                s.height[slot] = p.nBase;
            }
        } else {
#pragma unroll
            for (int e = 0; e < 4; ++e) s.R[(l16 + 16 * e) * kSlots + slot] = 0.0f;
            if (l16 == 0) {
                s.bc[slot] = -1;
                s.ngroups[slot] = 0;
            }
        }
        if (l16 == 0) s.best[slot] = 0ULL;
    };
    refill();

    // wait for the dictionary
    {
        uint32_t done = 0;
        while (!done) {
            asm volatile(
                "{\n\t.reg .pred P;\n\tmbarrier.try_wait.parity.shared::cta.b64 P, [%1], 0;\n\tselp.u32 %0, 1, 0, P;\n\t}"
                : "=r"(done)
                : "r"(smemAddr(&s.mbar))
                : "memory");
        }
    }
    __syncthreads();

    const int bcg = warp >> 2;               // step A: 8 slots [8*bcg, 8*bcg+8)
    const int rg = ((warp & 3) << 5) | lane;  // step A: atoms [4*rg, 4*rg+4)

    while (true) {
        // ================= step A: base atoms, 4 atoms x 8 slots per thread =================
        {
            float acc[4][8];
#pragma unroll
            for (int a = 0; a < 4; ++a)
#pragma unroll
                for (int b = 0; b < 8; ++b) acc[a][b] = 0.0f;
            const float4* dptr = reinterpret_cast<const float4*>(s.D) + rg;
            const float4* rptr = reinterpret_cast<const float4*>(s.R) + 2 * bcg;
#pragma unroll 4
            for (int j = 0; j < kN; ++j) {
                const float4 d4 = dptr[j * (kRowsPad / 4)];
                const float4 r0 = rptr[j * (kSlots / 4)];
                const float4 r1 = rptr[j * (kSlots / 4) + 1];
                const float d[4] = {d4.x, d4.y, d4.z, d4.w};
                const float r[8] = {r0.x, r0.y, r0.z, r0.w, r1.x, r1.y, r1.z, r1.w};
#pragma unroll
                for (int a = 0; a < 4; ++a)
#pragma unroll
                    for (int b = 0; b < 8; ++b) acc[a][b] = __fadd_rn(acc[a][b], __fmul_rn(d[a], r[b]));
            }
#pragma unroll
            for (int b = 0; b < 8; ++b) {
                unsigned long long k = 0ULL;
#pragma unroll
                for (int a = 0; a < 4; ++a) {
                    const int idx = 4 * rg + a;
                    if (idx < p.nBase) k = umax64(k, makeKey(acc[a][b], idx));
                }
                k = warpMax64(k);
                if (lane == 0) atomicMax(&s.best[8 * bcg + b], k);
            }
        }
        // ================= step B: appended detail groups, one warp per (slot, group) =================
        {
            const int ng = s.ngroups[lane];
            int incl = ng;
#pragma unroll
            for (int off = 1; off < 32; off <<= 1) {
                int t = __shfl_up_sync(kFull, incl, off);
                if (lane >= off) incl += t;
            }
            const int total = __shfl_sync(kFull, incl, 31);
            for (int item = warp; item < total; item += kMpThreads / 32) {
                const unsigned m = __ballot_sync(kFull, incl > item);
                const int sl = __ffs(m) - 1;
                const int gi = item - (__shfl_sync(kFull, incl, sl) - __shfl_sync(kFull, ng, sl));
                const int atom = s.groupAtom[sl * kMaxK + gi];
                const int first = s.groupFirst[sl * kMaxK + gi];
                const int nrows = s.detRows[atom];
                const int ch = s.bc[sl] % 3;
                const float* g = p.detT + (static_cast<size_t>(ch) * p.nBase + atom) * (kN * 64) + lane;
                float a0 = 0.0f, a1 = 0.0f;
#pragma unroll 8
                for (int j = 0; j < kN; ++j) {
                    const float d0 = __ldg(g + j * 64);
                    const float d1 = __ldg(g + j * 64 + 32);
                    const float r = s.R[j * kSlots + sl];
                    a0 = __fadd_rn(a0, __fmul_rn(d0, r));
                    a1 = __fadd_rn(a1, __fmul_rn(d1, r));
                }
                unsigned long long k = makeKey(a0, first + lane);  // rows 0..31 always exist (62/63 rows)
                if (lane + 32 < nrows) k = umax64(k, makeKey(a1, first + lane + 32));
                k = warpMax64(k);
                if (lane == 0) atomicMax(&s.best[sl], k);
            }
        }
        __syncthreads();
        // ================= step C: select, quantise, update (MatchingPursuit.cpp:86-104) =================
        int active = 0;
        {
            const int bc = s.bc[slot];
            if (bc >= 0) {
                const unsigned long long key = s.best[slot];
                const unsigned int lo = static_cast<unsigned int>(key);
                const int idx = static_cast<int>(0x7FFFFFFFu - (lo >> 1));
                const float proj = __uint_as_float(static_cast<unsigned int>(key >> 32) | (lo << 31));
                const int reps = s.reps[slot];
                const int prev = s.prev[slot];
                const int ng = s.ngroups[slot];
                const int height = s.height[slot];
                const int ch = bc % 3;
                const float qf = s.quant[ch * kMaxK + reps];
                const int qi = static_cast<int>(roundf(__fdiv_rn(proj, qf)));
                const uint32_t deltaId = (reps ? zigzag(idx - prev) : static_cast<uint32_t>(idx)) & 0xFFFFu;
                statRows += (l16 == 0) ? height : 0;
                int count = -1;  // >= 0: finished with that many atoms
                __syncwarp(halfMask);
                if (qi == 0) {
                    count = reps;
                } else {
                    const float c = __fmul_rn(qf, static_cast<float>(qi));
                    const float* row;
                    int rstride;
                    if (idx < p.nBase) {
                        row = s.D + idx;
                        rstride = kRowsPad;
                    } else {
                        int gi = ng - 1;
                        while (gi > 0 && s.groupFirst[slot * kMaxK + gi] > idx) --gi;
                        const int atom = s.groupAtom[slot * kMaxK + gi];
                        row = p.detT + (static_cast<size_t>(ch) * p.nBase + atom) * (kN * 64) + (idx - s.groupFirst[slot * kMaxK + gi]);
                        rstride = 64;
                    }
#pragma unroll
                    for (int e = 0; e < 4; ++e) {
                        const int j = l16 + 16 * e;
                        const float rv = s.R[j * kSlots + slot];
                        s.R[j * kSlots + slot] = __fsub_rn(rv, __fmul_rn(c, row[j * rstride]));
                    }
                    if (l16 == 0) {
                        s.pairs[slot * kMaxK + reps] = deltaId | ((zigzag(qi) & 0xFFFFu) << 16);
                        s.prev[slot] = idx;
                        s.reps[slot] = reps + 1;
                        s.best[slot] = 0ULL;
                        if (idx < p.nBase) {
                            s.groupAtom[slot * kMaxK + ng] = static_cast<uint16_t>(idx);
                            s.groupFirst[slot * kMaxK + ng] = static_cast<uint16_t>(height);
                            s.ngroups[slot] = ng + 1;
                            s.height[slot] = height + s.detRows[idx];
                        }
                    }
                    if (reps + 1 == p.K) count = p.K;
                }
                if (count >= 0) {
                    __syncwarp(halfMask);
                    if (l16 == 0) p.counts[bc] = static_cast<uint16_t>(count);
                    for (int i = l16; i < p.K; i += 16) p.pairs[static_cast<size_t>(bc) * p.K + i] = i < count ? s.pairs[slot * kMaxK + i] : 0u;
                    statPicks += (l16 == 0) ? count : 0;
                    __syncwarp(halfMask);
                    refill();
                }
                active = 1;
            }
        }
        if (!__syncthreads_or(active)) break;
    }
    // algorithmic-work accounting
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) {
        statRows += __shfl_xor_sync(kFull, statRows, off);
        statPicks += __shfl_xor_sync(kFull, statPicks, off);
    }
    if (lane == 0) {
        atomicAdd(&p.stats[0], statRows);
        atomicAdd(&p.stats[1], statPicks);
    }
}

}  // namespace

size_t matchingPursuitSmemBytes() { return sizeof(MpSmem) + 128; }

cudaError_t launchMatchingPursuit(const MpParams& p, int numSMs, cudaStream_t stream) {
    static bool configured = false;
    const size_t smem = matchingPursuitSmemBytes();
    if (!configured) {
        cudaError_t e = cudaFuncSetAttribute(matchingPursuitKernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
        if (e != cudaSuccess) return e;
        configured = true;
    }
    if (p.nbc == 0) return cudaSuccess;
    unsigned int grid = (p.nbc + kSlots - 1) / kSlots;
    if (grid > static_cast<unsigned int>(numSMs)) grid = numSMs;
    matchingPursuitKernel<<<grid, kMpThreads, smem, stream>>>(p);
    return cudaGetLastError();
}

}  // namespace mnenc
