// Matching-pursuit kernel, tensor-core version (sm_100a): the correlation of the 32 in-flight residuals with the
// 510 shared base atoms runs as a dense [512 x 64] x [64 x 32] contraction on the 5th-generation tensor cores
// (tcgen05.mma kind::tf32, accumulators in TMEM), used as a PRE-FILTER; every decision is then taken with the
// oracle's exact FP32 arithmetic:
//
//   1. TF32 scores p~ for all 512 atoms x 32 slots (32 tcgen05.mma of 128x32x8 per trip, issued by one thread,
//      operands read straight from the FP32 dictionary / residual images in shared memory).
//   2. Rigorous error bound per slot: operands are cut to 11 significant bits (|delta| <= 2^-10 each), so
//      |p~_i - p_i| <= (2*2^-10 + 2^-20 + accumulation terms) * sum_j |d_ij r_j| <= 2.0e-3 * ||d_i|| ||r||.
//      With ||d_i|| <= 1.0001 (unit-norm atoms) eps = 2.2e-3 * ||r||_2 bounds the difference between the TF32
//      score and the oracle's sequential FP32 dot product (which itself is within 4e-6 ||r|| of the exact value).
//   3. Every atom with |p~_i| >= max_i |p~_i| - 2 eps is re-scored exactly (acc = fl(acc + fl(d*r)), j ascending,
//      unfused; GeneralMatrixVector.h:108-215). The oracle's winner and everything that ties with it are always in
//      that set, so max over the exact keys (|p| bits, lowest index) reproduces SelectFast bit for bit
//      (MatchingPursuit.cpp:27-37, Visitor.h:227). Typically 2-3 atoms per slot survive the filter.
//   4. Detail (KLT) atoms, quantisation and the residual update are as in the FP32 kernel: exact.
//
// Shared-memory operand images use the canonical K-major no-swizzle UMMA layout (8-row x 16-byte core matrices):
//   byte offset(row, k) = (row / 8) * 2048 + (k / 4) * 128 + (row % 8) * 16 + (k % 4) * 4
// for both the dictionary (rows = atoms, 128 KB, written once by TMA bulk copies from a pre-arranged global
// image) and the residuals (rows = slots, 8 KB, written by the threads of step C). LBO = 128 B, SBO = 2048 B.
#include "mnenc_internal.h"

#include <cuda_runtime.h>

namespace mnenc {

namespace {

constexpr unsigned kFull = 0xFFFFFFFFu;
constexpr uint32_t kGroupBytes = 2048;  // 8 rows x 64 k x 4 B
constexpr uint32_t kChunkBytes = 128;   // one core matrix: 8 rows x 16 B

__host__ __device__ constexpr int umOff(int row, int k) {  // float index in a canonical operand image
    return (row >> 3) * 512 + (k >> 2) * 32 + (row & 7) * 4 + (k & 3);
}

struct __align__(1024) MpSmemTc {
    float D[kN * kRowsPad];            // 128 KB dictionary image
    float R[kN * kSlots];              // 8 KB residual image
    unsigned long long best[kSlots];   // exact arg-max key
    unsigned long long mbarDict, mbarMma;
    alignas(16) uint32_t approxMax[kSlots];  // max_i |p~_i| per slot, float bits (read as float4)
    alignas(16) float twoEps[kSlots];        // 2 * eps per slot (read as float4)
    uint32_t tmemBase;
    uint32_t pairs[kSlots * kMaxK];
    uint16_t groupAtom[kSlots * kMaxK];
    uint16_t groupFirst[kSlots * kMaxK];
    float quant[3 * kMaxK];
    int bc[kSlots];
    int reps[kSlots];
    int prev[kSlots];
    int ngroups[kSlots];
    int height[kSlots];
    uint8_t detRows[kRowsPad];
};

__device__ __forceinline__ unsigned long long makeKey(float p, int idx) {
    const unsigned int b = __float_as_uint(p);
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

__device__ __forceinline__ float yuvChannel(int ch, double R, double G, double B) {  // misc.cpp:12-21
    const double y = __dadd_rn(__dadd_rn(__dmul_rn(0.299, R), __dmul_rn(0.587, G)), __dmul_rn(0.114, B));
    if (ch == 0) return __double2float_rn(y);
    if (ch == 1) return __double2float_rn(__dmul_rn(0.436 / (1.0 - 0.114), __dsub_rn(B, y)));
    return __double2float_rn(__dmul_rn(0.615 / (1.0 - 0.299), __dsub_rn(R, y)));
}

// ---- tcgen05 / mbarrier wrappers ----
__device__ __forceinline__ void mbarInit(unsigned long long* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smemAddr(bar)), "r"(count));
}
__device__ __forceinline__ void mbarWait(unsigned long long* bar, uint32_t parity) {
    uint32_t done = 0;
    while (!done) {
        asm volatile("{\n\t.reg .pred P;\n\tmbarrier.try_wait.parity.shared::cta.b64 P, [%1], %2;\n\tselp.u32 %0, 1, 0, P;\n\t}"
                     : "=r"(done)
                     : "r"(smemAddr(bar)), "r"(parity)
                     : "memory");
    }
}
// K-major, no swizzle: start address, LBO = 128 B (between the two 16-byte k-chunks of one MMA), SBO = 2048 B
// (between 8-row groups), descriptor version 1 (Blackwell). Fields are in 16-byte units.
__device__ __forceinline__ unsigned long long umDesc(uint32_t saddr) {
    return static_cast<unsigned long long>((saddr & 0x3FFFFu) >> 4) | (static_cast<unsigned long long>(kChunkBytes >> 4) << 16) |
           (static_cast<unsigned long long>(kGroupBytes >> 4) << 32) | (1ull << 46);
}
// D[tmem] (+)= A[smem] * B[smem], kind::tf32, M = 128, N = 32, K = 8
__device__ __forceinline__ void ummaTf32(uint32_t tmemD, unsigned long long descA, unsigned long long descB, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmemD),
        "l"(descA), "l"(descB), "r"(idesc), "r"(accumulate)
        : "memory");
}
__device__ __forceinline__ void ummaCommit(unsigned long long* bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smemAddr(bar)) : "memory");
}
__device__ __forceinline__ void tmemLoad32(uint32_t taddr, uint32_t (&v)[32]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), "=r"(v[9]),
          "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]), "=r"(v[17]), "=r"(v[18]),
          "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]),
          "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
        : "r"(taddr)
        : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}
// M = 128, N = 32, A/B = TF32 K-major, D = F32 (cute::UMMA::InstrDescriptor bit layout)
constexpr uint32_t kIdesc = (1u << 4) | (2u << 7) | (2u << 10) | ((32u >> 3) << 17) | ((128u >> 4) << 24);

}  // namespace

namespace {

__global__ void __launch_bounds__(kMpThreads, 1) matchingPursuitTcKernel(const MpParams p) {
    extern __shared__ __align__(1024) unsigned char smemRaw[];
    MpSmemTc& s = *reinterpret_cast<MpSmemTc*>(smemRaw);
    const int tid = threadIdx.x;
    const int lane = tid & 31;
    const int warp = tid >> 5;

    // ---- one-time setup: barriers, TMEM (4 accumulator tiles x 32 columns), dictionary image via TMA bulk copies ----
    if (tid == 0) {
        mbarInit(&s.mbarDict, 1);
        mbarInit(&s.mbarMma, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smemAddr(&s.tmemBase)), "r"(128u) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    for (int i = tid; i < 3 * kMaxK; i += kMpThreads) s.quant[i] = p.quant[i];
    for (int i = tid; i < kRowsPad; i += kMpThreads) s.detRows[i] = i < p.nBase ? p.detRows[i] : 0;
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmemBase = s.tmemBase;
    if (tid == 0) {
        constexpr uint32_t total = kN * kRowsPad * sizeof(float);
        constexpr uint32_t chunk = 32768;
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smemAddr(&s.mbarDict)), "r"(total) : "memory");
        for (uint32_t o = 0; o < total; o += chunk)
            asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                             smemAddr(reinterpret_cast<unsigned char*>(s.D) + o)),
                         "l"(reinterpret_cast<const unsigned char*>(p.baseUm) + o), "r"(chunk), "r"(smemAddr(&s.mbarDict))
                         : "memory");
    }

    const int slot = tid >> 4;  // step C: half-warp `slot` owns slot `slot`; its thread l16 owns k = 4*l16 .. 4*l16+3
    const int l16 = tid & 15;
    const unsigned halfMask = 0xFFFFu << (lane & 16);
    float4* const myR = reinterpret_cast<float4*>(s.R + (slot >> 3) * 512 + l16 * 32 + (slot & 7) * 4);
    unsigned long long statRows = 0, statPicks = 0;

    auto halfSum = [&](float v) {
#pragma unroll
        for (int off = 8; off > 0; off >>= 1) v += __shfl_xor_sync(halfMask, v, off);
        return v;
    };

    // Claim the next block-channel for this slot and gather its pixels (CompressedImage.cpp:595-611).
    auto refill = [&]() {
        unsigned int next = 0;
        if (l16 == 0) next = atomicAdd(p.workCounter, 1u);
        next = __shfl_sync(halfMask, next, lane & 16);
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (next < p.nbc) {
            const int blk = next / 3, ch = next - 3 * blk;
            const int bx = blk / p.nby, by = blk - bx * p.nby;
            float e[4];
#pragma unroll
            for (int t = 0; t < 4; ++t) {
                const int px = 4 * l16 + t;  // block vector index dx + 8*dy
                const int x = bx * 8 + (px & 7), y = by * 8 + (px >> 3);
                e[t] = 0.0f;
                if (x < p.width && y < p.height) {
                    const uint8_t* q = p.rgb + static_cast<size_t>(y) * p.stride + 3 * static_cast<size_t>(x);
                    e[t] = yuvChannel(ch, static_cast<double>(q[0]), static_cast<double>(q[1]), static_cast<double>(q[2]));
                }
            }
            v = make_float4(e[0], e[1], e[2], e[3]);
        }
        *myR = v;
        const float ss = halfSum(v.x * v.x + v.y * v.y + v.z * v.z + v.w * v.w);
        if (l16 == 0) {
            s.bc[slot] = next < p.nbc ? static_cast<int>(next) : -1;
            s.reps[slot] = 0;
            s.prev[slot] = 0;
            s.ngroups[slot] = 0;
            s.height[slot] = p.nBase;
            s.best[slot] = 0ULL;
            s.approxMax[slot] = 0u;
            s.twoEps[slot] = 2.0f * p.epsRel * sqrtf(ss);
        }
    };
    refill();
    mbarWait(&s.mbarDict, 0);
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // residual image written by threads -> visible to the tensor core
    __syncthreads();

    const int tile = warp >> 2, quarter = warp & 3;               // epilogue: this warp owns atoms tile*128 + quarter*32 + lane
    const int myAtom = tile * 128 + quarter * 32 + lane;
    const uint32_t myTaddr = tmemBase + (static_cast<uint32_t>(quarter * 32) << 16) + static_cast<uint32_t>(tile * 32);
    const uint32_t dAddr = smemAddr(s.D), rAddr = smemAddr(s.R);
    uint32_t parity = 0;

    while (true) {
        // ================= step A1: TF32 scores of all base atoms on the tensor cores (asynchronous) =================
        if (tid == 0) {
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
#pragma unroll
            for (int t = 0; t < 4; ++t)
#pragma unroll
                for (int k = 0; k < 8; ++k)
                    ummaTf32(tmemBase + t * 32, umDesc(dAddr + t * 16 * kGroupBytes + k * 2 * kChunkBytes), umDesc(rAddr + k * 2 * kChunkBytes),
                             kIdesc, k > 0 ? 1u : 0u);
            ummaCommit(&s.mbarMma);
        }
        // ================= step B: appended detail groups, exact FP32, overlaps the tensor-core work =================
        {
            const int ng = s.ngroups[lane];
            int incl = ng;
#pragma unroll
            for (int off = 1; off < 32; off <<= 1) {
                const int t = __shfl_up_sync(kFull, incl, off);
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
                const float* rs = s.R + (sl >> 3) * 512 + (sl & 7) * 4;
                float a0 = 0.0f, a1 = 0.0f;
#pragma unroll 16
                for (int j = 0; j < kN; ++j) {
                    const float d0 = __ldg(g + j * 64);
                    const float d1 = __ldg(g + j * 64 + 32);
                    const float r = rs[(j >> 2) * 32 + (j & 3)];
                    a0 = __fadd_rn(a0, __fmul_rn(d0, r));
                    a1 = __fadd_rn(a1, __fmul_rn(d1, r));
                }
                unsigned long long k = makeKey(a0, first + lane);
                if (lane + 32 < nrows) k = umax64(k, makeKey(a1, first + lane + 32));
                k = warpMax64(k);
                if (lane == 0) atomicMax(&s.best[sl], k);
            }
        }
        // ================= step A2: filter + exact re-score =================
        mbarWait(&s.mbarMma, parity);
        parity ^= 1u;
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        float v[32];
        {
            uint32_t raw[32];
            tmemLoad32(myTaddr, raw);
#pragma unroll
            for (int i = 0; i < 32; ++i) v[i] = fabsf(__uint_as_float(raw[i]));
        }
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
        {   // per-slot maximum over the warp's 32 atoms: transposed butterfly, lane L ends with slot L
            float w16[16], w8[8], w4[4], w2[2];
            const bool b4 = lane & 16, b3 = lane & 8, b2 = lane & 4, b1 = lane & 2, b0 = lane & 1;
#pragma unroll
            for (int i = 0; i < 16; ++i) w16[i] = fmaxf(b4 ? v[i + 16] : v[i], __shfl_xor_sync(kFull, b4 ? v[i] : v[i + 16], 16));
#pragma unroll
            for (int i = 0; i < 8; ++i) w8[i] = fmaxf(b3 ? w16[i + 8] : w16[i], __shfl_xor_sync(kFull, b3 ? w16[i] : w16[i + 8], 8));
#pragma unroll
            for (int i = 0; i < 4; ++i) w4[i] = fmaxf(b2 ? w8[i + 4] : w8[i], __shfl_xor_sync(kFull, b2 ? w8[i] : w8[i + 4], 4));
#pragma unroll
            for (int i = 0; i < 2; ++i) w2[i] = fmaxf(b1 ? w4[i + 2] : w4[i], __shfl_xor_sync(kFull, b1 ? w4[i] : w4[i + 2], 2));
            const float m = fmaxf(b0 ? w2[1] : w2[0], __shfl_xor_sync(kFull, b0 ? w2[0] : w2[1], 1));
            atomicMax(&s.approxMax[lane], __float_as_uint(m));
        }
        __syncthreads();
        if (myAtom < p.nBase) {
            unsigned cand = 0;
            const float4* mx = reinterpret_cast<const float4*>(s.approxMax);
            const float4* te = reinterpret_cast<const float4*>(s.twoEps);
#pragma unroll
            for (int i4 = 0; i4 < 8; ++i4) {
                const float4 m4 = mx[i4];  // bit patterns of non-negative floats
                const float4 e4 = te[i4];
                cand |= (v[4 * i4 + 0] + e4.x >= m4.x ? 1u : 0u) << (4 * i4 + 0);
                cand |= (v[4 * i4 + 1] + e4.y >= m4.y ? 1u : 0u) << (4 * i4 + 1);
                cand |= (v[4 * i4 + 2] + e4.z >= m4.z ? 1u : 0u) << (4 * i4 + 2);
                cand |= (v[4 * i4 + 3] + e4.w >= m4.w ? 1u : 0u) << (4 * i4 + 3);
            }
            const float4* dp = reinterpret_cast<const float4*>(s.D + (myAtom >> 3) * 512 + (myAtom & 7) * 4);
            while (cand) {
                const int sl = __ffs(cand) - 1;
                cand &= cand - 1;
                const float4* rp = reinterpret_cast<const float4*>(s.R + (sl >> 3) * 512 + (sl & 7) * 4);
                float acc = 0.0f;
#pragma unroll
                for (int c = 0; c < 16; ++c) {  // exact: j = 4c .. 4c+3 ascending, unfused
                    const float4 d4 = dp[c * 8];
                    const float4 r4 = rp[c * 8];
                    acc = __fadd_rn(acc, __fmul_rn(d4.x, r4.x));
                    acc = __fadd_rn(acc, __fmul_rn(d4.y, r4.y));
                    acc = __fadd_rn(acc, __fmul_rn(d4.z, r4.z));
                    acc = __fadd_rn(acc, __fmul_rn(d4.w, r4.w));
                }
                atomicMax(&s.best[sl], makeKey(acc, myAtom));
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
                int count = -1;
                __syncwarp(halfMask);
                if (qi == 0) {
                    count = reps;
                } else {
                    const float c = __fmul_rn(qf, static_cast<float>(qi));
                    float4 row;
                    if (idx < p.nBase) {
                        row = *reinterpret_cast<const float4*>(s.D + (idx >> 3) * 512 + l16 * 32 + (idx & 7) * 4);
                    } else {
                        int gi = ng - 1;
                        while (gi > 0 && s.groupFirst[slot * kMaxK + gi] > idx) --gi;
                        const int atom = s.groupAtom[slot * kMaxK + gi];
                        const float* g = p.detT + (static_cast<size_t>(ch) * p.nBase + atom) * (kN * 64) + (idx - s.groupFirst[slot * kMaxK + gi]) +
                                         4 * l16 * 64;
                        row = make_float4(__ldg(g), __ldg(g + 64), __ldg(g + 128), __ldg(g + 192));
                    }
                    float4 r = *myR;
                    r.x = __fsub_rn(r.x, __fmul_rn(c, row.x));
                    r.y = __fsub_rn(r.y, __fmul_rn(c, row.y));
                    r.z = __fsub_rn(r.z, __fmul_rn(c, row.z));
                    r.w = __fsub_rn(r.w, __fmul_rn(c, row.w));
                    *myR = r;
                    const float ss = halfSum(r.x * r.x + r.y * r.y + r.z * r.z + r.w * r.w);
                    if (l16 == 0) {
                        s.pairs[slot * kMaxK + reps] = deltaId | ((zigzag(qi) & 0xFFFFu) << 16);
                        s.prev[slot] = idx;
                        s.reps[slot] = reps + 1;
                        s.best[slot] = 0ULL;
                        s.approxMax[slot] = 0u;
                        s.twoEps[slot] = 2.0f * p.epsRel * sqrtf(ss);
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
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        if (!__syncthreads_or(active)) break;
    }
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) {
        statRows += __shfl_xor_sync(kFull, statRows, off);
        statPicks += __shfl_xor_sync(kFull, statPicks, off);
    }
    if (lane == 0) {
        atomicAdd(&p.stats[0], statRows);
        atomicAdd(&p.stats[1], statPicks);
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmemBase), "r"(128u) : "memory");
}

}  // namespace

void buildUmmaImage(const float* baseRowMajor, int nBase, float* image) {
    for (int i = 0; i < kN * kRowsPad; ++i) image[i] = 0.0f;
    for (int a = 0; a < nBase; ++a)
        for (int k = 0; k < kN; ++k) image[umOff(a, k)] = baseRowMajor[static_cast<size_t>(a) * kN + k];
}

cudaError_t launchMatchingPursuitTc(const MpParams& p, int numSMs, cudaStream_t stream) {
    static bool configured = false;
    const size_t smem = sizeof(MpSmemTc) + 1024;
    if (!configured) {
        cudaError_t e = cudaFuncSetAttribute(matchingPursuitTcKernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
        if (e != cudaSuccess) return e;
        configured = true;
    }
    if (p.nbc == 0) return cudaSuccess;
    unsigned int grid = (p.nbc + kSlots - 1) / kSlots;
    if (grid > static_cast<unsigned int>(numSMs)) grid = numSMs;
    matchingPursuitTcKernel<<<grid, kMpThreads, smem, stream>>>(p);
    return cudaGetLastError();
}

}  // namespace mnenc
