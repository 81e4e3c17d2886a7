// Matching-pursuit kernel, version 3 (sm_100a): BOTH correlations of a greedy iteration run on the 5th-generation
// tensor cores as a pre-filter, every decision is then taken with the oracle's exact FP32 arithmetic.
//
//   base atoms   [512 x 64] x [64 x 32 slots]   tcgen05.mma kind::f16, dictionary image resident in shared memory
//   KLT groups   [2 x 64 rows x 64] x [64 x 32] one MMA tile per PAIR of (slot, appended group) items; the 8 KB fp16
//                group images stream from L2 through a shared-memory ring filled by TMA bulk copies
//                (dynamicBasisFast, CompressedImage.cpp:252-289, becomes index arithmetic + a copy engine)
//
// Operands are fp16 images; the residual is split r = hi + lo (two fp16 images, two accumulating MMA passes), so the
// only first-order error is the fp16 rounding of the dictionary: |p~ - p| <= epsRel * ||r||_2 with
// epsRel = max_atom ||d - fp16(d)||_2 + small terms (computed on the host from the imported dictionary,
// mnenc_api.cu). Every atom with |p~| >= max|p~| - 2 eps goes to a candidate list and is re-scored exactly
// (acc = fl(acc + fl(d*r)), j ascending, unfused: Eigen GEMV without FMA, GeneralMatrixVector.h:108-215); the
// maximum over exact 64-bit keys (|p| bits, lowest index) reproduces SelectFast bit for bit
// (MatchingPursuit.cpp:27-37, Visitor.h:227). Quantisation and the residual update are exact as before.
//
// Warp roles (576 threads): warps 0-15 consumers (epilogues, exact re-score, step C), warp 16 TMA producer,
// warp 17 MMA issuer. TMEM: columns 0-127 base accumulators (4 tiles x 32 slots), 128-511 twelve pair accumulators.
//
// Operand images use the canonical K-major no-swizzle UMMA layout with 8-row x 16-byte core matrices:
//   byte offset(row, k) = (row / 8) * 1024 + (k / 8) * 128 + (row % 8) * 16 + (k % 8) * 2      LBO 128 B, SBO 1024 B
#include "mnenc_internal.h"

#include <cuda_fp16.h>
#include <cuda_runtime.h>

#include <algorithm>
#include <cmath>
#include <cstdio>

namespace mnenc {

namespace {

constexpr unsigned kFull = 0xFFFFFFFFu;
constexpr int kConsumerThreads = 512;
constexpr int kTmaWarp = 16, kMmaWarp = 17;
constexpr int kThreadsF16 = 576;
constexpr int kStages = 4;   // ring stages of 16 KB = two group images
constexpr int kAccs = 9;     // pair accumulators (32 TMEM columns each)
constexpr uint32_t kGroupImgBytes = 8192;
constexpr uint32_t kSbo = 1024, kLbo = 128, kKStepBytes = 256;  // one MMA consumes K = 16 = two 128-byte chunks
constexpr int kMaxItems = kSlots * kMaxK;
constexpr int kCandCap = 512;
constexpr int kRStride = 68;  // floats per slot in the FP32 residual array (272 B: float4 reads of 8 slots are conflict free)
constexpr uint32_t kDcAccCol = 128;   // 3 resident-group tiles (one per channel) x 32 slots
constexpr uint32_t kDetAccCol = 224;  // streamed pair accumulators

struct __align__(1024) SmemF16 {
    __half D16[kRowsPad * kN];                                      // 64 KB base dictionary image
    __half R16[2 * kSlots * kN];                                    // 8 KB residual image: rows 0-31 hi, 32-63 lo
    alignas(1024) unsigned char dc[3][2 * kGroupImgBytes];          // 48 KB resident KLT groups of atoms 0 and nBase-1 (Y, U, V)
    alignas(1024) unsigned char ring[kStages][2 * kGroupImgBytes];  // 64 KB
    alignas(16) float R32[kSlots * kRStride];                       // exact residuals
    unsigned long long best[kSlots];                                // exact arg-max key per slot
    unsigned long long mbarDict, mbarBase;
    unsigned long long full[kStages], empty[kStages];
    unsigned long long accFull[kAccs], accEmpty[kAccs];
    alignas(16) uint32_t approxMax[kSlots];  // max |p~| per slot, float bits
    alignas(16) float twoEps[kSlots];
    uint32_t tmemBase;
    uint32_t nItems[2], nDc[2];
    uint32_t nCand;
    uint16_t items[2][kMaxItems];    // streamed groups: slot | group position << 5
    uint16_t dcItems[2][kMaxItems];  // resident groups: slot | group position << 5 | (atom != 0) << 10
    uint2 cand[kCandCap];          // {slot | dictionary index << 8, row id in rows32}
    uint32_t pairs[kSlots * kMaxK];
    uint16_t groupAtom[kSlots * kMaxK];
    uint16_t groupFirst[kSlots * kMaxK];
    float quant[3 * kMaxK];
    int bc[kSlots];
    int reps[kSlots];
    int prev[kSlots];
    int ngroups[kSlots];
    int height[kSlots];
    uint8_t zeroRes[kSlots];
    uint8_t detRows[kRowsPad];
};

static_assert(sizeof(SmemF16) + 1024 <= 232448, "shared memory of the fp16 matching-pursuit kernel exceeds 227 KB");

__device__ __forceinline__ unsigned long long makeKey(float p, int idx) {
    const unsigned int b = __float_as_uint(p);
    return (static_cast<unsigned long long>(b & 0x7FFFFFFFu) << 32) | (static_cast<unsigned long long>(0x7FFFFFFFu - idx) << 1) |
           (b >> 31);
}
__device__ __forceinline__ uint32_t zigzag(int x) { return (static_cast<uint32_t>(x) << 1) ^ static_cast<uint32_t>(x >> 31); }
__device__ __forceinline__ uint32_t smemAddr(const void* p) { return static_cast<uint32_t>(__cvta_generic_to_shared(p)); }

__device__ __forceinline__ float yuvChannel(int ch, double R, double G, double B) {  // misc.cpp:12-21
    const double y = __dadd_rn(__dadd_rn(__dmul_rn(0.299, R), __dmul_rn(0.587, G)), __dmul_rn(0.114, B));
    if (ch == 0) return __double2float_rn(y);
    if (ch == 1) return __double2float_rn(__dmul_rn(0.436 / (1.0 - 0.114), __dsub_rn(B, y)));
    return __double2float_rn(__dmul_rn(0.615 / (1.0 - 0.299), __dsub_rn(R, y)));
}

// ---- mbarrier / tcgen05 / TMA wrappers ----
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
__device__ __forceinline__ void mbarArrive(unsigned long long* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smemAddr(bar)) : "memory");
}
__device__ __forceinline__ void mbarExpectTx(unsigned long long* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smemAddr(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulkCopy(void* dstSmem, const void* srcGlobal, uint32_t bytes, unsigned long long* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smemAddr(dstSmem)),
                 "l"(srcGlobal), "r"(bytes), "r"(smemAddr(bar))
                 : "memory");
}
__device__ __forceinline__ unsigned long long umDesc(uint32_t saddr) {  // K-major, no swizzle, descriptor version 1
    return static_cast<unsigned long long>((saddr & 0x3FFFFu) >> 4) | (static_cast<unsigned long long>(kLbo >> 4) << 16) |
           (static_cast<unsigned long long>(kSbo >> 4) << 32) | (1ull << 46);
}
// D[tmem] (+)= A[smem] * B[smem], kind::f16, M = 128, N = 32, K = 16
__device__ __forceinline__ void ummaF16(uint32_t tmemD, unsigned long long descA, unsigned long long descB, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmemD),
        "l"(descA), "l"(descB), "r"(idesc), "r"(accumulate)
        : "memory");
}
__device__ __forceinline__ void ummaCommit(unsigned long long* bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smemAddr(bar)) : "memory");
}
__device__ __forceinline__ void fenceBefore() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void fenceAfter() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
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
__device__ __forceinline__ uint32_t tmemLoad1(uint32_t taddr) {
    uint32_t v;
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x1.b32 {%0}, [%1];" : "=r"(v) : "r"(taddr) : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
    return v;
}
__device__ __forceinline__ void consumerBarrier() { asm volatile("bar.sync 1, %0;" ::"n"(kConsumerThreads) : "memory"); }
// M = 128, N = 32, A/B = F16 K-major, D = F32 (cute::UMMA::InstrDescriptor bit layout)
constexpr uint32_t kIdescF16 = (1u << 4) | (0u << 7) | (0u << 10) | ((32u >> 3) << 17) | ((128u >> 4) << 24);

// exact oracle dot product of one dictionary row (global, row-major FP32) with the residual of one slot
__device__ __forceinline__ float exactDot(const float* row, const float* r) {
    const float4* dp = reinterpret_cast<const float4*>(row);
    const float4* rp = reinterpret_cast<const float4*>(r);
    float acc = 0.0f;
#pragma unroll
    for (int c = 0; c < 16; ++c) {
        const float4 d4 = __ldg(dp + c);
        const float4 r4 = rp[c];
        acc = __fadd_rn(acc, __fmul_rn(d4.x, r4.x));
        acc = __fadd_rn(acc, __fmul_rn(d4.y, r4.y));
        acc = __fadd_rn(acc, __fmul_rn(d4.z, r4.z));
        acc = __fadd_rn(acc, __fmul_rn(d4.w, r4.w));
    }
    return acc;
}

}  // namespace

namespace {

__global__ void __launch_bounds__(kThreadsF16, 1) matchingPursuitF16Kernel(const MpParams p) {
    extern __shared__ __align__(1024) unsigned char smemRaw[];
    SmemF16& s = *reinterpret_cast<SmemF16*>(smemRaw);  // no pointer arithmetic on integers: accesses stay LDS/STS/ATOMS
    const int tid = threadIdx.x;
    const int lane = tid & 31;
    const int warp = tid >> 5;

    // ---- one-time setup: barriers, TMEM (all 512 columns), small tables, dictionary image via TMA bulk copies ----
    if (tid == 0) {
        mbarInit(&s.mbarDict, 1);
        mbarInit(&s.mbarBase, 1);
        for (int i = 0; i < kStages; ++i) mbarInit(&s.full[i], 1), mbarInit(&s.empty[i], 1);
        for (int i = 0; i < kAccs; ++i) mbarInit(&s.accFull[i], 1), mbarInit(&s.accEmpty[i], 4);
        s.nItems[0] = s.nItems[1] = 0;
        s.nDc[0] = s.nDc[1] = 0;
        s.nCand = 0;
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smemAddr(&s.tmemBase)), "r"(512u) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    for (int i = tid; i < 3 * kMaxK; i += kThreadsF16) s.quant[i] = p.quant[i];
    for (int i = tid; i < kRowsPad; i += kThreadsF16) s.detRows[i] = i < p.nBase ? p.detRows[i] : 0;
    fenceBefore();
    __syncthreads();
    fenceAfter();
    const uint32_t tmemBase = s.tmemBase;
    if (tid == 0) {
        constexpr uint32_t total = kRowsPad * kN * sizeof(__half);
        mbarExpectTx(&s.mbarDict, total + 6 * kGroupImgBytes);
        for (uint32_t o = 0; o < total; o += 32768)
            bulkCopy(reinterpret_cast<unsigned char*>(s.D16) + o, reinterpret_cast<const unsigned char*>(p.baseUm16) + o, 32768, &s.mbarDict);
        for (int ch = 0; ch < 3; ++ch)  // resident groups: atom 0 -> rows 0-63 of the channel's tile, atom nBase-1 -> rows 64-127
            for (int h = 0; h < 2; ++h)
                bulkCopy(s.dc[ch] + h * kGroupImgBytes,
                         reinterpret_cast<const unsigned char*>(p.detUm16) + (static_cast<size_t>(ch) * p.nBase + (h ? p.nBase - 1 : 0)) * kGroupImgBytes,
                         kGroupImgBytes, &s.mbarDict);
    }

    // ---- consumer-thread geometry (step C: half-warp `slot` owns slot `slot`; thread l16 owns pixels 4*l16 .. 4*l16+3) ----
    const int slot = (tid >> 4) & (kSlots - 1);
    const int l16 = tid & 15;
    const unsigned halfMask = 0xFFFFu << (lane & 16);
    float4* const myR32 = reinterpret_cast<float4*>(s.R32 + slot * kRStride + 4 * l16);
    unsigned char* const myR16 = reinterpret_cast<unsigned char*>(s.R16) + (slot >> 3) * kSbo + (l16 >> 1) * kLbo + (slot & 7) * 16 + (l16 & 1) * 8;
    unsigned long long statRows = 0, statPicks = 0;

    auto halfSum = [&](float v) {
#pragma unroll
        for (int off = 8; off > 0; off >>= 1) v += __shfl_xor_sync(halfMask, v, off);
        return v;
    };
    // residual of this thread's 4 pixels -> exact FP32 copy + fp16 hi / lo operand images; returns ||r||^2 and r != 0
    auto storeResidual = [&](const float4 r, bool& nonZero) {
        *myR32 = r;
        const __half hx = __float2half_rn(r.x), hy = __float2half_rn(r.y), hz = __float2half_rn(r.z), hw = __float2half_rn(r.w);
        const __half lx = __float2half_rn(r.x - __half2float(hx)), ly = __float2half_rn(r.y - __half2float(hy)),
                     lz = __float2half_rn(r.z - __half2float(hz)), lw = __float2half_rn(r.w - __half2float(hw));
        uint2 hi, lo;
        hi.x = static_cast<uint32_t>(__half_as_ushort(hx)) | (static_cast<uint32_t>(__half_as_ushort(hy)) << 16);
        hi.y = static_cast<uint32_t>(__half_as_ushort(hz)) | (static_cast<uint32_t>(__half_as_ushort(hw)) << 16);
        lo.x = static_cast<uint32_t>(__half_as_ushort(lx)) | (static_cast<uint32_t>(__half_as_ushort(ly)) << 16);
        lo.y = static_cast<uint32_t>(__half_as_ushort(lz)) | (static_cast<uint32_t>(__half_as_ushort(lw)) << 16);
        *reinterpret_cast<uint2*>(myR16) = hi;
        *reinterpret_cast<uint2*>(myR16 + kSlots * kN * sizeof(__half)) = lo;
        const bool nz = (r.x != 0.0f) | (r.y != 0.0f) | (r.z != 0.0f) | (r.w != 0.0f);
        nonZero = (__ballot_sync(halfMask, nz) & halfMask) != 0u;
        return halfSum(r.x * r.x + r.y * r.y + r.z * r.z + r.w * r.w);
    };
    auto epsOf = [&](float ss) { return 2.0f * (p.epsRel16 * sqrtf(ss) * 1.0001f + p.epsAbs16); };

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
        bool nonZero;
        const float ss = storeResidual(v, nonZero);
        if (l16 == 0) {
            s.bc[slot] = next < p.nbc ? static_cast<int>(next) : -1;
            s.reps[slot] = 0;
            s.prev[slot] = 0;
            s.ngroups[slot] = 0;
            s.height[slot] = p.nBase;
            s.best[slot] = 0ULL;
            s.approxMax[slot] = 0u;
            s.twoEps[slot] = epsOf(ss);
            s.zeroRes[slot] = nonZero ? 0 : 1;
        }
        return next < p.nbc;
    };
    auto pushCandidate = [&](int sl, int idx, uint32_t rowId) {
        const uint32_t pos = atomicAdd(&s.nCand, 1u);
        if (pos < kCandCap) {
            s.cand[pos] = make_uint2(static_cast<uint32_t>(sl) | (static_cast<uint32_t>(idx) << 8), rowId);
        } else {  // list full (pathological near-tie floods): re-score on the spot
            const float a = exactDot(p.rows32 + static_cast<size_t>(rowId) * kN, s.R32 + sl * kRStride);
            atomicMax(&s.best[sl], makeKey(a, idx));
        }
    };

    int active = 0;
    if (warp < kTmaWarp) {
        active = refill() ? 1 : 0;
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // residual images -> visible to the tensor core
    }

    const int tile = warp >> 2, quarter = warp & 3;  // TMEM lane quarter of this warp; base tile of a consumer warp
    const int myAtom = tile * 128 + quarter * 32 + lane;
    const uint32_t laneBase = static_cast<uint32_t>(quarter * 32) << 16;
    const uint32_t dAddr = smemAddr(s.D16), rAddr = smemAddr(s.R16);
    uint32_t pairBase = 0, trip = 0;
    // optional phase timing (MNENC_MP_TIMING=1): cycles of consumer thread 0 of CTA 0 per phase, printed at the end
    long long tPhase[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    long long tLast = 0;
    unsigned long long nItemsSum = 0, nDcSum = 0, nCandSum = 0;
    const bool timing = p.debugTiming && tid == 0 && blockIdx.x == 0;
#define MNENC_TICK(k)                     \
    if (timing) {                         \
        const long long now_ = clock64(); \
        tPhase[k] += now_ - tLast;        \
        tLast = now_;                     \
    }
    if (timing) tLast = clock64();

    while (true) {
        // ---- trip barrier: residual images, slot state and the item list of this trip are complete ----
        fenceBefore();
        if (!__syncthreads_or(active)) break;
        fenceAfter();
        const uint32_t cur = trip & 1u;
        const uint32_t nItems = s.nItems[cur];
        const uint32_t nDc = s.nDc[cur];
        MNENC_TICK(0)  // trip barrier
        if (timing) nItemsSum += nItems, nDcSum += nDc;
        const uint32_t nPairs = (nItems + 1u) >> 1;

        if (warp == kMmaWarp) {
            // ================= MMA issuer: base tiles, then one tile per landed pair of group images =================
            if (lane == 0) {
                if (trip == 0) mbarWait(&s.mbarDict, 0);
#pragma unroll
                for (int t = 0; t < 4; ++t)
#pragma unroll
                    for (int h = 0; h < 2; ++h)
#pragma unroll
                        for (int ks = 0; ks < 4; ++ks)
                            ummaF16(tmemBase + t * 32, umDesc(dAddr + t * 16 * kSbo + ks * kKStepBytes),
                                    umDesc(rAddr + h * (kSlots * kN * 2) + ks * kKStepBytes), kIdescF16, (h | ks) ? 1u : 0u);
                if (nDc) {
                    const uint32_t cAddr = smemAddr(s.dc[0]);
#pragma unroll
                    for (int t = 0; t < 3; ++t)
#pragma unroll
                        for (int h = 0; h < 2; ++h)
#pragma unroll
                            for (int ks = 0; ks < 4; ++ks)
                                ummaF16(tmemBase + kDcAccCol + t * 32, umDesc(cAddr + t * 2 * kGroupImgBytes + ks * kKStepBytes),
                                        umDesc(rAddr + h * (kSlots * kN * 2) + ks * kKStepBytes), kIdescF16, (h | ks) ? 1u : 0u);
                }
                ummaCommit(&s.mbarBase);
                for (uint32_t q = 0; q < nPairs; ++q) {
                    const uint32_t ctr = pairBase + q;
                    const uint32_t stage = ctr % kStages, su = ctr / kStages, acc = ctr % kAccs, au = ctr / kAccs;
                    mbarWait(&s.full[stage], su & 1u);
                    mbarWait(&s.accEmpty[acc], (au & 1u) ^ 1u);
                    fenceAfter();
                    const uint32_t aAddr = smemAddr(s.ring[stage]);
#pragma unroll
                    for (int h = 0; h < 2; ++h)
#pragma unroll
                        for (int ks = 0; ks < 4; ++ks)
                            ummaF16(tmemBase + kDetAccCol + acc * 32, umDesc(aAddr + ks * kKStepBytes),
                                    umDesc(rAddr + h * (kSlots * kN * 2) + ks * kKStepBytes), kIdescF16, (h | ks) ? 1u : 0u);
                    ummaCommit(&s.empty[stage]);   // ring stage free once these MMAs have read it
                    ummaCommit(&s.accFull[acc]);   // scores of the pair are in TMEM
                }
            }
        } else if (warp == kTmaWarp) {
            // ================= TMA producer: group images of this trip's items, L2 -> shared-memory ring =================
            if (lane == 0) {
                for (uint32_t q = 0; q < nPairs; ++q) {
                    const uint32_t ctr = pairBase + q;
                    const uint32_t stage = ctr % kStages, su = ctr / kStages;
                    mbarWait(&s.empty[stage], (su & 1u) ^ 1u);
                    const uint32_t n = min(2u, nItems - 2u * q);
                    mbarExpectTx(&s.full[stage], n * kGroupImgBytes);
                    for (uint32_t i = 0; i < n; ++i) {
                        const uint32_t it = s.items[cur][2u * q + i];
                        const uint32_t sl = it & 31u, gi = it >> 5;
                        const uint32_t atom = s.groupAtom[sl * kMaxK + gi];
                        const uint32_t ch = static_cast<uint32_t>(s.bc[sl] % 3);
                        bulkCopy(s.ring[stage] + i * kGroupImgBytes,
                                 reinterpret_cast<const unsigned char*>(p.detUm16) + (static_cast<size_t>(ch) * p.nBase + atom) * kGroupImgBytes,
                                 kGroupImgBytes, &s.full[stage]);
                    }
                }
            }
        } else {
            // ================= consumers =================
            if (tid == 0) {
                s.nItems[cur ^ 1u] = 0;
                s.nDc[cur ^ 1u] = 0;
                s.nCand = 0;
            }
            const unsigned liveMask = __ballot_sync(kFull, s.bc[lane] >= 0 && !s.zeroRes[lane]);
            // ---- (a) base scores: TMEM -> registers, per-slot maximum ----
            mbarWait(&s.mbarBase, trip & 1u);
            fenceAfter();
            float v[32];
            {
                uint32_t raw[32];
                tmemLoad32(tmemBase + laneBase + static_cast<uint32_t>(tile * 32), raw);
#pragma unroll
                for (int i = 0; i < 32; ++i) v[i] = fabsf(__uint_as_float(raw[i]));
            }
            fenceBefore();
            MNENC_TICK(1)  // wait for the base MMA + TMEM load
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
            consumerBarrier();
            MNENC_TICK(2)  // butterfly + barrier
            // ---- (b) base atoms that can still be the exact maximum -> candidate list ----
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
                cand &= liveMask;
                while (cand) {  // re-scored on the spot: this time is otherwise spent waiting for the streamed groups
                    const int sl = __ffs(cand) - 1;
                    cand &= cand - 1;
                    const float a = exactDot(p.rows32 + static_cast<size_t>(myAtom) * kN, s.R32 + sl * kRStride);
                    atomicMax(&s.best[sl], makeKey(a, myAtom));
                }
            }
            MNENC_TICK(3)  // base filter + inline re-score
            // ---- (c0) resident groups (atoms 0 / nBase-1): scores of ALL slots come from one shared tile per channel ----
            for (uint32_t i = static_cast<uint32_t>(tile); i < nDc; i += 4) {
                const uint32_t it = s.dcItems[cur][i];
                if (static_cast<int>(it >> 10) != (quarter >> 1)) continue;  // warp-uniform: lanes 0-63 atom 0, 64-127 atom nBase-1
                const int sl = static_cast<int>(it & 31u), gi = static_cast<int>((it >> 5) & 31u);
                const int ch = s.bc[sl] % 3;
                const uint32_t raw = tmemLoad1(tmemBase + laneBase + kDcAccCol + static_cast<uint32_t>(ch * 32 + sl));
                const int atom = s.groupAtom[sl * kMaxK + gi];
                const int first = s.groupFirst[sl * kMaxK + gi];
                const int row = (quarter & 1) * 32 + lane;
                const bool rowOk = row < s.detRows[atom];
                const float a = rowOk ? fabsf(__uint_as_float(raw)) : 0.0f;
                float wmax = a;
#pragma unroll
                for (int off = 16; off > 0; off >>= 1) wmax = fmaxf(wmax, __shfl_xor_sync(kFull, wmax, off));
                uint32_t seen = 0;
                if (lane == 0) seen = max(atomicMax(&s.approxMax[sl], __float_as_uint(wmax)), __float_as_uint(wmax));
                seen = __shfl_sync(kFull, seen, 0);
                if (rowOk && ((liveMask >> sl) & 1u) && a + s.twoEps[sl] >= __uint_as_float(seen))
                    pushCandidate(sl, first + row, static_cast<uint32_t>(kRowsPad + (ch * p.nBase + atom) * 64 + row));
            }
            fenceBefore();
            MNENC_TICK(4)  // resident groups
            // ---- (c) detail pairs: warpgroup g takes pairs g, g+4, ...; quarters 0,1 -> first item, 2,3 -> second ----
            for (uint32_t q = static_cast<uint32_t>(tile); q < nPairs; q += 4) {
                const uint32_t ctr = pairBase + q;
                const uint32_t acc = ctr % kAccs, au = ctr / kAccs;
                const uint32_t itemIdx = 2u * q + static_cast<uint32_t>(quarter >> 1);
                const bool valid = itemIdx < nItems;
                const uint32_t it = valid ? s.items[cur][itemIdx] : 0u;
                const int sl = static_cast<int>(it & 31u), gi = static_cast<int>(it >> 5);
                mbarWait(&s.accFull[acc], au & 1u);
                fenceAfter();
                const uint32_t raw = tmemLoad1(tmemBase + laneBase + kDetAccCol + acc * 32 + static_cast<uint32_t>(sl));
                fenceBefore();
                __syncwarp();
                if (lane == 0) mbarArrive(&s.accEmpty[acc]);
                if (valid) {
                    const int atom = s.groupAtom[sl * kMaxK + gi];
                    const int first = s.groupFirst[sl * kMaxK + gi];
                    const int row = (quarter & 1) * 32 + lane;
                    const bool rowOk = row < s.detRows[atom];
                    const float a = rowOk ? fabsf(__uint_as_float(raw)) : 0.0f;
                    float wmax = a;
#pragma unroll
                    for (int off = 16; off > 0; off >>= 1) wmax = fmaxf(wmax, __shfl_xor_sync(kFull, wmax, off));
                    uint32_t seen = 0;
                    if (lane == 0) seen = max(atomicMax(&s.approxMax[sl], __float_as_uint(wmax)), __float_as_uint(wmax));
                    seen = __shfl_sync(kFull, seen, 0);
                    if (rowOk && ((liveMask >> sl) & 1u) && a + s.twoEps[sl] >= __uint_as_float(seen)) {
                        const int ch = s.bc[sl] % 3;
                        pushCandidate(sl, first + row, static_cast<uint32_t>(kRowsPad + (ch * p.nBase + atom) * 64 + row));
                    }
                }
            }
            consumerBarrier();
            MNENC_TICK(5)  // streamed pairs + barrier
            // ---- (d) exact re-score of the candidates, one per thread ----
            {
                const uint32_t n = min(s.nCand, static_cast<uint32_t>(kCandCap));
                if (timing) nCandSum += s.nCand;
                for (uint32_t i = tid; i < n; i += kConsumerThreads) {
                    const uint2 c = s.cand[i];
                    const int sl = static_cast<int>(c.x & 31u);
                    const float a = exactDot(p.rows32 + static_cast<size_t>(c.y) * kN, s.R32 + sl * kRStride);
                    atomicMax(&s.best[sl], makeKey(a, static_cast<int>(c.x >> 8)));
                }
            }
            consumerBarrier();
            MNENC_TICK(6)  // candidate re-score + barrier
            // ================= step C: select, quantise, update (MatchingPursuit.cpp:86-104) =================
            active = 0;
            const int bc = s.bc[slot];
            if (bc >= 0) {
                const unsigned long long key = s.zeroRes[slot] ? makeKey(0.0f, 0) : s.best[slot];  // r == 0: every score is +0, atom 0 wins
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
                    uint32_t rowId = static_cast<uint32_t>(idx);
                    if (idx >= p.nBase) {
                        int gi = ng - 1;
                        while (gi > 0 && s.groupFirst[slot * kMaxK + gi] > idx) --gi;
                        const int atom = s.groupAtom[slot * kMaxK + gi];
                        rowId = static_cast<uint32_t>(kRowsPad + (ch * p.nBase + atom) * 64 + (idx - s.groupFirst[slot * kMaxK + gi]));
                    }
                    const float4 row = __ldg(reinterpret_cast<const float4*>(p.rows32 + static_cast<size_t>(rowId) * kN) + l16);
                    float4 r = *myR32;
                    r.x = __fsub_rn(r.x, __fmul_rn(c, row.x));
                    r.y = __fsub_rn(r.y, __fmul_rn(c, row.y));
                    r.z = __fsub_rn(r.z, __fmul_rn(c, row.z));
                    r.w = __fsub_rn(r.w, __fmul_rn(c, row.w));
                    bool nonZero;
                    const float ss = storeResidual(r, nonZero);
                    const int ngNew = ng + (idx < p.nBase ? 1 : 0);
                    if (l16 == 0) {
                        s.pairs[slot * kMaxK + reps] = deltaId | ((zigzag(qi) & 0xFFFFu) << 16);
                        s.prev[slot] = idx;
                        s.reps[slot] = reps + 1;
                        s.best[slot] = 0ULL;
                        s.approxMax[slot] = 0u;
                        s.twoEps[slot] = epsOf(ss);
                        s.zeroRes[slot] = nonZero ? 0 : 1;
                        if (idx < p.nBase) {
                            s.groupAtom[slot * kMaxK + ng] = static_cast<uint16_t>(idx);
                            s.groupFirst[slot * kMaxK + ng] = static_cast<uint16_t>(height);
                            s.ngroups[slot] = ngNew;
                            s.height[slot] = height + s.detRows[idx];
                        }
                    }
                    if (reps + 1 == p.K) {
                        count = p.K;
                    } else if (ngNew > 0) {  // this slot's (slot, group) items of the next trip: resident or streamed
                        __syncwarp(halfMask);
                        for (int g0 = 0; g0 < ngNew; g0 += 16) {
                            const int gi = g0 + l16;
                            const bool valid = gi < ngNew;
                            const int atom = valid ? s.groupAtom[slot * kMaxK + gi] : -1;
                            const bool isDc = valid && (atom == 0 || atom == p.nBase - 1);
                            const unsigned mDc = (__ballot_sync(halfMask, isDc) & halfMask) >> (lane & 16);
                            const unsigned mSt = (__ballot_sync(halfMask, valid && !isDc) & halfMask) >> (lane & 16);
                            uint32_t atDc = 0, atSt = 0;
                            if (l16 == 0) {
                                if (mDc) atDc = atomicAdd(&s.nDc[cur ^ 1u], static_cast<uint32_t>(__popc(mDc)));
                                if (mSt) atSt = atomicAdd(&s.nItems[cur ^ 1u], static_cast<uint32_t>(__popc(mSt)));
                            }
                            atDc = __shfl_sync(halfMask, atDc, lane & 16);
                            atSt = __shfl_sync(halfMask, atSt, lane & 16);
                            const unsigned below = (1u << l16) - 1u;
                            if (isDc)
                                s.dcItems[cur ^ 1u][atDc + __popc(mDc & below)] = static_cast<uint16_t>(slot | (gi << 5) | ((atom != 0 ? 1 : 0) << 10));
                            else if (valid)
                                s.items[cur ^ 1u][atSt + __popc(mSt & below)] = static_cast<uint16_t>(slot | (gi << 5));
                        }
                    }
                }
                if (count >= 0) {
                    __syncwarp(halfMask);
                    if (l16 == 0) p.counts[bc] = static_cast<uint16_t>(count);
                    for (int i = l16; i < p.K; i += 16) p.pairs[static_cast<size_t>(bc) * p.K + i] = i < count ? s.pairs[slot * kMaxK + i] : 0u;
                    statPicks += (l16 == 0) ? count : 0;
                    __syncwarp(halfMask);
                    active = refill() ? 1 : 0;
                } else {
                    active = 1;
                }
            }
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            MNENC_TICK(7)  // step C
        }
        pairBase += nPairs;
        ++trip;
    }
    if (warp < kTmaWarp) {
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
    if (timing)
        printf("mp f16 timing, CTA 0: trips %u; cycles per trip: barrier %lld, baseMMA+ld %lld, max+bar %lld, base filter %lld, resident %lld, "
               "streamed+bar %lld, cand+bar %lld, stepC %lld; per trip: streamed items %.1f, resident items %.1f, listed candidates %.1f\n",
               trip, tPhase[0] / trip, tPhase[1] / trip, tPhase[2] / trip, tPhase[3] / trip, tPhase[4] / trip, tPhase[5] / trip, tPhase[6] / trip,
               tPhase[7] / trip, double(nItemsSum) / trip, double(nDcSum) / trip, double(nCandSum) / trip);
#undef MNENC_TICK
    fenceBefore();
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmemBase), "r"(512u) : "memory");
}

}  // namespace

// ------------------------------------------------------------------------------------------------
// Host: operand images. umOff16 = element index in a canonical K-major image of fp16 (see file header).
// ------------------------------------------------------------------------------------------------
static inline size_t umOff16(int row, int k) { return static_cast<size_t>(row >> 3) * 512 + (k >> 3) * 64 + (row & 7) * 8 + (k & 7); }

void buildF16Images(const float* baseRowMajor, int nBase, const int32_t* detailRows, const float* const detail[3], uint16_t* baseUm16,
                    uint16_t* detUm16, float* rows32, double* maxRoundingNorm, double* maxNorm) {
    double worst = 0.0, worstNorm = 0.0;
    auto put = [&](uint16_t* image, int row, const float* src, float* exactRow) {
        double err = 0.0, nrm = 0.0;
        for (int k = 0; k < kN; ++k) {
            const __half h = __float2half_rn(src[k]);
            image[umOff16(row, k)] = __half_as_ushort(h);
            const double e = static_cast<double>(src[k]) - static_cast<double>(__half2float(h));
            err += e * e;
            nrm += static_cast<double>(src[k]) * src[k];
            exactRow[k] = src[k];
        }
        worst = std::max(worst, std::sqrt(err));
        worstNorm = std::max(worstNorm, std::sqrt(nrm));
    };
    std::fill(baseUm16, baseUm16 + static_cast<size_t>(kRowsPad) * kN, uint16_t(0));
    std::fill(rows32, rows32 + (static_cast<size_t>(kRowsPad) + static_cast<size_t>(3) * nBase * 64) * kN, 0.0f);
    std::fill(detUm16, detUm16 + static_cast<size_t>(3) * nBase * 64 * kN, uint16_t(0));
    for (int a = 0; a < nBase; ++a) put(baseUm16, a, baseRowMajor + static_cast<size_t>(a) * kN, rows32 + static_cast<size_t>(a) * kN);
    for (int ch = 0; ch < 3; ++ch) {
        size_t rowOff = 0;
        for (int a = 0; a < nBase; ++a) {
            const size_t g = static_cast<size_t>(ch) * nBase + a;
            for (int r = 0; r < detailRows[a]; ++r)
                put(detUm16 + g * 64 * kN, r, detail[ch] + (rowOff + r) * kN, rows32 + (static_cast<size_t>(kRowsPad) + g * 64 + r) * kN);
            rowOff += detailRows[a];
        }
    }
    *maxRoundingNorm = worst;
    *maxNorm = worstNorm;
}

size_t f16Rows32Count(int nBase) { return static_cast<size_t>(kRowsPad) + static_cast<size_t>(3) * nBase * 64; }

cudaError_t launchMatchingPursuitF16(const MpParams& p, int numSMs, cudaStream_t stream) {
    static bool configured = false;
    const size_t smem = sizeof(SmemF16) + 1024;
    if (!configured) {
        cudaError_t e = cudaFuncSetAttribute(matchingPursuitF16Kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
        if (e != cudaSuccess) return e;
        configured = true;
    }
    if (p.nbc == 0) return cudaSuccess;
    unsigned int grid = (p.nbc + kSlots - 1) / kSlots;
    if (grid > static_cast<unsigned int>(numSMs)) grid = numSMs;
    matchingPursuitF16Kernel<<<grid, kThreadsF16, smem, stream>>>(p);
    return cudaGetLastError();
}

}  // namespace mnenc
