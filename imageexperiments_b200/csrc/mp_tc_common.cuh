// Device helpers shared by the tensor-core matching-pursuit kernels: mbarrier / TMA bulk copy / tcgen05 wrappers
// (inline PTX for sm_100a), the exact oracle arithmetic and the arg-max key.
#pragma once

#include <cuda_fp16.h>
#include <cuda_runtime.h>

#include <cstdint>

namespace mnenc {
namespace tc {

constexpr unsigned kFullMask = 0xFFFFFFFFu;

// 64-bit arg-max key: max(key) == "largest |p|, lowest index on ties" == Eigen's visitor with strict > (Visitor.h:227)
__device__ __forceinline__ unsigned long long makeKey(float p, int idx) {
    const unsigned int b = __float_as_uint(p);
    return (static_cast<unsigned long long>(b & 0x7FFFFFFFu) << 32) | (static_cast<unsigned long long>(0x7FFFFFFFu - idx) << 1) |
           (b >> 31);
}
__device__ __forceinline__ int keyIndex(unsigned long long key) { return static_cast<int>(0x7FFFFFFFu - (static_cast<unsigned int>(key) >> 1)); }
__device__ __forceinline__ float keyValue(unsigned long long key) {
    return __uint_as_float(static_cast<unsigned int>(key >> 32) | (static_cast<unsigned int>(key) << 31));
}
__device__ __forceinline__ uint32_t zigzag(int x) { return (static_cast<uint32_t>(x) << 1) ^ static_cast<uint32_t>(x >> 31); }
__device__ __forceinline__ uint32_t smemAddr(const void* p) { return static_cast<uint32_t>(__cvta_generic_to_shared(p)); }

// img::YUVFromRGB in double, left-associative, then cast (ImageHelper/src/misc.cpp:12-21, misc.h:7-11)
__device__ __forceinline__ float yuvChannel(int ch, double R, double G, double B) {
    const double y = __dadd_rn(__dadd_rn(__dmul_rn(0.299, R), __dmul_rn(0.587, G)), __dmul_rn(0.114, B));
    if (ch == 0) return __double2float_rn(y);
    if (ch == 1) return __double2float_rn(__dmul_rn(0.436 / (1.0 - 0.114), __dsub_rn(B, y)));
    return __double2float_rn(__dmul_rn(0.615 / (1.0 - 0.299), __dsub_rn(R, y)));
}

// exact oracle dot product of one dictionary row (global, row-major FP32) with a residual in shared memory:
// acc = fl(acc + fl(d*r)), j ascending, never fused (Eigen GEMV without FMA, GeneralMatrixVector.h:108-215)
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

// ---- mbarrier ----
__device__ __forceinline__ void mbarInit(unsigned long long* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smemAddr(bar)), "r"(count));
}
__device__ __forceinline__ bool mbarTest(unsigned long long* bar, uint32_t parity) {
    uint32_t done;
    asm volatile("{\n\t.reg .pred P;\n\tmbarrier.try_wait.parity.shared::cta.b64 P, [%1], %2;\n\tselp.u32 %0, 1, 0, P;\n\t}"
                 : "=r"(done)
                 : "r"(smemAddr(bar)), "r"(parity)
                 : "memory");
    return done != 0;
}
__device__ __forceinline__ void mbarWait(unsigned long long* bar, uint32_t parity) {
    while (!mbarTest(bar, parity)) {
    }
}
__device__ __forceinline__ void mbarArrive(unsigned long long* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smemAddr(bar)) : "memory");
}
__device__ __forceinline__ void mbarExpectTx(unsigned long long* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smemAddr(bar)), "r"(bytes) : "memory");
}
// TMA bulk copy global -> shared, completion counted in bytes on `bar` (SASS: UBLKCP.S.G)
__device__ __forceinline__ void bulkCopy(void* dstSmem, const void* srcGlobal, uint32_t bytes, unsigned long long* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smemAddr(dstSmem)),
                 "l"(srcGlobal), "r"(bytes), "r"(smemAddr(bar))
                 : "memory");
}
// warp-convergent variants (every lane executes the call with warp-uniform operands, one elected lane issues; see ummaF16Elect)
__device__ __forceinline__ void mbarExpectTxElect(unsigned long long* bar, uint32_t bytes) {
    asm volatile("{\n\t.reg .pred q;\n\telect.sync _|q, 0xffffffff;\n\t@q mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;\n\t}" ::"r"(smemAddr(bar)),
                 "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void bulkCopyElect(void* dstSmem, const void* srcGlobal, uint32_t bytes, unsigned long long* bar) {
    asm volatile(
        "{\n\t.reg .pred q;\n\telect.sync _|q, 0xffffffff;\n\t"
        "@q cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];\n\t}" ::"r"(smemAddr(dstSmem)),
        "l"(srcGlobal), "r"(bytes), "r"(smemAddr(bar))
        : "memory");
}
__device__ __forceinline__ void fenceProxyAsync() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// ---- tcgen05 ----
// shared-memory matrix descriptor: K-major, no swizzle, version 1 (cute::UMMA::SmemDescriptor); offsets in 16-byte units
__device__ __forceinline__ unsigned long long umDesc(uint32_t saddr, uint32_t lboBytes, uint32_t sboBytes) {
    return static_cast<unsigned long long>((saddr & 0x3FFFFu) >> 4) | (static_cast<unsigned long long>(lboBytes >> 4) << 16) |
           (static_cast<unsigned long long>(sboBytes >> 4) << 32) | (1ull << 46);
}
// D[tmem] (+)= A[smem] * B[smem], kind::f16 (SASS: UTCHMMA)
__device__ __forceinline__ void ummaF16(uint32_t tmemD, unsigned long long descA, unsigned long long descB, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmemD),
        "l"(descA), "l"(descB), "r"(idesc), "r"(accumulate)
        : "memory");
}
// A warp-uniform value the compiler KNOWS is uniform: REDUX delivers its result in a uniform register, so descriptors built
// from it stay in the uniform datapath (no R2UR.BROADCAST per MMA).
__device__ __forceinline__ uint32_t warpUniform(uint32_t v) { return __reduce_max_sync(kFullMask, v); }
// Warp-convergent variants: EVERY lane of the issuing warp executes the call with warp-uniform operands, one elected lane
// issues. With the whole issue path inside `if (lane == 0)` the compiler cannot keep descriptors in uniform registers and
// wraps each UTCHMMA in an ELECT / R2UR.BROADCAST / BRA.U.ANY loop: ~140 cycles per MMA, measured - the issuing thread, not
// the tensor pipe, was the limit.
__device__ __forceinline__ void ummaF16Elect(uint32_t tmemD, unsigned long long descA, unsigned long long descB, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p, q;\n\tsetp.ne.b32 p, %4, 0;\n\telect.sync _|q, 0xffffffff;\n\t"
        "@q tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmemD),
        "l"(descA), "l"(descB), "r"(idesc), "r"(accumulate)
        : "memory");
}
__device__ __forceinline__ void ummaCommitElect(unsigned long long* bar) {
    asm volatile(
        "{\n\t.reg .pred q;\n\telect.sync _|q, 0xffffffff;\n\t"
        "@q tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n\t}" ::"r"(smemAddr(bar))
        : "memory");
}
__device__ __forceinline__ void ummaCommit(unsigned long long* bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smemAddr(bar)) : "memory");
}
__device__ __forceinline__ void fenceBefore() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void fenceAfter() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
// 32 lanes x 32 columns of 32 bit: thread = TMEM lane, register i = column i (SASS: LDTM.x32)
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
__device__ __forceinline__ void tmemLoad32NoWait(uint32_t taddr, uint32_t (&v)[32]) {  // issue several, then one tmemWaitLd()
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
}
// 32 lanes x 16 columns (SASS: LDTM.x16)
__device__ __forceinline__ void tmemLoad16(uint32_t taddr, uint32_t (&v)[16]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), "=r"(v[9]),
          "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
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
// same without the wait: issue several, then one tmemWaitLd()
__device__ __forceinline__ uint32_t tmemLoad1NoWait(uint32_t taddr) {
    uint32_t v;
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x1.b32 {%0}, [%1];" : "=r"(v) : "r"(taddr) : "memory");
    return v;
}
__device__ __forceinline__ void tmemWaitLd() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
template <int kId, int kThreads>
__device__ __forceinline__ void namedBarrier() {
    asm volatile("bar.sync %0, %1;" ::"n"(kId), "n"(kThreads) : "memory");
}
__device__ __forceinline__ void namedBarrier(int id, int threads) { asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(threads) : "memory"); }

// M = 128, N = 32, A/B = F16 K-major, D = F32 (cute::UMMA::InstrDescriptor bit layout)
constexpr uint32_t kIdescF16M128N16 = (1u << 4) | (0u << 7) | (0u << 10) | ((16u >> 3) << 17) | ((128u >> 4) << 24);
constexpr uint32_t kIdescF16M128N32 = (1u << 4) | (0u << 7) | (0u << 10) | ((32u >> 3) << 17) | ((128u >> 4) << 24);

}  // namespace tc
}  // namespace mnenc
