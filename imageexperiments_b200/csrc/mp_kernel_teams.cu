// Matching-pursuit kernel, version 4 (sm_100a): two independent TEAMS per CTA work out of phase, so the latency
// chain of one greedy iteration (MMA -> TMEM read -> arg-max -> exact re-score -> residual update) of one team is
// covered by the other team's work. Arithmetic, pre-filter bound and exactness argument are those of
// mp_kernel_f16.cu:
//   * approximate scores of ALL dictionary rows on the tensor cores (tcgen05.mma kind::f16, fp16 dictionary images,
//     residual rounded to fp16 (hi part only by default; MNENC_MP_LO=1 adds the lo part as a second pass), FP32 accumulators in TMEM),
//   * every row within 2*eps of the approximate maximum is re-scored with the oracle's exact FP32 arithmetic
//     (tc::exactDot) and the exact 64-bit key decides (SelectFast, MatchingPursuit.cpp:27-37),
//   * quantisation and residual update exact (MatchingPursuit.cpp:86-104).
//
// A team = 8 consumer warps + 1 TMA warp + 1 MMA warp, 32 block-channels ("slots") in flight, all of ONE colour
// channel at a time (Y, then U, then V): the KLT groups of atoms 0 and nBase-1 of that channel stay resident in
// shared memory as one extra MMA tile shared by all slots; every other appended group (dynamicBasisFast,
// CompressedImage.cpp:252-289) streams from L2 through a 2-stage TMA ring, two 8 KB group images per MMA tile.
// Block indices come from a per-channel global counter through a small shared-memory FIFO kept full by the TMA warp;
// the pixels of a slot's NEXT block are prefetched into registers while the current one is being coded
// (block gather + YUVFromRGB: CompressedImage.cpp:595-611, misc.cpp:12-21).
//
// TMEM (512 columns): team t owns columns [256 t, 256 t + 256): 4 base tiles x 32 slots, 1 resident-group tile,
// 6 accumulators (16 columns: one half of the slots) for streamed pairs. Shared memory: 64 KB base image + 2 x 16 KB resident tiles + 2 x 61 KB team state.
#include "mnenc_internal.h"
#include "mp_tc_common.cuh"

#include <algorithm>
#include <cstdio>

namespace mnenc {

namespace {

using namespace tc;

constexpr int kTeams = 2;
constexpr int kTeamConsumers = 256;                 // 8 warps
constexpr int kTeamThreads = kTeamConsumers + 64;   // + TMA warp + MMA warp
constexpr int kThreadsV4 = kTeams * kTeamThreads;   // 640: warps 0-15 consumers, 16/17 TMA, 18/19 MMA
constexpr int kTStages = 2;                         // ring stages of 16 KB (two group images)
constexpr int kTAccs = 6;                           // pair accumulators per team (16 columns each: a pair serves one half of the slots)
constexpr uint32_t kImgBytes = 8192;                // one fp16 group image: 64 rows x 64
constexpr uint32_t kSbo = 1024, kLbo = 128, kKStep = 256;
constexpr int kItemCap = kSlots * kMaxK;            // streamed items grow from index 0, resident items from the top
constexpr int kCandCapT = 256;
constexpr int kRStr = 68;                           // floats per slot in the FP32 residual array
constexpr int kFifo = 128;                          // block-index FIFO entries
constexpr uint32_t kNone = 0xFFFFFFFFu;
constexpr uint32_t kColDc = 128, kColAcc = 160, kColsPerTeam = 256;

struct TeamSmem {
    alignas(1024) __half R16[2 * kSlots * kN];               // residual image: rows 0-31 hi, 32-63 lo (8 KB)
    alignas(1024) unsigned char ring[kTStages][2 * kImgBytes];
    alignas(16) float R32[kSlots * kRStr];                   // exact residuals
    unsigned long long best[kSlots];                         // exact arg-max key per slot
    unsigned long long mbarBase, mbarBase2, mbarDc, mbarDcLoad;  // scores complete: base tiles 0-1 / base tiles 2-3 / resident tile
    unsigned long long full[kTStages], empty[kTStages];
    unsigned long long accFull[kTAccs], accEmpty[kTAccs];
    alignas(16) uint32_t approxMax[kSlots];                  // max |p~| per slot (float bits)
    alignas(16) float twoEps[kSlots];
    uint32_t nItems[2], nItemsHi[2], nActive[2];           // streamed items of slots 0-15 (list grows from 0) / 16-31 (from the top)
    uint16_t dcFirst[2][kSlots];                             // dictionary index of row 0 of the slot's FIRST group of atom 0 / atom nBase-1
                                                             // (0xFFFF: none). A repeated group has identical rows, so on the exact
                                                             // ties it produces the lower index - the first occurrence - always wins.
    uint32_t nCand;
    uint32_t fifoHead, fifoTail;
    uint32_t fifo[kFifo];
    uint16_t items[2][kItemCap];                             // slot | group position << 5
    uint2 cand[kCandCapT];                                   // {slot | dictionary index << 8, row id in rows32}
    uint16_t groupAtom[kSlots * kMaxK];
    uint16_t groupFirst[kSlots * kMaxK];
    int bc[kSlots];
    int reps[kSlots];
    int prev[kSlots];
    int ngroups[kSlots];
    int height[kSlots];
    uint8_t zeroRes[kSlots];
};

struct __align__(1024) SmemV4 {
    __half D16[kRowsPad * kN];                               // 64 KB base dictionary image (shared by both teams)
    alignas(1024) unsigned char dc[kTeams][2 * kImgBytes];   // resident tile of each team's current channel
    TeamSmem team[kTeams];
    unsigned long long mbarDict;
    uint32_t tmemBase;
    float quant[3 * kMaxK];
    uint8_t detRows[kRowsPad];
};
static_assert(sizeof(SmemV4) + 1024 <= 232448, "shared memory of the team kernel exceeds 227 KB");

}  // namespace

namespace {

__device__ __forceinline__ uint32_t rawByte(const uint32_t (&w)[6], int k) { return (w[k >> 2] >> (8 * (k & 3))) & 0xFFu; }

__global__ void __launch_bounds__(kThreadsV4, 1) matchingPursuitTeamsKernel(const MpParams p) {
    extern __shared__ __align__(1024) unsigned char smemRaw[];
    SmemV4& s = *reinterpret_cast<SmemV4*>(smemRaw);  // no integer pointer arithmetic: accesses stay LDS/STS/ATOMS
    const int tid = threadIdx.x;
    const int lane = tid & 31;
    const int warp = tid >> 5;
    const bool isConsumer = warp < 16;
    const bool isTma = warp == 16 || warp == 17;
    const bool isMma = warp >= 18;
    const int team = isConsumer ? (warp >> 3) : ((warp - 16) & 1);
    TeamSmem& T = s.team[team];
    const uint32_t nblocks = static_cast<uint32_t>(p.nbx) * static_cast<uint32_t>(p.nby);

    // ---- one-time setup ----
    if (tid == 0) {
        mbarInit(&s.mbarDict, 1);
        for (int t = 0; t < kTeams; ++t) {
            TeamSmem& X = s.team[t];
            mbarInit(&X.mbarBase, 1);
            mbarInit(&X.mbarBase2, 1);
            mbarInit(&X.mbarDc, 1);
            mbarInit(&X.mbarDcLoad, 1);
            for (int i = 0; i < kTStages; ++i) mbarInit(&X.full[i], 1), mbarInit(&X.empty[i], 1);
            for (int i = 0; i < kTAccs; ++i) mbarInit(&X.accFull[i], 1), mbarInit(&X.accEmpty[i], 4);
            X.nItems[0] = X.nItems[1] = X.nItemsHi[0] = X.nItemsHi[1] = X.nActive[0] = X.nActive[1] = 0;
            X.nCand = 0;
            X.fifoHead = X.fifoTail = 0;
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smemAddr(&s.tmemBase)), "r"(512u) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    for (int i = tid; i < 3 * kMaxK; i += kThreadsV4) s.quant[i] = p.quant[i];
    for (int i = tid; i < kRowsPad; i += kThreadsV4) s.detRows[i] = i < p.nBase ? p.detRows[i] : 0;
    for (int i = tid; i < kTeams * kSlots; i += kThreadsV4) {
        TeamSmem& X = s.team[i / kSlots];
        const int sl = i % kSlots;
        X.bc[sl] = -1, X.reps[sl] = 0, X.prev[sl] = 0, X.ngroups[sl] = 0, X.height[sl] = p.nBase;
        X.best[sl] = 0ULL, X.approxMax[sl] = 0u, X.twoEps[sl] = 0.0f, X.zeroRes[sl] = 1;
        X.dcFirst[0][sl] = X.dcFirst[1][sl] = 0xFFFFu;
    }
    fenceBefore();
    __syncthreads();
    fenceAfter();
    const uint32_t tmemT = s.tmemBase + static_cast<uint32_t>(team) * kColsPerTeam;
    if (tid == 0) {
        constexpr uint32_t total = kRowsPad * kN * sizeof(__half);
        mbarExpectTx(&s.mbarDict, total);
        for (uint32_t o = 0; o < total; o += 32768)
            bulkCopy(reinterpret_cast<unsigned char*>(s.D16) + o, reinterpret_cast<const unsigned char*>(p.baseUm16) + o, 32768, &s.mbarDict);
    }

    // ---- consumer geometry: 8 threads per slot, thread l8 owns pixel row l8 of the block = vector elements 8*l8 .. 8*l8+7 ----
    const int tidT = tid & (kTeamConsumers - 1);
    const int slot = tidT >> 3;
    const int l8 = tidT & 7;
    const unsigned groupMask = 0xFFu << (lane & 24);
    const int leaderLane = lane & 24;
    float* const myR32 = T.R32 + slot * kRStr + 8 * l8;
    unsigned char* const myR16 = reinterpret_cast<unsigned char*>(T.R16) + (slot >> 3) * kSbo + l8 * kLbo + (slot & 7) * 16;
    const int quarter = warp & 3, wset = (warp >> 2) & 1;
    const uint32_t laneBase = static_cast<uint32_t>(quarter * 32) << 16;
    const uint32_t rAddr = smemAddr(T.R16);
    unsigned long long statRows = 0, statPicks = 0;
    uint32_t nextBlk = kNone;  // block whose pixels sit in `pix` (prefetched), kNone = nothing prefetched
    uint32_t pix[6] = {0, 0, 0, 0, 0, 0};
    int ch = -1;

    auto groupSum = [&](float v) {
        v += __shfl_xor_sync(groupMask, v, 4);
        v += __shfl_xor_sync(groupMask, v, 2);
        v += __shfl_xor_sync(groupMask, v, 1);
        return v;
    };
    // residual elements of this thread -> exact FP32 copy + fp16 hi / lo operand images; returns ||r||^2 of the slot
    auto storeResidual = [&](const float (&r)[8], bool& nonZero) {
        reinterpret_cast<float4*>(myR32)[0] = make_float4(r[0], r[1], r[2], r[3]);
        reinterpret_cast<float4*>(myR32)[1] = make_float4(r[4], r[5], r[6], r[7]);
        uint32_t hi[4], lo[4] = {0u, 0u, 0u, 0u};
        float ss = 0.0f;
        bool nz = false;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const __half h0 = __float2half_rn(r[2 * i]), h1 = __float2half_rn(r[2 * i + 1]);
            hi[i] = static_cast<uint32_t>(__half_as_ushort(h0)) | (static_cast<uint32_t>(__half_as_ushort(h1)) << 16);
            if (p.residualLo) {
                const __half l0 = __float2half_rn(r[2 * i] - __half2float(h0)), l1 = __float2half_rn(r[2 * i + 1] - __half2float(h1));
                lo[i] = static_cast<uint32_t>(__half_as_ushort(l0)) | (static_cast<uint32_t>(__half_as_ushort(l1)) << 16);
            }
            ss += r[2 * i] * r[2 * i] + r[2 * i + 1] * r[2 * i + 1];
            nz |= (r[2 * i] != 0.0f) | (r[2 * i + 1] != 0.0f);
        }
        *reinterpret_cast<uint4*>(myR16) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
        if (p.residualLo) *reinterpret_cast<uint4*>(myR16 + kSlots * kN * sizeof(__half)) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
        nonZero = (__ballot_sync(groupMask, nz) & groupMask) != 0u;
        return groupSum(ss);
    };
    auto epsOf = [&](float ss) { return 2.0f * (p.epsRel16 * sqrtf(ss) * 1.0001f + p.epsAbs16); };
    // 24 raw bytes of pixel row l8 of block `blk` (zeros outside the image: CompressedImage.cpp:599-609)
    auto loadPixels = [&](uint32_t blk, uint32_t (&w)[6]) {
#pragma unroll
        for (int i = 0; i < 6; ++i) w[i] = 0u;
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
    };
    // Start block-channel (blk, ch) in this slot from raw pixels; returns nothing, state written by the group leader.
    auto startBlock = [&](uint32_t blk, const uint32_t (&w)[6]) {
        float r[8];
#pragma unroll
        for (int i = 0; i < 8; ++i)
            r[i] = yuvChannel(ch, static_cast<double>(rawByte(w, 3 * i)), static_cast<double>(rawByte(w, 3 * i + 1)),
                              static_cast<double>(rawByte(w, 3 * i + 2)));
        bool nonZero;
        const float ss = storeResidual(r, nonZero);
        if (l8 == 0) {
            T.bc[slot] = static_cast<int>(3u * blk + static_cast<uint32_t>(ch));
            T.reps[slot] = 0;
            T.prev[slot] = 0;
            T.ngroups[slot] = 0;
            T.height[slot] = p.nBase;
            T.best[slot] = 0ULL;
            T.approxMax[slot] = 0u;
            T.twoEps[slot] = epsOf(ss);
            T.zeroRes[slot] = nonZero ? 0 : 1;
            T.dcFirst[0][slot] = T.dcFirst[1][slot] = 0xFFFFu;
        }
    };
    auto idleSlot = [&]() {
        const float z[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
        bool nonZero;
        storeResidual(z, nonZero);
        if (l8 == 0) {
            T.bc[slot] = -1;
            T.ngroups[slot] = 0;
            T.best[slot] = 0ULL;
            T.approxMax[slot] = 0u;
            T.twoEps[slot] = 0.0f;
            T.zeroRes[slot] = 1;
            T.dcFirst[0][slot] = T.dcFirst[1][slot] = 0xFFFFu;
        }
    };
    // next block index of this team's channel (shared-memory FIFO kept full by the TMA warp)
    auto popBlock = [&]() {
        uint32_t blk = 0;
        if (l8 == 0) blk = T.fifo[atomicAdd(&T.fifoHead, 1u) & (kFifo - 1)];
        return __shfl_sync(groupMask, blk, leaderLane);
    };
    auto pushCandidate = [&](int sl, int idx, uint32_t rowId) {
        const uint32_t pos = atomicAdd(&T.nCand, 1u);
        if (pos < kCandCapT) {
            T.cand[pos] = make_uint2(static_cast<uint32_t>(sl) | (static_cast<uint32_t>(idx) << 8), rowId);
        } else {  // list full (near-tie floods): re-score on the spot
            const float a = exactDot(p.rows32 + static_cast<size_t>(rowId) * kN, T.R32 + sl * kRStr);
            atomicMax(&T.best[sl], makeKey(a, idx));
        }
    };
    auto teamBarrier = [&]() { namedBarrier(1 + team, kTeamThreads); };
    auto consumerBarrier = [&]() { namedBarrier(3 + team, kTeamConsumers); };

    uint32_t trip = 0, normalTrips = 0, dcLoads = 0, pairBase = 0;
    bool dcLoadPending = false, dictReady = false, exhausted = false;
    // optional phase timing (MNENC_MP_TIMING=1): cycles of consumer thread 0 of CTA 0 per phase
    long long tPhase[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    long long tLast = 0;
    unsigned long long nItemsSum = 0, nCandSum = 0;
    unsigned long long nFlagSum = 0;
    long long mmaIssue = 0, mmaDone = 0;  // timing only: cycles of the two tmemLoad32 of phase (a)
    const bool timing = p.debugTiming && tid == 0 && blockIdx.x == 0;
#define MNENC_TICK(k)                     \
    if (timing) {                         \
        const long long now_ = clock64(); \
        tPhase[k] += now_ - tLast;        \
        tLast = now_;                     \
    }
    if (timing) tLast = clock64();

    // FIFO top-up by the whole TMA warp: entries [tail, head + kFifo) <- next block indices of channel `ch`
    auto topUpFifo = [&]() {
        uint32_t tail = 0, n = 0, first = 0;
        if (lane == 0) {
            tail = T.fifoTail;
            n = *reinterpret_cast<volatile uint32_t*>(&T.fifoHead) + kFifo - tail;
            first = nblocks;
            if (n && !exhausted) first = atomicAdd(&p.workCounter[ch], n);
            T.fifoTail = tail + n;
        }
        tail = __shfl_sync(kFullMask, tail, 0);
        n = __shfl_sync(kFullMask, n, 0);
        first = __shfl_sync(kFullMask, first, 0);
        if (n && (first >= nblocks || first + n >= nblocks)) exhausted = true;  // the channel's counter has run past the last block
        for (uint32_t i = lane; i < n; i += 32) T.fifo[(tail + i) & (kFifo - 1)] = (first < nblocks && first + i < nblocks) ? first + i : kNone;
    };

    while (true) {
        // ---- team barrier: residual images, slot state and the item lists of this trip are complete ----
        fenceBefore();
        teamBarrier();
        fenceAfter();
        const uint32_t cur = trip & 1u;
        MNENC_TICK(0)
        if (T.nActive[cur] == 0) {
            // ===== channel exhausted for this team (or kernel start): next channel, refill every slot =====
            ++ch;
            if (ch == 3) break;
            exhausted = false;
            if (isTma) {
                if (lane == 0) {
                    if (dcLoads) mbarWait(&T.mbarDcLoad, (dcLoads - 1u) & 1u);  // never two tile loads in flight
                    mbarExpectTx(&T.mbarDcLoad, 2 * kImgBytes);
                    const unsigned char* g = reinterpret_cast<const unsigned char*>(p.detUm16);
                    bulkCopy(s.dc[team], g + (static_cast<size_t>(ch) * p.nBase) * kImgBytes, kImgBytes, &T.mbarDcLoad);
                    bulkCopy(s.dc[team] + kImgBytes, g + (static_cast<size_t>(ch) * p.nBase + p.nBase - 1) * kImgBytes, kImgBytes, &T.mbarDcLoad);
                    T.fifoTail = T.fifoHead;  // whatever is left holds only kNone
                    T.nActive[cur ^ 1u] = 0;
                    T.nItems[cur ^ 1u] = 0;
                    T.nItemsHi[cur ^ 1u] = 0;
                }
                __syncwarp();
                topUpFifo();
            }
            ++dcLoads;
            dcLoadPending = true;
            teamBarrier();
            if (isConsumer) {
                const uint32_t blk = popBlock();
                if (blk != kNone) {
                    uint32_t w[6];
                    loadPixels(blk, w);
                    startBlock(blk, w);
                    if (l8 == 0) atomicAdd(&T.nActive[cur ^ 1u], 1u);
                } else {
                    idleSlot();
                }
                nextBlk = blk != kNone ? popBlock() : kNone;
                if (nextBlk != kNone) loadPixels(nextBlk, pix);
                fenceProxyAsync();
            }
            ++trip;
            continue;
        }
        // Streamed (slot, group) items are paired within one half of the slots, so that a pair's MMA needs only N = 16 columns
        // and six accumulators fit where three did: TMA and MMA run further ahead of the consumers.
        const uint32_t nLo = T.nItems[cur], nHi = T.nItemsHi[cur];
        const uint32_t nPLo = (nLo + 1u) >> 1, nPairs = nPLo + ((nHi + 1u) >> 1);
        if (timing) nItemsSum += nLo + nHi;
        auto pairCount = [&](uint32_t q) { return q < nPLo ? min(2u, nLo - 2u * q) : min(2u, nHi - 2u * (q - nPLo)); };
        auto pairItem = [&](uint32_t q, uint32_t i) -> uint32_t {
            return q < nPLo ? T.items[cur][2u * q + i] : T.items[cur][kItemCap - 1 - (2u * (q - nPLo) + i)];
        };

        if (isMma) {
            // ================= MMA issuer: the whole warp runs the issue path with warp-uniform operands, one elected lane
            // issues (see ummaF16Elect) =================
            const int nH = p.residualLo ? 2 : 1;  // residual = fp16 hi (+ fp16 lo): one or two accumulating passes
            {
                if (!dictReady) mbarWait(&s.mbarDict, 0), dictReady = true;
                const uint32_t dAddr = warpUniform(smemAddr(s.D16));
                const uint32_t rA = warpUniform(rAddr);
                const uint32_t tm = warpUniform(tmemT);
                // descriptors differ only in the 16-byte-granular start address (low 14 bits, no carry: shared memory < 256 KB)
                const unsigned long long descD = umDesc(dAddr, kLbo, kSbo), descR = umDesc(rA, kLbo, kSbo);
                const bool mmaTiming = p.debugTiming && blockIdx.x == 0 && team == 0;
                const long long tm0 = mmaTiming ? clock64() : 0;
                // one commit per pair of tiles: warp set 0 (tiles 0, 1) starts its epilogue while tiles 2, 3 are still being issued
#pragma unroll
                for (int t = 0; t < 4; ++t) {
                    for (int h = 0; h < nH; ++h)
#pragma unroll
                        for (int ks = 0; ks < 4; ++ks)
                            ummaF16Elect(tm + t * 32, descD + ((t * 16 * kSbo + ks * kKStep) >> 4),
                                         descR + ((h * (kSlots * kN * 2) + ks * kKStep) >> 4), kIdescF16M128N32, (h | ks) ? 1u : 0u);
                    if (t == 1) ummaCommitElect(&T.mbarBase);
                    if (t == 3) ummaCommitElect(&T.mbarBase2);
                }
                {   // resident tile: KLT groups of atoms 0 and nBase-1 of the team's channel against every slot
                    if (dcLoadPending) mbarWait(&T.mbarDcLoad, (dcLoads - 1u) & 1u), dcLoadPending = false;
                    const unsigned long long descC = umDesc(warpUniform(smemAddr(s.dc[team])), kLbo, kSbo);
                    for (int h = 0; h < nH; ++h)
#pragma unroll
                        for (int ks = 0; ks < 4; ++ks)
                            ummaF16Elect(tm + kColDc, descC + ((ks * kKStep) >> 4),
                                         descR + ((h * (kSlots * kN * 2) + ks * kKStep) >> 4), kIdescF16M128N32, (h | ks) ? 1u : 0u);
                }
                const long long tm1 = mmaTiming ? clock64() : 0;
                ummaCommitElect(&T.mbarDc);
                if (mmaTiming) {  // timing only: how long the issue takes, and when the scores are complete
                    mbarWait(&T.mbarDc, normalTrips & 1u);
                    mmaIssue += tm1 - tm0;
                    mmaDone += clock64() - tm0;
                }
                for (uint32_t q = 0; q < nPairs; ++q) {
                    const uint32_t ctr = pairBase + q;
                    const uint32_t stage = ctr % kTStages, su = ctr / kTStages, acc = ctr % kTAccs, au = ctr / kTAccs;
                    mbarWait(&T.full[stage], su & 1u);
                    mbarWait(&T.accEmpty[acc], (au & 1u) ^ 1u);
                    fenceAfter();
                    const unsigned long long descA = umDesc(warpUniform(smemAddr(T.ring[stage])), kLbo, kSbo);
                    const unsigned long long descRh = descR + (q < nPLo ? 0u : (2u * kSbo) >> 4);  // residual rows of slots 16-31
                    for (int h = 0; h < nH; ++h)
#pragma unroll
                        for (int ks = 0; ks < 4; ++ks)
                            ummaF16Elect(tm + kColAcc + acc * 16, descA + ((ks * kKStep) >> 4),
                                         descRh + ((h * (kSlots * kN * 2) + ks * kKStep) >> 4), kIdescF16M128N16, (h | ks) ? 1u : 0u);
                    ummaCommitElect(&T.empty[stage]);   // ring stage free once these MMAs have read it
                    ummaCommitElect(&T.accFull[acc]);   // scores of the pair are in TMEM
                }
            }
        } else if (isTma) {
            // ================= TMA producer: streamed group images, then the block-index FIFO. The whole warp runs the
            // issue path (warp-uniform operands, one elected lane issues): with two ring stages the issue latency of a pair
            // adds directly to the time per pair =================
            for (uint32_t q = 0; q < nPairs; ++q) {
                const uint32_t ctr = pairBase + q;
                const uint32_t stage = ctr % kTStages, su = ctr / kTStages;
                mbarWait(&T.empty[stage], (su & 1u) ^ 1u);
                const uint32_t n = pairCount(q);
                mbarExpectTxElect(&T.full[stage], n * kImgBytes);
                for (uint32_t i = 0; i < n; ++i) {
                    const uint32_t it = pairItem(q, i);
                    const uint32_t atom = T.groupAtom[(it & 31u) * kMaxK + (it >> 5)];
                    bulkCopyElect(T.ring[stage] + i * kImgBytes,
                                  reinterpret_cast<const unsigned char*>(p.detUm16) + (static_cast<size_t>(ch) * p.nBase + atom) * kImgBytes, kImgBytes,
                                  &T.full[stage]);
                }
            }
            __syncwarp();
            topUpFifo();
        } else {
            // ================= consumers =================
            if (tidT == 0) {
                T.nItems[cur ^ 1u] = 0;
                T.nItemsHi[cur ^ 1u] = 0;
                T.nActive[cur ^ 1u] = 0;
                T.nCand = 0;
            }
            const unsigned liveMask = __ballot_sync(kFullMask, T.bc[lane] >= 0 && !T.zeroRes[lane]);
            // ---- (a) base scores: per-slot maximum over this warp's two tiles ----
            mbarWait(wset ? &T.mbarBase2 : &T.mbarBase, normalTrips & 1u);
            fenceAfter();
            MNENC_TICK(1)
            float lm0 = 0.0f, lm1 = 0.0f, lmDc = 0.0f;  // this warp's LOCAL maxima (lane L: slot L) of its two tiles / the resident tile
#pragma unroll 1
            for (int tt = 0; tt < 2; ++tt) {
                float v[32];
                {
                    uint32_t raw[32];
                    tmemLoad32(tmemT + laneBase + static_cast<uint32_t>((wset * 2 + tt) * 32), raw);
#pragma unroll
                    for (int i = 0; i < 32; ++i) v[i] = fabsf(__uint_as_float(raw[i]));
                }
                // transposed butterfly: lane L ends with the maximum of slot L over the warp's 32 atoms
                float w16[16], w8[8], w4[4], w2[2];
                const bool b4 = lane & 16, b3 = lane & 8, b2 = lane & 4, b1 = lane & 2, b0 = lane & 1;
#pragma unroll
                for (int i = 0; i < 16; ++i) w16[i] = fmaxf(b4 ? v[i + 16] : v[i], __shfl_xor_sync(kFullMask, b4 ? v[i] : v[i + 16], 16));
#pragma unroll
                for (int i = 0; i < 8; ++i) w8[i] = fmaxf(b3 ? w16[i + 8] : w16[i], __shfl_xor_sync(kFullMask, b3 ? w16[i] : w16[i + 8], 8));
#pragma unroll
                for (int i = 0; i < 4; ++i) w4[i] = fmaxf(b2 ? w8[i + 4] : w8[i], __shfl_xor_sync(kFullMask, b2 ? w8[i] : w8[i + 4], 4));
#pragma unroll
                for (int i = 0; i < 2; ++i) w2[i] = fmaxf(b1 ? w4[i + 2] : w4[i], __shfl_xor_sync(kFullMask, b1 ? w4[i] : w4[i + 2], 2));
                const float m = fmaxf(b0 ? w2[1] : w2[0], __shfl_xor_sync(kFullMask, b0 ? w2[0] : w2[1], 1));
                atomicMax(&T.approxMax[lane], __float_as_uint(m));
                if (tt == 0) lm0 = m;
                else lm1 = m;
            }
            // resident tile: this warp's 32 rows (quarter) x 16 slots (warp set); rows 0-63 atom 0, 64-127 atom nBase-1
            const int dcK = quarter >> 1, dcRow = (quarter & 1) * 32 + lane;
            const int dcAtom = dcK ? p.nBase - 1 : 0;
            const bool dcRowOk = dcRow < s.detRows[dcAtom];
            const unsigned dcHas = __ballot_sync(kFullMask, lane < 16 && T.dcFirst[dcK][wset * 16 + (lane & 15)] != 0xFFFFu) & 0xFFFFu &
                                   (liveMask >> (wset * 16));
            mbarWait(&T.mbarDc, normalTrips & 1u);  // every warp: the next trip's MMAs may only start once all have seen this one's
            fenceAfter();
            if (dcHas) {
                uint32_t raw[16];
                tmemLoad16(tmemT + laneBase + kColDc + static_cast<uint32_t>(wset * 16), raw);
                float u[16], w8[8], w4[4], w2[2];
                const bool b3 = lane & 8, b2 = lane & 4, b1 = lane & 2, b0 = lane & 1;
#pragma unroll
                for (int i = 0; i < 16; ++i) {
                    const float a = (dcRowOk && ((dcHas >> i) & 1u)) ? fabsf(__uint_as_float(raw[i])) : 0.0f;
                    u[i] = fmaxf(a, __shfl_xor_sync(kFullMask, a, 16));
                }
#pragma unroll
                for (int i = 0; i < 8; ++i) w8[i] = fmaxf(b3 ? u[i + 8] : u[i], __shfl_xor_sync(kFullMask, b3 ? u[i] : u[i + 8], 8));
#pragma unroll
                for (int i = 0; i < 4; ++i) w4[i] = fmaxf(b2 ? w8[i + 4] : w8[i], __shfl_xor_sync(kFullMask, b2 ? w8[i] : w8[i + 4], 4));
#pragma unroll
                for (int i = 0; i < 2; ++i) w2[i] = fmaxf(b1 ? w4[i + 2] : w4[i], __shfl_xor_sync(kFullMask, b1 ? w4[i] : w4[i + 2], 2));
                const float m = fmaxf(b0 ? w2[1] : w2[0], __shfl_xor_sync(kFullMask, b0 ? w2[0] : w2[1], 1));
                if (lane < 16 && ((dcHas >> lane) & 1u)) atomicMax(&T.approxMax[wset * 16 + lane], __float_as_uint(m)), lmDc = m;
            }
            consumerBarrier();
            MNENC_TICK(2)
            // ---- (b) rows that can still be the exact maximum -> candidate list. A slot can only have such a row in this
            // warp's part of a tile if the warp's LOCAL maximum of that slot is within 2 eps of the slot's global maximum:
            // that is true for ~3 of the 32 slots, so only those columns are read from TMEM again. ----
            auto filterColumns = [&](unsigned need, uint32_t colBase, int slBase, auto&& push) {
                while (need) {
                    const int i0 = __ffs(need) - 1;
                    need &= need - 1;
                    const int i1 = need ? __ffs(need) - 1 : -1;
                    if (i1 >= 0) need &= need - 1;
                    const uint32_t r0 = tmemLoad1NoWait(tmemT + laneBase + colBase + static_cast<uint32_t>(i0));
                    const uint32_t r1 = i1 >= 0 ? tmemLoad1NoWait(tmemT + laneBase + colBase + static_cast<uint32_t>(i1)) : 0u;
                    tmemWaitLd();
                    const int s0 = slBase + i0, s1 = slBase + i1;
                    if (fabsf(__uint_as_float(r0)) + T.twoEps[s0] >= __uint_as_float(T.approxMax[s0])) push(s0);
                    if (i1 >= 0 && fabsf(__uint_as_float(r1)) + T.twoEps[s1] >= __uint_as_float(T.approxMax[s1])) push(s1);
                }
            };
#pragma unroll 1
            for (int tt = 0; tt < 2; ++tt) {
                const int tile = wset * 2 + tt;
                const int atom = tile * 128 + quarter * 32 + lane;
                const float lm = tt ? lm1 : lm0;
                const unsigned need = __ballot_sync(kFullMask, lm + T.twoEps[lane] >= __uint_as_float(T.approxMax[lane])) & liveMask;
                if (timing) nFlagSum += __popc(need);
                filterColumns(need, static_cast<uint32_t>(tile * 32), 0, [&](int sl) {
                    if (atom < p.nBase) pushCandidate(sl, atom, static_cast<uint32_t>(atom));
                });
            }
            if (dcHas) {  // rows of the resident groups
                const int sl = wset * 16 + (lane & 15);
                const unsigned need = __ballot_sync(kFullMask, lane < 16 && lmDc + T.twoEps[sl] >= __uint_as_float(T.approxMax[sl])) & dcHas;
                filterColumns(need, kColDc + static_cast<uint32_t>(wset * 16), wset * 16, [&](int sl2) {
                    if (dcRowOk)
                        pushCandidate(sl2, static_cast<int>(T.dcFirst[dcK][sl2]) + dcRow, static_cast<uint32_t>(kRowsPad + (ch * p.nBase + dcAtom) * 64 + dcRow));
                });
            }
            fenceBefore();
            MNENC_TICK(3)
            MNENC_TICK(4)
            // One (slot, group) item whose 64 scores sit in TMEM column `col`: lanes of this warp hold rows rowBase + lane.
            auto scoreItem = [&](uint32_t it, uint32_t col, int rowBase) {
                const int sl = static_cast<int>(it & 31u), gi = static_cast<int>(it >> 5);
                const uint32_t raw = tmemLoad1(tmemT + laneBase + col + static_cast<uint32_t>(sl & 15));
                const int atom = T.groupAtom[sl * kMaxK + gi];
                const int first = T.groupFirst[sl * kMaxK + gi];
                const int row = rowBase + lane;
                const bool rowOk = row < s.detRows[atom];
                const float a = rowOk ? fabsf(__uint_as_float(raw)) : 0.0f;
                const uint32_t wmax = __reduce_max_sync(kFullMask, __float_as_uint(a));
                uint32_t seen = 0;
                if (lane == 0) seen = max(atomicMax(&T.approxMax[sl], wmax), wmax);
                seen = __shfl_sync(kFullMask, seen, 0);
                if (rowOk && ((liveMask >> sl) & 1u) && a + T.twoEps[sl] >= __uint_as_float(seen))
                    pushCandidate(sl, first + row, static_cast<uint32_t>(kRowsPad + (ch * p.nBase + atom) * 64 + row));
            };
            // ---- (c) streamed pairs: warp set w takes pairs w, w+2, ...; quarters 0,1 -> first item, 2,3 -> second.
            // (Batching several pairs per warp for ILP was measured slower: the phase is bound by the TMA ring's two stages,
            // and a batch cannot start before its last pair has arrived.) ----
            for (uint32_t q = static_cast<uint32_t>(wset); q < nPairs; q += 2) {
                const uint32_t ctr = pairBase + q;
                const uint32_t acc = ctr % kTAccs, au = ctr / kTAccs;
                const uint32_t itemSel = static_cast<uint32_t>(quarter >> 1);
                mbarWait(&T.accFull[acc], au & 1u);
                fenceAfter();
                if (itemSel < pairCount(q)) scoreItem(pairItem(q, itemSel), kColAcc + acc * 16, (quarter & 1) * 32);
                fenceBefore();
                __syncwarp();
                if (lane == 0) mbarArrive(&T.accEmpty[acc]);
            }
            consumerBarrier();
            MNENC_TICK(5)
            // ---- (d) exact re-score of the listed candidates, one per thread ----
            {
                const uint32_t n = min(T.nCand, static_cast<uint32_t>(kCandCapT));
                if (timing) nCandSum += T.nCand;
                for (uint32_t i = tidT; i < n; i += kTeamConsumers) {
                    const uint2 c = T.cand[i];
                    const int sl = static_cast<int>(c.x & 31u);
                    const float a = exactDot(p.rows32 + static_cast<size_t>(c.y) * kN, T.R32 + sl * kRStr);
                    atomicMax(&T.best[sl], makeKey(a, static_cast<int>(c.x >> 8)));
                }
            }
            consumerBarrier();
            MNENC_TICK(6)
            // ================= step C: select, quantise, update (MatchingPursuit.cpp:86-104), 8 threads per slot =================
            const int bc = T.bc[slot];
            if (bc >= 0) {
                const unsigned long long key = T.zeroRes[slot] ? makeKey(0.0f, 0) : T.best[slot];  // r == 0: every score is +0, atom 0 wins
                const int idx = keyIndex(key);
                const float proj = keyValue(key);
                const int reps = T.reps[slot];
                const int prev = T.prev[slot];
                const int ng = T.ngroups[slot];
                const int height = T.height[slot];
                const float qf = s.quant[ch * kMaxK + reps];
                const int qi = static_cast<int>(roundf(__fdiv_rn(proj, qf)));
                const uint32_t deltaId = (reps ? zigzag(idx - prev) : static_cast<uint32_t>(idx)) & 0xFFFFu;
                statRows += (l8 == 0) ? height : 0;
                int count = -1;
                __syncwarp(groupMask);
                if (qi == 0) {
                    count = reps;
                } else {
                    const float c = __fmul_rn(qf, static_cast<float>(qi));
                    uint32_t rowId = static_cast<uint32_t>(idx);
                    if (idx >= p.nBase) {
                        int gi = ng - 1;
                        while (gi > 0 && T.groupFirst[slot * kMaxK + gi] > idx) --gi;
                        const int atom = T.groupAtom[slot * kMaxK + gi];
                        rowId = static_cast<uint32_t>(kRowsPad + (ch * p.nBase + atom) * 64 + (idx - T.groupFirst[slot * kMaxK + gi]));
                    }
                    const float4* rowp = reinterpret_cast<const float4*>(p.rows32 + static_cast<size_t>(rowId) * kN) + 2 * l8;
                    const float4 d0 = __ldg(rowp), d1 = __ldg(rowp + 1);
                    const float4 r0 = reinterpret_cast<const float4*>(myR32)[0], r1 = reinterpret_cast<const float4*>(myR32)[1];
                    float r[8];
                    r[0] = __fsub_rn(r0.x, __fmul_rn(c, d0.x));
                    r[1] = __fsub_rn(r0.y, __fmul_rn(c, d0.y));
                    r[2] = __fsub_rn(r0.z, __fmul_rn(c, d0.z));
                    r[3] = __fsub_rn(r0.w, __fmul_rn(c, d0.w));
                    r[4] = __fsub_rn(r1.x, __fmul_rn(c, d1.x));
                    r[5] = __fsub_rn(r1.y, __fmul_rn(c, d1.y));
                    r[6] = __fsub_rn(r1.z, __fmul_rn(c, d1.z));
                    r[7] = __fsub_rn(r1.w, __fmul_rn(c, d1.w));
                    bool nonZero;
                    const float ss = storeResidual(r, nonZero);
                    const int ngNew = ng + (idx < p.nBase ? 1 : 0);
                    if (l8 == 0) {
                        p.pairs[static_cast<size_t>(bc) * p.K + reps] = deltaId | ((zigzag(qi) & 0xFFFFu) << 16);
                        T.prev[slot] = idx;
                        T.reps[slot] = reps + 1;
                        T.best[slot] = 0ULL;
                        T.approxMax[slot] = 0u;
                        T.twoEps[slot] = epsOf(ss);
                        T.zeroRes[slot] = nonZero ? 0 : 1;
                        if (idx < p.nBase) {
                            if (idx == 0 && T.dcFirst[0][slot] == 0xFFFFu) T.dcFirst[0][slot] = static_cast<uint16_t>(height);
                            if (idx == p.nBase - 1 && T.dcFirst[1][slot] == 0xFFFFu) T.dcFirst[1][slot] = static_cast<uint16_t>(height);
                            T.groupAtom[slot * kMaxK + ng] = static_cast<uint16_t>(idx);
                            T.groupFirst[slot * kMaxK + ng] = static_cast<uint16_t>(height);
                            T.ngroups[slot] = ngNew;
                            T.height[slot] = height + s.detRows[idx];
                        }
                    }
                    if (reps + 1 == p.K) {
                        count = p.K;
                    } else if (ngNew > 0) {  // this slot's streamed (slot, group) items of the next trip (resident groups need no list)
                        __syncwarp(groupMask);
                        for (int g0 = 0; g0 < ngNew; g0 += 8) {
                            const int gi = g0 + l8;
                            const bool valid = gi < ngNew;
                            const int atom = valid ? T.groupAtom[slot * kMaxK + gi] : -1;
                            const bool isDc = valid && (atom == 0 || atom == p.nBase - 1);
                            const unsigned mSt = (__ballot_sync(groupMask, valid && !isDc) & groupMask) >> leaderLane;
                            uint32_t atSt = 0;
                            const bool hiHalf = slot >= 16;
                            if (l8 == 0 && mSt) atSt = atomicAdd(hiHalf ? &T.nItemsHi[cur ^ 1u] : &T.nItems[cur ^ 1u], static_cast<uint32_t>(__popc(mSt)));
                            atSt = __shfl_sync(groupMask, atSt, leaderLane);
                            const unsigned below = (1u << l8) - 1u;
                            const uint32_t at = atSt + __popc(mSt & below);
                            if (valid && !isDc) T.items[cur ^ 1u][hiHalf ? kItemCap - 1 - at : at] = static_cast<uint16_t>(slot | (gi << 5));
                        }
                    }
                }
                int stillActive = 1;
                if (count >= 0) {
                    if (l8 == 0) p.counts[bc] = static_cast<uint16_t>(count);
                    for (int i = count + l8; i < p.K; i += 8) p.pairs[static_cast<size_t>(bc) * p.K + i] = 0u;
                    statPicks += (l8 == 0) ? count : 0;
                    __syncwarp(groupMask);
                    if (nextBlk != kNone) {  // pixels were prefetched while the previous block was being coded
                        startBlock(nextBlk, pix);
                        nextBlk = popBlock();
                        if (nextBlk != kNone) loadPixels(nextBlk, pix);
                    } else {
                        idleSlot();
                        stillActive = 0;
                    }
                }
                if (stillActive && l8 == 0) atomicAdd(&T.nActive[cur ^ 1u], 1u);
            }
            fenceProxyAsync();
            MNENC_TICK(7)
        }
        pairBase += nPairs;
        ++trip;
        ++normalTrips;
    }
    if (isConsumer) {
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) {
            statRows += __shfl_xor_sync(kFullMask, statRows, off);
            statPicks += __shfl_xor_sync(kFullMask, statPicks, off);
        }
        if (lane == 0) {
            atomicAdd(&p.stats[0], statRows);
            atomicAdd(&p.stats[1], statPicks);
        }
    }
    if (p.debugTiming && blockIdx.x == 0 && team == 0 && isMma && lane == 0 && normalTrips)
        printf("mp teams timing, MMA thread: base + resident MMAs per trip: %lld cycles to issue, %lld cycles until complete\n",
               mmaIssue / normalTrips, mmaDone / normalTrips);
    if (timing && normalTrips)
        printf("mp teams timing, CTA 0 team 0: %u trips; cycles per trip: barrier %lld, baseMMA wait %lld, max+bar %lld, base filter %lld, "
               "resident %lld, streamed+bar %lld, cand+bar %lld, stepC %lld; per trip: streamed items %.1f, flagged columns of warp 0 in (b) %.1f, candidates %.1f\n",
               normalTrips, tPhase[0] / normalTrips, tPhase[1] / normalTrips, tPhase[2] / normalTrips, tPhase[3] / normalTrips,
               tPhase[4] / normalTrips, tPhase[5] / normalTrips, tPhase[6] / normalTrips, tPhase[7] / normalTrips,
               double(nItemsSum) / normalTrips, double(nFlagSum) / normalTrips, double(nCandSum) / normalTrips);
#undef MNENC_TICK
    fenceBefore();
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(s.tmemBase), "r"(512u) : "memory");
}

}  // namespace

cudaError_t launchMatchingPursuitTeams(const MpParams& p, int numSMs, cudaStream_t stream) {
    static bool configured = false;
    const size_t smem = sizeof(SmemV4) + 1024;
    if (!configured) {
        cudaError_t e = cudaFuncSetAttribute(matchingPursuitTeamsKernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
        if (e != cudaSuccess) return e;
        configured = true;
    }
    if (p.nbc == 0) return cudaSuccess;
    const unsigned int nblocks = p.nbc / 3;
    unsigned int grid = (nblocks + kTeams * kSlots - 1) / (kTeams * kSlots);
    if (grid > static_cast<unsigned int>(numSMs)) grid = numSMs;
    matchingPursuitTeamsKernel<<<grid, kThreadsV4, smem, stream>>>(p);
    return cudaGetLastError();
}

}  // namespace mnenc
