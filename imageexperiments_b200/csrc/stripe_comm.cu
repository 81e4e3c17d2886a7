// ONE image sharded by block columns over the GPUs of one box, exchanges INSIDE the library (BASELINE.json
// configs[3]): the six phases of the stripe protocol (stripe_api.cu) driven in C++ with NCCL over NVLink / NVSwitch between
// them, so that the reference-side callers — the C++ shim and the `Compression` front end (Compression.cpp:117-123) — get
// stripes from a plain encode call. Two ways to get a communicator:
//   * one process per GPU (torchrun, MPI, ...): mnenc_comm_unique_id on one rank, the 128 bytes handed to the others by any
//     means, mnenc_comm_init_rank on every rank, then mnenc_encode_striped_rgb8 collectively;
//   * one process, several GPUs: mnenc_create_multi (ncclCommInitAll); mnenc_encode_rgb8 on that context runs one host thread
//     per GPU through the same code.
// Exchanges (north_star: "NCCL ... only for the global symbol-histogram allreduce and the offset scan and gather of per-GPU
// bitstream segments"):
//   all-gather  stream sizes + last raw DC value; boundary run records                      (two small host-staged gathers)
//   histogram   each rank compacts its local histogram into (sequence, symbol, count, first position) entries; the entry
//               lists are all-gathered and merged on every GPU (sum / min / max) - ~1 MB instead of three dense
//               [193][65536] all-reduces (150 MB)
//   code tables the 193 table builds are dealt over the ranks (largest alphabets first), every rank builds its share on
//               its host threads, the tables are all-gathered
//   offset scan all-gather of the payload bits per (rank, sequence)
//   segments    reduce(SUM) of the packed words to rank 0: every rank packs its slices at their FINAL bit positions, the bits
//               are disjoint, so the sum is the concatenation of the per-GPU bitstream segments
#include <algorithm>
#include <condition_variable>
#include <cstring>
#include <mutex>
#include <numeric>
#include <thread>

#include "mnenc_ctx.cuh"
#include "nccl_dyn.h"

namespace mnenc {

// One process, several GPUs: the host threads of the group meet here right before every NCCL launch, after the buffers of that
// exchange have been allocated. Without it thread A's NCCL kernel (spinning until its peer's kernel starts) and thread B's
// cudaHostAlloc / cudaMalloc on the way to that launch (which wait for every running kernel of the process) wait for each other.
// A rank that fails aborts the meeting point, so its peers return an error instead of launching a collective nobody completes.
struct GroupSync {
    std::mutex m;
    std::condition_variable cv;
    int n = 1, waiting = 0;
    uint64_t generation = 0;
    bool aborted = false;
    bool arrive() {
        std::unique_lock<std::mutex> lk(m);
        if (aborted) return false;
        const uint64_t g0 = generation;
        if (++waiting == n) {
            waiting = 0;
            ++generation;
            cv.notify_all();
            return true;
        }
        cv.wait(lk, [&] { return generation != g0 || aborted; });
        return generation != g0;
    }
    void abort() {
        std::lock_guard<std::mutex> lk(m);
        aborted = true;
        cv.notify_all();
    }
    void reset() {
        std::lock_guard<std::mutex> lk(m);
        aborted = false;
        waiting = 0;
    }
};

struct StripeComm {
    std::shared_ptr<GroupSync> sync;  // mnenc_create_multi only
    ncclComm_t comm = nullptr;
    int rank = 0, nranks = 1;
    DevBuf stage, gathered;   // host-staged all-gathers; all-gathered histogram entries
    PinBuf hStage;
    float ms[16] = {};         // host clock per phase / exchange of the last striped encode
};

void destroyStripeComm(StripeComm* sc) {
    if (!sc) return;
    if (sc->comm) {
        std::string err;
        if (const NcclApi* n = ncclApi(&err)) n->commDestroy(sc->comm);
    }
    delete sc;
}

int stripeRanks(const mnenc_ctx* c) { return c->comm ? c->comm->nranks : 1; }

namespace {

#define NC(expr)                                                                                                       \
    do {                                                                                                               \
        ncclResult_t r_ = (expr);                                                                                      \
        if (r_ != ncclSuccess) return ::mnenc::fail(MNENC_ERR_CUDA, std::string(#expr) + ": " + nccl->getErrorString(r_)); \
    } while (0)

#define MEET(sc)                                                                                                  \
    do {                                                                                                          \
        if ((sc)->sync && !(sc)->sync->arrive()) return ::mnenc::fail(MNENC_ERR_STATE, "another GPU of the group failed"); \
    } while (0)

// Failures are collective. A rank whose local work failed must not simply return: its peers would sit in the next NCCL
// collective forever (there is no time-out, and across processes nobody can abort them). So every host-staged gather carries,
// in a 16-byte header in front of the payload, the result code of the sender's local work since the previous exchange; after
// the gather all ranks see all codes and, if any is non-zero, ALL of them return an error from this same point - the failing
// rank its own code and text, the others MNENC_ERR_STATE naming the failed rank. What this cannot cover: a sticky CUDA error
// (the rank can no longer launch anything, the collective included) and a rank that dies; both need the process group torn down.
constexpr size_t kStatusBytes = 16;
constexpr size_t kStageSlot = size_t(1) << 20;  // reserved per rank at communicator creation, so that the gathers allocate nothing

int groupVerdict(const StripeComm* sc, const unsigned char* slots, size_t slotBytes, int localRc) {
    for (int r = 0; r < sc->nranks; ++r) {
        int32_t code = 0;
        std::memcpy(&code, slots + slotBytes * r, sizeof(code));
        if (code == 0) continue;
        if (localRc) return localRc;  // this rank failed too: keep its own error text
        return fail(MNENC_ERR_STATE, "rank " + std::to_string(r) + " of the group failed (code " + std::to_string(code) +
                                         "); this rank stopped with it before the next exchange");
    }
    return localRc;
}

// all-gather of `bytes` host bytes per rank (small records): H2D, ncclAllGather, D2H on the context's stream. localRc: see above;
// when it is non-zero the payload may be garbage (src may even be null), only the header matters.
int allGatherHost(mnenc_ctx* c, const NcclApi* nccl, const void* src, void* dst, size_t bytes, int localRc) {
    StripeComm* sc = c->comm.get();
    if (sc->nranks == 1) {
        if (!localRc && bytes) std::memcpy(dst, src, bytes);
        return localRc;
    }
    const size_t padded = kStatusBytes + (bytes + 15) / 16 * 16;
    // beyond the reservation only for very large code tables; a failure HERE still strands the peers (documented in DESIGN 7.2)
    CU(sc->stage.ensure(padded * (sc->nranks + 1)));
    CU(sc->hStage.ensure(padded * (sc->nranks + 1)));
    unsigned char* h = sc->hStage.as<unsigned char>();
    unsigned char* d = sc->stage.as<unsigned char>();
    const int32_t code = localRc;
    std::memset(h, 0, kStatusBytes);
    std::memcpy(h, &code, sizeof(code));
    if (!localRc && bytes) std::memcpy(h + kStatusBytes, src, bytes);
    MEET(sc);
    CU(cudaMemcpyAsync(d, h, padded, cudaMemcpyHostToDevice, c->stream));
    NC(nccl->allGather(d, d + padded, padded, ncclUint8, sc->comm, c->stream));
    CU(cudaMemcpyAsync(h + padded, d + padded, padded * sc->nranks, cudaMemcpyDeviceToHost, c->stream));
    CU(waitStream(c));
    if (int rc = groupVerdict(sc, h + padded, padded, localRc)) return rc;
    if (bytes)
        for (int r = 0; r < sc->nranks; ++r) std::memcpy(static_cast<unsigned char*>(dst) + r * bytes, h + padded * (r + 1) + kStatusBytes, bytes);
    return MNENC_OK;
}

// Before an exchange that carries no host-staged record of its own (the reduce of the segments): gather the codes alone.
int agreeStatus(mnenc_ctx* c, const NcclApi* nccl, int localRc) { return allGatherHost(c, nccl, nullptr, nullptr, 0, localRc); }

int reserveStage(StripeComm* sc) {
    CU(sc->stage.ensure(kStageSlot * (sc->nranks + 1)));
    CU(sc->hStage.ensure(kStageSlot * (sc->nranks + 1)));
    return MNENC_OK;
}

}  // namespace

// Sparse exchange of the symbol histogram: local dense tables -> compact entries -> all-gather -> merge into the dense
// tables of every rank (identical everywhere afterwards, like the three dense all-reduces it replaces).
int exchangeHistogramSparse(mnenc_ctx* c, BackendBuffers* bp, int nSeq, int localRc) {
    StripeComm* sc = c->comm.get();
    std::string err;
    const NcclApi* nccl = ncclApi(&err);
    if (!nccl) return fail(MNENC_ERR_CUDA, err);
    cudaStream_t st = c->stream;
    const int nr = sc->nranks;
    if (!localRc && !bp) localRc = fail(MNENC_ERR_STATE, "stripe protocol called out of order");
    if (!localRc) c->launches += launchCompact(*bp, nSeq, st);  // hist bins -> entries (bins cleared, firstPos = entry index)
    // {entries, result code of the local work} of every rank: the code rides along, so a rank whose histogram phase failed
    // still joins this gather (with 0 entries) and every rank stops after it, together (see allGatherHost)
    uint32_t* dMine = sc->stage.as<uint32_t>();       // reserved at communicator creation
    uint32_t* dAll = dMine + 4;
    uint32_t* hMine = sc->hStage.as<uint32_t>();
    uint32_t* hAll = hMine + 4;
    hMine[0] = 0, hMine[1] = static_cast<uint32_t>(localRc);
    CU(cudaMemcpyAsync(dMine, hMine, 8, cudaMemcpyHostToDevice, st));
    if (!localRc) CU(cudaMemcpyAsync(dMine, bp->nEntries, 4, cudaMemcpyDeviceToDevice, st));
    MEET(sc);
    NC(nccl->allGather(dMine, dAll, 2, ncclUint32, sc->comm, st));
    CU(cudaMemcpyAsync(hAll, dAll, 8 * nr, cudaMemcpyDeviceToHost, st));
    CU(waitStream(c));
    for (int r = 0; r < nr; ++r)
        if (hAll[2 * r + 1]) {
            if (localRc) return localRc;
            return fail(MNENC_ERR_STATE, "rank " + std::to_string(r) + " of the group failed (code " + std::to_string(static_cast<int32_t>(hAll[2 * r + 1])) +
                                             "); this rank stopped with it before the histogram exchange");
        }
    BackendBuffers& b = *bp;
    uint32_t nMax = 0;
    for (int r = 0; r < nr; ++r) nMax = std::max(nMax, hAll[2 * r]);
    // the same decision on every rank (all see the same counts), so nobody is left alone in the next gather
    if (nMax > c->entryCap) return fail(MNENC_ERR_INTERNAL, "more than 2^20 distinct (sequence, symbol) pairs on one rank");
    const uint32_t mine = hAll[2 * sc->rank];
    c->launches += launchResetHistogram(b, mine, st);  // firstPos back to "not seen", maxSym and nEntries to 0; the entry list stays
    if (nMax) {
        // grow-only; sized for the worst case the first time, so that this allocation cannot fail on one rank in a later image
        CU(sc->gathered.ensure(static_cast<size_t>(c->entryCap) * nr * sizeof(HistEntry)));
        MEET(sc);
        NC(nccl->allGather(b.entries, sc->gathered.p, static_cast<size_t>(nMax) * sizeof(HistEntry), ncclUint8, sc->comm, st));
        c->launches += launchMergeEntries(b, sc->gathered.as<HistEntry>(), dAll, 2, nr, nMax, st);
    }
    return MNENC_OK;
}

namespace {

// Serialised code table of one sequence, as exchanged between the ranks.
struct PlanWire {
    uint32_t seq, golombM;  // golombM: 0 = Huffman
    uint32_t eofCode, eofLen, maxCodeLen, nSym;
    uint64_t payloadBits, headerBits;
    uint32_t headerBytes, pad;
};

}  // namespace

// Which sequences rank `me` builds: largest alphabets first, dealt round-robin (the same list on every rank: the alphabet
// sizes are those of the merged histogram).
static std::vector<int> shareOf(const std::vector<std::vector<SymbolStat>>& stats, int me, int nr) {
    const int nSeq = static_cast<int>(stats.size());
    std::vector<int> byLoad(nSeq);
    std::iota(byLoad.begin(), byLoad.end(), 0);
    std::stable_sort(byLoad.begin(), byLoad.end(), [&](int a, int b2) { return stats[a].size() > stats[b2].size(); });
    std::vector<int> mine;
    for (int i = me; i < nSeq; i += nr) mine.push_back(byLoad[i]);
    return mine;
}

// This rank's share of the code tables: built into plans[q] and serialised for the peers.
void planShardBuild(const std::vector<std::vector<SymbolStat>>& stats, int me, int nr, HostPool* pool, std::vector<SequencePlan>& plans,
                    std::vector<unsigned char>& wire) {
    const std::vector<int> mineSeqs = shareOf(stats, me, nr);
    auto build = [&](int i) { planSequence(stats[mineSeqs[i]], plans[mineSeqs[i]]); };
    if (pool) pool->parallelFor(static_cast<int>(mineSeqs.size()), build);
    else
        for (int i = 0; i < static_cast<int>(mineSeqs.size()); ++i) build(i);
    wire.clear();
    for (int q : mineSeqs) {
        const SequencePlan& p = plans[q];
        PlanWire w{};
        w.seq = static_cast<uint32_t>(q);
        w.golombM = p.golomb ? p.golombM : 0u;
        w.eofCode = p.eofCode, w.eofLen = p.eofLen, w.maxCodeLen = p.maxCodeLen;
        w.nSym = p.golomb ? 0u : static_cast<uint32_t>(stats[q].size());
        w.payloadBits = p.payloadBits, w.headerBits = p.header.nbits;
        w.headerBytes = static_cast<uint32_t>(p.header.bytes.size());
        const size_t at = wire.size(), hb = (w.headerBytes + 7) / 8 * 8;
        wire.resize(at + sizeof(PlanWire) + hb + static_cast<size_t>(w.nSym) * 8);
        std::memcpy(wire.data() + at, &w, sizeof(w));
        std::memcpy(wire.data() + at + sizeof(w), p.header.bytes.data(), w.headerBytes);
        uint32_t* sym = reinterpret_cast<uint32_t*>(wire.data() + at + sizeof(w) + hb);
        for (uint32_t k = 0; k < w.nSym; ++k) {
            sym[2 * k] = stats[q][k].symbol | (static_cast<uint32_t>(p.len[k]) << 16);
            sym[2 * k + 1] = p.code[k];
        }
    }
}

// A peer's serialised tables -> plans[q] in THIS rank's entry order (the peers hold the same set of (sequence, symbol) entries
// after the histogram merge, possibly in another order: codes are matched by symbol).
int planShardApply(const std::vector<std::vector<SymbolStat>>& stats, const unsigned char* p0, size_t size, std::vector<SequencePlan>& plans) {
    const int nSeq = static_cast<int>(stats.size());
    std::vector<uint32_t> slot(65536);
    size_t at = 0;
    while (at + sizeof(PlanWire) <= size) {
        PlanWire w;
        std::memcpy(&w, p0 + at, sizeof(w));
        if (w.seq >= static_cast<uint32_t>(nSeq)) return fail(MNENC_ERR_INTERNAL, "bad code table from a peer");
        const size_t hb = (static_cast<size_t>(w.headerBytes) + 7) / 8 * 8;
        if (at + sizeof(PlanWire) + hb + static_cast<size_t>(w.nSym) * 8 > size) return fail(MNENC_ERR_INTERNAL, "truncated code table from a peer");
        SequencePlan& p = plans[w.seq];
        const std::vector<SymbolStat>& mineStats = stats[w.seq];
        p = SequencePlan{};
        p.golomb = w.golombM != 0;
        p.golombM = w.golombM;
        p.eofCode = w.eofCode, p.eofLen = static_cast<uint8_t>(w.eofLen), p.maxCodeLen = w.maxCodeLen;
        p.payloadBits = w.payloadBits;
        p.header.nbits = w.headerBits;
        p.header.bytes.assign(p0 + at + sizeof(w), p0 + at + sizeof(w) + w.headerBytes);
        p.code.assign(mineStats.size(), 0u);
        p.len.assign(mineStats.size(), 0);
        if (w.nSym) {
            if (w.nSym != mineStats.size()) return fail(MNENC_ERR_INTERNAL, "code table of a peer does not match the merged histogram");
            for (uint32_t k = 0; k < w.nSym; ++k) slot[mineStats[k].symbol & 0xFFFFu] = k;
            const uint32_t* sym = reinterpret_cast<const uint32_t*>(p0 + at + sizeof(w) + hb);
            for (uint32_t k = 0; k < w.nSym; ++k) {
                const uint32_t s = sym[2 * k] & 0xFFFFu, mk = slot[s];
                if (mk >= w.nSym || mineStats[mk].symbol != s) return fail(MNENC_ERR_INTERNAL, "code table of a peer names an unknown symbol");
                p.code[mk] = sym[2 * k + 1];
                p.len[mk] = static_cast<uint8_t>(sym[2 * k] >> 16);
            }
        }
        at += sizeof(PlanWire) + hb + static_cast<size_t>(w.nSym) * 8;
    }
    return MNENC_OK;
}

// Code tables of all sequences with the builds dealt over the ranks. stats[q] are this rank's entries of sequence q (same SET
// on every rank after the histogram merge, order may differ); on return plans[q].code/len follow this rank's order.
int planSharded(mnenc_ctx* c, const std::vector<std::vector<SymbolStat>>& stats, std::vector<SequencePlan>& plans, int localRc) {
    StripeComm* sc = c->comm.get();
    std::string err;
    const NcclApi* nccl = ncclApi(&err);
    if (!nccl) return fail(MNENC_ERR_CUDA, err);
    const int nr = sc->nranks, me = sc->rank;
    std::vector<unsigned char> wire;
    if (!localRc) {
        try {
            planShardBuild(stats, me, nr, c->pool.get(), plans, wire);
        } catch (...) {
            localRc = abiCaught();  // joins the gather below with its code; every rank stops there
        }
    }
    // sizes (with the result codes of the local work, see allGatherHost), then the tables
    std::vector<uint64_t> sizes(nr);
    const uint64_t mySize = wire.size();
    if (int rc = allGatherHost(c, nccl, &mySize, sizes.data(), 8, localRc)) return rc;
    const size_t maxSize = static_cast<size_t>(*std::max_element(sizes.begin(), sizes.end()));
    std::vector<unsigned char> all;
    int rcLocal = MNENC_OK;
    try {
        wire.resize(maxSize);
        all.resize(maxSize * nr);
    } catch (...) {
        rcLocal = abiCaught();
    }
    if (int rc = allGatherHost(c, nccl, wire.data(), all.data(), maxSize, rcLocal)) return rc;
    for (int r = 0; r < nr; ++r) {
        if (r == me) continue;
        if (int rc = planShardApply(stats, all.data() + maxSize * r, static_cast<size_t>(sizes[r]), plans)) return rc;
    }
    return MNENC_OK;
}

// reduce(SUM) of the packed words to rank 0, in up to four chunks: while chunk i + 1 is being reduced over NVLink, rank 0 already
// copies chunk i to the host on a second stream (the 56 MB stream of a 16384^2 image is 1 ms of PCIe time, as long as the rest of
// the tail together).
// Rank 0's fallible set-up for reduceSegments (second stream, events, the host result buffer), done BEFORE the ranks agree
// to enter the reduce: an out-of-memory here is reported to the peers instead of leaving them inside ncclReduce.
int prepareReduce(mnenc_ctx* c, uint8_t** res) {
    StripeComm* sc = c->comm.get();
    *res = nullptr;
    if (c->stripeResult) resultRelease(c->stripeResult), c->stripeResult = nullptr;
    if (sc->rank != 0) return MNENC_OK;
    if (!c->copyStream) CU(cudaStreamCreateWithFlags(&c->copyStream, cudaStreamNonBlocking));
    for (auto& e : c->evChunk)
        if (!e) CU(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    *res = resultAcquire(stripeTotalBytes(c));
    if (!*res) return fail(MNENC_ERR_INTERNAL, "out of host memory");
    return MNENC_OK;
}

int reduceSegments(mnenc_ctx* c, void* words, size_t nWords, uint8_t* res) {
    StripeComm* sc = c->comm.get();
    std::string err;
    const NcclApi* nccl = ncclApi(&err);
    if (!nccl) return fail(MNENC_ERR_CUDA, err);
    const size_t totalBytes = stripeTotalBytes(c);
    const int chunks = nWords >= (size_t(1) << 20) ? 4 : 1;
    auto bail = [&](int rc) {
        resultRelease(res);
        return rc;
    };
    if (sc->sync && !sc->sync->arrive()) return bail(fail(MNENC_ERR_STATE, "another GPU of the group failed"));
    uint32_t* w = static_cast<uint32_t*>(words);
    const size_t per = (nWords + chunks - 1) / chunks;
    for (int i = 0; i < chunks; ++i) {
        const size_t off = per * i, n = off < nWords ? std::min(per, nWords - off) : 0;
        if (n == 0) break;
        if (ncclResult_t r = nccl->reduce(w + off, w + off, n, ncclUint32, ncclSum, 0, sc->comm, c->stream); r != ncclSuccess)
            return bail(fail(MNENC_ERR_CUDA, std::string("ncclReduce: ") + nccl->getErrorString(r)));
        if (sc->rank == 0 && 4 * off < totalBytes) {
            const size_t bytes = std::min(4 * n, totalBytes - 4 * off);
            cudaError_t e = cudaEventRecord(c->evChunk[i], c->stream);
            if (e == cudaSuccess) e = cudaStreamWaitEvent(c->copyStream, c->evChunk[i], 0);
            if (e == cudaSuccess) e = cudaMemcpyAsync(res + 4 * off, w + off, bytes, cudaMemcpyDeviceToHost, c->copyStream);
            if (e != cudaSuccess) return bail(fail(MNENC_ERR_CUDA, std::string("D2H of the stream: ") + cudaGetErrorString(e)));
        }
    }
    if (cudaError_t e = waitStream(c); e != cudaSuccess) return bail(fail(MNENC_ERR_CUDA, std::string("reduce of the segments: ") + cudaGetErrorString(e)));
    if (sc->rank == 0) {
        if (cudaError_t e = cudaStreamSynchronize(c->copyStream); e != cudaSuccess)
            return bail(fail(MNENC_ERR_CUDA, std::string("D2H of the stream: ") + cudaGetErrorString(e)));
        c->stripeResult = res;
    }
    return MNENC_OK;
}

int encodeStriped(mnenc_ctx* c, const uint8_t* rgb, int onDevice, size_t width, size_t height, size_t stride, uint32_t bx0, uint32_t bx1,
                  uint8_t** out, size_t* outBytes) {
    NvtxRange nvtxRange("mnenc:encode_striped");
    StripeComm* sc = c->comm.get();
    if (!sc) return fail(MNENC_ERR_STATE, "no communicator on this context (mnenc_comm_init_rank / mnenc_create_multi)");
    std::string err;
    const NcclApi* nccl = ncclApi(&err);
    if (!nccl) return fail(MNENC_ERR_CUDA, err);
    const int K = c->K, nr = sc->nranks, nSeq = 6 * K + 1;
    std::memset(sc->ms, 0, sizeof(sc->ms));
    double t = nowMs();
    auto lap = [&](int i) {
        const double now = nowMs();
        sc->ms[i] += static_cast<float>(now - t);
        t = now;
    };
    // Every phase's result code goes into the exchange that follows it (allGatherHost): a failing rank still joins that exchange
    // and all ranks return from there together, instead of the healthy ones waiting in NCCL for a peer that has left.
    struct InLibrary {  // mnenc_stripe_plan deals the table builds over the communicator only while this call runs
        mnenc_ctx* c;
        explicit InLibrary(mnenc_ctx* x) : c(x) { c->stripeInLibrary = true; }
        ~InLibrary() { c->stripeInLibrary = false; }
    } inLibrary(c);
    std::vector<uint32_t> rec1(3 * K + 3), all1(static_cast<size_t>(nr) * (3 * K + 3));
    int rc = mnenc_stripe_begin(c, rgb, onDevice, width, height, stride, sc->rank, nr, bx0, bx1, rec1.data());
    lap(0);
    if (int g = allGatherHost(c, nccl, rec1.data(), all1.data(), rec1.size() * 4, rc)) return g;
    lap(1);
    std::vector<uint32_t> rec2(static_cast<size_t>(6) * K * 6), all2(rec2.size() * nr);
    rc = mnenc_stripe_rle(c, all1.data(), rec2.data());
    lap(2);
    if (int g = allGatherHost(c, nccl, rec2.data(), all2.data(), rec2.size() * 4, rc)) return g;
    lap(3);
    void *hist = nullptr, *first = nullptr, *msym = nullptr;
    size_t nh = 0, nm = 0;
    rc = mnenc_stripe_histogram(c, all2.data(), &hist, &first, &nh, &msym, &nm);
    lap(4);
    if (nr > 1) {
        if (int g = stripeExchangeHistogram(c, rc)) return g;
    } else if (rc) {
        return rc;
    }
    lap(5);
    std::vector<uint64_t> bits(nSeq), allBits(static_cast<size_t>(nSeq) * nr);
    rc = mnenc_stripe_plan(c, bits.data());  // code tables dealt over the ranks; joins its own exchanges even when it fails
    lap(6);
    if (int g = allGatherHost(c, nccl, bits.data(), allBits.data(), bits.size() * 8, rc)) return g;
    lap(7);
    void* words = nullptr;
    size_t nWords = 0, nBytes = 0;
    rc = mnenc_stripe_pack(c, allBits.data(), &words, &nWords, &nBytes);
    lap(8);
    if (nr > 1) {
        uint8_t* res = nullptr;
        if (!rc) rc = prepareReduce(c, &res);
        if (int g = agreeStatus(c, nccl, rc)) {
            resultRelease(res);
            return g;
        }
        if (int g = reduceSegments(c, words, nWords, res)) return g;
    } else if (rc) {
        return rc;
    }
    lap(9);
    if (int rc2 = mnenc_stripe_finish(c, out, outBytes)) return rc2;
    lap(10);
    return MNENC_OK;
}

}  // namespace mnenc

using namespace mnenc;

extern "C" {

int mnenc_comm_unique_id(uint8_t* id128) try {
    if (!id128) return fail(MNENC_ERR_ARG, "null argument");
    std::string err;
    const NcclApi* nccl = ncclApi(&err);
    if (!nccl) return fail(MNENC_ERR_CUDA, err);
    static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId is 128 bytes");
    ncclUniqueId id;
    NC(nccl->getUniqueId(&id));
    std::memcpy(id128, &id, 128);
    return MNENC_OK;
} catch (...) {
    return mnenc::abiCaught();
}

int mnenc_comm_init_rank(mnenc_t* c, int rank, int nranks, const uint8_t* id128) try {
    if (!c || !id128 || nranks < 1 || rank < 0 || rank >= nranks) return fail(MNENC_ERR_ARG, "bad argument");
    std::string err;
    const NcclApi* nccl = ncclApi(&err);
    if (!nccl) return fail(MNENC_ERR_CUDA, err);
    CU(cudaSetDevice(c->device));
    std::unique_ptr<StripeComm, void (*)(StripeComm*)> sc(new StripeComm, destroyStripeComm);
    sc->rank = rank, sc->nranks = nranks;
    ncclUniqueId id;
    std::memcpy(&id, id128, 128);
    NC(nccl->commInitRank(&sc->comm, nranks, id, rank));
    if (int rc = reserveStage(sc.get())) return rc;
    c->comm = std::move(sc);
    return MNENC_OK;
} catch (...) {
    return mnenc::abiCaught();
}

int mnenc_comm_destroy(mnenc_t* c) try {
    if (!c) return fail(MNENC_ERR_ARG, "null argument");
    c->comm.reset();
    return MNENC_OK;
} catch (...) {
    return mnenc::abiCaught();
}

int mnenc_encode_striped_rgb8(mnenc_t* c, const uint8_t* rgb_stripe, int on_device, size_t width, size_t height, size_t row_stride_bytes,
                              uint32_t bx0, uint32_t bx1, uint8_t** out, size_t* out_bytes) try {
    if (!c || !out || !out_bytes) return fail(MNENC_ERR_ARG, "null argument");
    *out = nullptr;
    *out_bytes = 0;
    return encodeStriped(c, rgb_stripe, on_device, width, height, row_stride_bytes, bx0, bx1, out, out_bytes);
} catch (...) {
    return mnenc::abiCaught();
}

int mnenc_last_stripe_timing(mnenc_t* c, float* ms16) try {
    if (!c || !ms16) return fail(MNENC_ERR_ARG, "null argument");
    if (!c->comm) return fail(MNENC_ERR_STATE, "no communicator on this context");
    std::memcpy(ms16, c->comm->ms, sizeof(c->comm->ms));
    return MNENC_OK;
} catch (...) {
    return mnenc::abiCaught();
}

int mnenc_create_multi(mnenc_t** out, const int* devices, int n_devices, int K, int block_size, const float* base, int n_base,
                       const int32_t* detail_rows, const float* detailY, const float* detailU, const float* detailV) try {
    if (!out || !devices || n_devices < 1) return fail(MNENC_ERR_ARG, "bad argument");
    std::string err;
    const NcclApi* nccl = ncclApi(&err);
    if (!nccl) return fail(MNENC_ERR_CUDA, err);
    std::vector<mnenc_t*> ctxs;
    auto bail = [&](int rc) {
        for (mnenc_t* x : ctxs) mnenc_destroy(x);
        return rc;
    };
    for (int i = 0; i < n_devices; ++i) {
        mnenc_t* x = nullptr;
        if (int rc = mnenc_create(&x, devices[i], K, block_size, base, n_base, detail_rows, detailY, detailU, detailV)) return bail(rc);
        ctxs.push_back(x);
    }
    std::vector<ncclComm_t> comms(n_devices);
    if (ncclResult_t r = nccl->commInitAll(comms.data(), n_devices, devices); r != ncclSuccess)
        return bail(fail(MNENC_ERR_CUDA, std::string("ncclCommInitAll: ") + nccl->getErrorString(r)));
    auto sync = std::make_shared<GroupSync>();
    sync->n = n_devices;
    for (int i = 0; i < n_devices; ++i) {
        std::unique_ptr<StripeComm, void (*)(StripeComm*)> sc(new StripeComm, destroyStripeComm);
        sc->rank = i, sc->nranks = n_devices, sc->comm = comms[i], sc->sync = sync;
        CU(cudaSetDevice(devices[i]));
        if (int rc = reserveStage(sc.get())) return bail(rc);
        ctxs[i]->comm = std::move(sc);
        if (i) ctxs[0]->groupPeers.push_back(ctxs[i]);
    }
    *out = ctxs[0];
    return MNENC_OK;
} catch (...) {
    return mnenc::abiCaught();
}

int mnenc_debug_plan_sharded(const uint32_t* symbols, const uint32_t* freqs, const uint32_t* first_pos, const uint32_t* seq_start, int n_seq,
                             int nranks, uint32_t* codes, uint8_t* lens, uint64_t* payload_bits, uint32_t* golomb_m) try {
    if (!symbols || !freqs || !first_pos || !seq_start || !codes || !lens || !payload_bits || !golomb_m || n_seq < 1 || nranks < 1)
        return fail(MNENC_ERR_ARG, "bad argument");
    // rank 0 holds the entries in the caller's order, every other rank in reverse order (the merge kernel gives no order guarantee)
    std::vector<std::vector<SymbolStat>> fwd(n_seq), rev(n_seq);
    for (int q = 0; q < n_seq; ++q) {
        for (uint32_t e = seq_start[q]; e < seq_start[q + 1]; ++e) fwd[q].push_back(SymbolStat{symbols[e], freqs[e], first_pos[e]});
        rev[q].assign(fwd[q].rbegin(), fwd[q].rend());
    }
    std::vector<SequencePlan> plans(n_seq);
    std::vector<unsigned char> wire;
    planShardBuild(fwd, 0, nranks, nullptr, plans, wire);
    for (int r = 1; r < nranks; ++r) {
        std::vector<SequencePlan> theirs(n_seq);
        planShardBuild(rev, r, nranks, nullptr, theirs, wire);
        if (int rc = planShardApply(fwd, wire.data(), wire.size(), plans)) return rc;
    }
    for (int q = 0; q < n_seq; ++q) {
        payload_bits[q] = plans[q].payloadBits;
        golomb_m[q] = plans[q].golomb ? plans[q].golombM : 0u;
        for (uint32_t e = seq_start[q]; e < seq_start[q + 1]; ++e) {
            codes[e] = plans[q].golomb ? 0u : plans[q].code[e - seq_start[q]];
            lens[e] = plans[q].golomb ? 0 : plans[q].len[e - seq_start[q]];
        }
    }
    return MNENC_OK;
} catch (...) {
    return mnenc::abiCaught();
}

int mnenc_create_auto(mnenc_t** out, int K, int block_size, const float* base, int n_base, const int32_t* detail_rows, const float* detailY,
                      const float* detailU, const float* detailV) try {
    std::vector<int> devices;
    if (const char* list = std::getenv("MNENC_DEVICES")) {
        if (std::string(list) == "all") {
            int n = 0;
            CU(cudaGetDeviceCount(&n));
            for (int i = 0; i < n; ++i) devices.push_back(i);
        } else {
            for (const char* p = list; *p;) {
                char* end = nullptr;
                const long v = std::strtol(p, &end, 10);
                if (end == p || v < 0 || v > 1023) return fail(MNENC_ERR_ARG, "MNENC_DEVICES: expected a comma-separated list of device numbers or \"all\"");
                devices.push_back(static_cast<int>(v));
                p = *end == ',' ? end + 1 : end;
                if (*end && *end != ',') return fail(MNENC_ERR_ARG, "MNENC_DEVICES: expected a comma-separated list of device numbers or \"all\"");
            }
        }
    }
    if (devices.size() > 1) return mnenc_create_multi(out, devices.data(), static_cast<int>(devices.size()), K, block_size, base, n_base, detail_rows, detailY, detailU, detailV);
    int device = devices.empty() ? 0 : devices[0];
    if (devices.empty())
        if (const char* d = std::getenv("MNENC_DEVICE")) device = std::atoi(d);
    return mnenc_create(out, device, K, block_size, base, n_base, detail_rows, detailY, detailU, detailV);
} catch (...) {
    return mnenc::abiCaught();
}

}  // extern "C"

namespace mnenc {

// mnenc_encode_rgb8 on a context made by mnenc_create_multi: one host thread per GPU, each takes its block columns of the
// host image through the striped encode; the stream comes back from the first GPU.
int encodeGroup(mnenc_ctx* c, const uint8_t* rgb, size_t width, size_t height, size_t stride, uint8_t** out, size_t* outBytes) {
    NvtxRange nvtxRange("mnenc:encode_group");
    std::vector<mnenc_ctx*> ranks{c};
    ranks.insert(ranks.end(), c->groupPeers.begin(), c->groupPeers.end());
    const int nr = static_cast<int>(ranks.size());
    const uint32_t nbx = static_cast<uint32_t>((width + 7) / 8);
    std::vector<int> rcs(nr, MNENC_OK);
    std::vector<std::string> errs(nr);
    std::vector<uint8_t*> outs(nr, nullptr);
    std::vector<size_t> sizes(nr, 0);
    auto work = [&](int r) {
        const uint32_t bx0 = static_cast<uint32_t>(static_cast<uint64_t>(nbx) * r / nr), bx1 = static_cast<uint32_t>(static_cast<uint64_t>(nbx) * (r + 1) / nr);
        rcs[r] = mnenc_encode_striped_rgb8(ranks[r], rgb + 3 * 8 * static_cast<size_t>(bx0), 0, width, height, stride, bx0, bx1, &outs[r], &sizes[r]);
        if (rcs[r]) {
            errs[r] = mnenc_last_error();
            if (c->comm->sync) c->comm->sync->abort();
        }
    };
    if (c->comm->sync) c->comm->sync->reset();
    std::vector<std::thread> threads;
    for (int r = 1; r < nr; ++r) threads.emplace_back(work, r);
    work(0);
    for (auto& th : threads) th.join();
    int bad = -1;  // the rank that failed first, not one that gave up at the meeting point because of it
    for (int r = 0; r < nr; ++r)
        if (rcs[r] && (bad < 0 || (errs[bad].find("of the group failed") != std::string::npos &&  // "another GPU ..." / "rank r ..."
                                   errs[r].find("of the group failed") == std::string::npos)))
            bad = r;
    if (bad >= 0) {
        for (uint8_t* p : outs) resultRelease(p);
        return fail(rcs[bad], "GPU " + std::to_string(ranks[bad]->device) + ": " + errs[bad]);
    }
    *out = outs[0];
    *outBytes = sizes[0];
    return MNENC_OK;
}

}  // namespace mnenc
