// C-ABI layer (include/mnenc.h) and the per-image orchestration of the encoder:
//   H2D pixels -> matching-pursuit kernel -> stream scatter / DC delta / RLE -> histogram -> (host) code
//   tables -> bit packing -> D2H -> header merge.
// The host work between the two GPU phases is the table build of entropy_host.cpp (alphabet-sized, never
// symbol-sized). Buffers are grow-only and live in the context, so steady-state encodes allocate nothing.
#include <cuda_runtime.h>
#include <sched.h>

#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <mutex>
#include <numeric>
#include <thread>
#include <string>
#include <vector>

#include "mnenc_ctx.cuh"

namespace mnenc {

namespace {
thread_local std::string g_error;
}

int fail(int code, const std::string& msg) {
    g_error = msg;
    return code;
}

// No exception crosses the ABI (include/mnenc.h): every int-returning entry point is a function-try-block that ends here.
// The context may be left mid-call (its scratch buffers are rewritten by the next call; the histogram may hold stale bins
// after an exception between the histogram and its reset, so the caller should destroy the context after MNENC_ERR_INTERNAL).
int abiCaught() noexcept {
    try {
        try {
            throw;
        } catch (const std::bad_alloc&) {
            return fail(MNENC_ERR_INTERNAL, "out of host memory");
        } catch (const std::exception& e) {
            return fail(MNENC_ERR_INTERNAL, std::string("unexpected exception: ") + e.what());
        } catch (...) {
            return fail(MNENC_ERR_INTERNAL, "unexpected exception");
        }
    } catch (...) {  // not even the text could be stored
        return MNENC_ERR_INTERNAL;
    }
}

void mergeOverlays(uint8_t* res, const std::vector<Overlay>& overlays) {
    for (const Overlay& o : overlays) {
        const uint64_t shift = o.bitPos & 7;
        size_t byte = static_cast<size_t>(o.bitPos >> 3);
        if (shift == 0) {
            for (size_t i = 0; i < o.bits.bytes.size(); ++i) res[byte + i] |= o.bits.bytes[i];
        } else {
            for (size_t i = 0; i < o.bits.bytes.size(); ++i) {
                res[byte + i] |= static_cast<uint8_t>(o.bits.bytes[i] >> shift);
                const uint8_t spill = static_cast<uint8_t>(o.bits.bytes[i] << (8 - shift));
                if (spill) res[byte + i + 1] |= spill;
            }
        }
    }
}

namespace {
struct ResultPool {
    struct Buf {
        uint8_t* p;
        size_t cap;
        bool inUse;
    };
    std::mutex m;
    std::vector<Buf> bufs;
    // Idle buffers kept for reuse. cudaHostAlloc / cudaFreeHost synchronise the device, so the limits are generous: a batch of
    // 16 streams in flight must recycle without ever touching the allocator in steady state.
    // Decoded images are 48 MB each at 16 MP: a batch of 32 holds 1.7 GB at once, hence 4 GiB (MNENC_POOL_MB overrides).
    // The count limit only bounds the linear searches: a batch of 1024 small frames (BASELINE.json configs[4]) returns 1024
    // buffers at once - with 64 slots every batch freed and re-pinned most of them, and those calls serialised the lanes
    // (192 frames of 1080p: 2131 -> ... MP/s, profiles/r02_batch1080p.md).
    static constexpr size_t kMaxIdle = 2048;
    size_t maxIdleBytes = size_t(4) << 30;
    ResultPool() {
        if (const char* e = std::getenv("MNENC_POOL_MB")) maxIdleBytes = static_cast<size_t>(std::max(0L, std::atol(e))) << 20;
    }
};
ResultPool& resultPool() {
    static ResultPool* pool = new ResultPool;  // leaked on purpose: no CUDA calls from static destructors
    return *pool;
}
}  // namespace

uint8_t* resultAcquire(size_t bytes) {
    ResultPool& rp = resultPool();
    {
        std::lock_guard<std::mutex> g(rp.m);
        ResultPool::Buf* best = nullptr;
        for (auto& b : rp.bufs)
            if (!b.inUse && b.cap >= bytes && (!best || b.cap < best->cap)) best = &b;
        if (best && best->cap <= 4 * bytes + (size_t(1) << 20)) {  // do not pin a 56 MB buffer under a 600-byte stream
            best->inUse = true;
            return best->p;
        }
    }
    void* p = nullptr;
    const size_t cap = bytes + bytes / 8 + 4096;
    if (cudaHostAlloc(&p, cap, cudaHostAllocPortable) != cudaSuccess) {
        cudaGetLastError();
        return static_cast<uint8_t*>(std::malloc(bytes ? bytes : 1));  // plain memory still works as a copy target
    }
    std::lock_guard<std::mutex> g(rp.m);
    rp.bufs.push_back({static_cast<uint8_t*>(p), cap, true});
    return static_cast<uint8_t*>(p);
}

void resultRelease(uint8_t* p) {
    if (!p) return;
    ResultPool& rp = resultPool();
    uint8_t* drop = nullptr;
    {
        std::lock_guard<std::mutex> g(rp.m);
        bool mine = false;
        size_t idle = 0, idleBytes = 0;
        for (auto& b : rp.bufs) {
            if (b.p == p) b.inUse = false, mine = true;
            if (!b.inUse) ++idle, idleBytes += b.cap;
        }
        if (!mine) {
            std::free(p);
            return;
        }
        if (idle > ResultPool::kMaxIdle || idleBytes > rp.maxIdleBytes) {
            size_t k = rp.bufs.size();
            for (size_t i = 0; i < rp.bufs.size(); ++i)
                if (!rp.bufs[i].inUse && (k == rp.bufs.size() || rp.bufs[i].cap < rp.bufs[k].cap)) k = i;
            drop = rp.bufs[k].p;
            rp.bufs.erase(rp.bufs.begin() + static_cast<long>(k));
        }
    }
    if (drop) cudaFreeHost(drop);
}

int ensureHistogram(mnenc_ctx* c) {
    if (c->histDirty) c->histReady = false, c->histDirty = false;  // an earlier encode failed between the histogram and its reset
    if (c->histReady) return MNENC_OK;
    const size_t bins = static_cast<size_t>(kMaxSeq) * kHistBins;
    CU(c->hist.ensure(bins * 4));
    CU(c->firstPos.ensure(bins * 4));
    CU(c->maxSym.ensure(kMaxSeq * 4));
    CU(c->nEntries.ensure(4));
    CU(cudaMemsetAsync(c->hist.p, 0, bins * 4, c->stream));
    launchFillU32(c->firstPos.as<uint32_t>(), bins, kNoPos, c->stream);
    CU(cudaMemsetAsync(c->maxSym.p, 0, kMaxSeq * 4, c->stream));
    CU(cudaMemsetAsync(c->nEntries.p, 0, 4, c->stream));
    if (c->entryCap == 0) {
        c->entryCap = 1u << 16;
        CU(c->entries.ensure(static_cast<size_t>(c->entryCap) * sizeof(HistEntry)));
    }
    c->histReady = true;
    return MNENC_OK;
}

// Everything after the matching pursuit: counts/pairs (device) -> .mn bytes.
int runBackend(mnenc_ctx* c, size_t width, size_t height, uint8_t** out, size_t* outBytes) {
    NvtxRange nvtxRange("mnenc:backend (RLE, Huffman/Golomb plan, bit packing)");
    const int K = c->K;
    const int nStreams = 6 * K, nSeq = nStreams + 1;
    const uint32_t nbx = static_cast<uint32_t>((width + 7) / 8), nby = static_cast<uint32_t>((height + 7) / 8);
    const uint32_t nblocks = nbx * nby;
    const uint32_t ntiles = (nblocks + kTileBlocks - 1) / kTileBlocks;
    cudaStream_t st = c->stream;

    BackendBuffers b{};
    b.counts = c->counts.as<uint16_t>() + c->frameBc;  // frameBc: this frame's slice when several frames shared one pursuit
    b.pairs = c->pairs.as<uint32_t>() + static_cast<size_t>(c->frameBc) * K;
    b.K = K;
    b.nblocks = nblocks;
    b.ntiles = ntiles;
    CU(c->tileCounts.ensure(static_cast<size_t>(ntiles) * 3 * K * 4 + 4));
    CU(c->totals.ensure(3 * K * 4));
    CU(c->dcTmp.ensure(static_cast<size_t>(nblocks) * 3 * 2 + 2));
    CU(c->hTotals.ensure(3 * K * 4));
    b.tileCounts = c->tileCounts.as<uint32_t>();
    b.totals = c->totals.as<uint32_t>();
    b.dcTmp = c->dcTmp.as<uint16_t>();

    // ---- phase 1: stream sizes ----
    CU(cudaEventRecord(c->ev[2], st));
    c->launches += launchTileCounts(b, st);
    CU(cudaMemcpyAsync(c->hTotals.p, b.totals, 3 * K * 4, cudaMemcpyDeviceToHost, st));
    CU(waitStream(c));
    StreamOffsets so{};
    uint64_t N64 = 0;
    for (int ch = 0; ch < 3; ++ch)
        for (int d = 0; d < K; ++d) {
            const uint32_t n = c->hTotals.as<uint32_t>()[ch * K + d];
            so.off[2 * (ch * K + d)] = static_cast<uint32_t>(N64);
            so.off[2 * (ch * K + d) + 1] = static_cast<uint32_t>(N64 + n);
            N64 += 2ull * n;
        }
    if (N64 + 3ull * nblocks >= 0xFFFFFFF0ull) return fail(MNENC_ERR_ARG, "image too large: more than 2^32 symbols");
    const uint32_t N = static_cast<uint32_t>(N64);
    so.off[nStreams] = N;

    // ---- phase 2: scatter, DC delta, RLE, decision, histogram ----
    const uint32_t upperM = N + 3 * nblocks;
    CU(c->arena.ensure(static_cast<size_t>(N) * 2 + 16));
    CU(c->flags.ensure(static_cast<size_t>(N) + 16));
    CU(c->runIdx.ensure((static_cast<size_t>(N) + 2) * 4));
    CU(c->startPos.ensure((static_cast<size_t>(N) + 2) * 4));
    CU(c->outOff.ensure((static_cast<size_t>(N) + 2) * 4));
    CU(c->rleArena.ensure((static_cast<size_t>(N) + N / 2 + 16) * 2));
    CU(c->sums32.ensure((static_cast<size_t>(upperM) / 2048 + 4) * 4));
    CU(c->sums64.ensure((static_cast<size_t>(upperM) / 2048 + 4) * 8));
    CU(c->seqTable.ensure(sizeof(SeqTable)));
    CU(c->hSeqTable.ensure(sizeof(SeqTable)));
    CU(c->hNEntries.ensure(4));
    if (int rc = ensureHistogram(c)) return rc;
    b.arena = c->arena.as<uint16_t>();
    b.flags = c->flags.as<uint8_t>();
    b.runIdx = c->runIdx.as<uint32_t>();
    b.startPos = c->startPos.as<uint32_t>();
    b.outOff = c->outOff.as<uint32_t>();
    b.rleArena = c->rleArena.as<uint16_t>();
    b.blockSums32 = c->sums32.as<uint32_t>();
    b.blockSums64 = c->sums64.as<unsigned long long>();
    b.seqTable = c->seqTable.as<SeqTable>();
    b.hist = c->hist.as<uint32_t>();
    b.firstPos = c->firstPos.as<uint32_t>();
    b.maxSym = c->maxSym.as<uint32_t>();
    b.entries = c->entries.as<HistEntry>();
    b.nEntries = c->nEntries.as<uint32_t>();
    b.entryCap = c->entryCap;

    c->launches += launchScatter(b, so, st);
    c->launches += launchRleAndDecide(b, so, N, st);
    c->launches += launchDecide(b, so, nullptr, st);
    CU(cudaEventRecord(c->ev[3], st));
    c->histDirty = true;  // until the touched bins have been reset: an early return leaves stale counts (ensureHistogram)
    c->launches += launchHistogram(b, upperM, nSeq, st);
    CU(cudaEventRecord(c->ev[4], st));
    if (c->failNextBackend) {  // test hook (mnenc_debug_fail_next_backend): leave through an ordinary error return with the bins populated
        c->failNextBackend = false;
        return fail(MNENC_ERR_INTERNAL, "injected failure after the histogram (test hook)");
    }
    const uint32_t firstChunk = 1u << 16;
    CU(c->hEntries.ensure(static_cast<size_t>(firstChunk) * sizeof(HistEntry)));
    CU(cudaMemcpyAsync(c->hSeqTable.p, b.seqTable, sizeof(SeqTable), cudaMemcpyDeviceToHost, st));
    CU(cudaMemcpyAsync(c->hNEntries.p, b.nEntries, 4, cudaMemcpyDeviceToHost, st));
    CU(cudaMemcpyAsync(c->hEntries.p, b.entries, static_cast<size_t>(firstChunk) * sizeof(HistEntry), cudaMemcpyDeviceToHost, st));
    CU(waitStream(c));
    uint32_t nEntries = *c->hNEntries.as<uint32_t>();
    if (nEntries > c->entryCap) {
        // Rare: more distinct (sequence, symbol) pairs than the list holds. Grow and redo the histogram.
        c->entryCap = nEntries + nEntries / 4;
        CU(c->entries.ensure(static_cast<size_t>(c->entryCap) * sizeof(HistEntry)));
        c->histReady = false;
        if (int rc = ensureHistogram(c)) return rc;
        c->histDirty = true;
        b.entries = c->entries.as<HistEntry>();
        b.entryCap = c->entryCap;
        c->launches += launchHistogram(b, upperM, nSeq, st);
        CU(cudaMemcpyAsync(c->hNEntries.p, b.nEntries, 4, cudaMemcpyDeviceToHost, st));
        CU(waitStream(c));
        nEntries = *c->hNEntries.as<uint32_t>();
        if (nEntries > c->entryCap) return fail(MNENC_ERR_INTERNAL, "histogram entry list overflow");
    }
    if (nEntries > firstChunk || c->hEntries.cap < static_cast<size_t>(nEntries) * sizeof(HistEntry)) {
        CU(c->hEntries.ensure(static_cast<size_t>(nEntries) * sizeof(HistEntry)));
        CU(cudaMemcpyAsync(c->hEntries.p, b.entries, static_cast<size_t>(nEntries) * sizeof(HistEntry), cudaMemcpyDeviceToHost, st));
        CU(waitStream(c));
    }

    // ---- host: code tables, Huffman-vs-Golomb choice, bit layout ----
    const double tHost0 = nowMs();
    const SeqTable& seq = *c->hSeqTable.as<SeqTable>();
    const HistEntry* ent = c->hEntries.as<HistEntry>();
    std::vector<std::vector<SymbolStat>> stats(nSeq);
    std::vector<std::vector<uint32_t>> entryOf(nSeq);
    for (uint32_t e = 0; e < nEntries; ++e) {
        if (ent[e].seq >= static_cast<uint32_t>(nSeq)) return fail(MNENC_ERR_INTERNAL, "bad histogram entry");
        stats[ent[e].seq].push_back(SymbolStat{ent[e].symbol, ent[e].freq, ent[e].firstPos});
        entryOf[ent[e].seq].push_back(e);
    }
    std::vector<SequencePlan> plans(nSeq);
    {   // largest alphabets first so the dynamic distribution over the pool ends evenly
        std::vector<int> byLoad(nSeq);
        std::iota(byLoad.begin(), byLoad.end(), 0);
        std::sort(byLoad.begin(), byLoad.end(), [&](int a, int b) { return stats[a].size() > stats[b].size(); });
        c->pool->parallelFor(nSeq, [&](int i) { planSequence(stats[byLoad[i]], plans[byLoad[i]]); });
    }

    CU(c->hEntryCodes.ensure(static_cast<size_t>(nEntries) * 8 + 8));
    CU(c->hCodeTable.ensure(sizeof(CodeTable)));
    CU(c->entryCodes.ensure(static_cast<size_t>(nEntries) * 8 + 8));
    CU(c->codeTable.ensure(sizeof(CodeTable)));
    unsigned long long* hCodes = c->hEntryCodes.as<unsigned long long>();
    CodeTable& hTable = *c->hCodeTable.as<CodeTable>();
    std::memset(&hTable, 0, sizeof(hTable));

    std::vector<Overlay> overlays;
    overlays.reserve(2 * nSeq + 1);
    uint64_t cursor = 0;
    {  // file header (CompressedImage.cpp:472-486)
        Overlay h;
        h.bitPos = 0;
        h.bits.put(0x4D4E3234u, 32);
        h.bits.put(static_cast<uint32_t>(width), 32);
        h.bits.put(static_cast<uint32_t>(height), 32);
        h.bits.put(static_cast<uint32_t>(K) & 0xFF, 8);
        h.bits.put(8, 8);
        for (int i = 0; i < 3 * K; ++i) h.bits.put(static_cast<uint16_t>(c->hQuant[(i / K) * kMaxK + (i % K)]), 16);
        cursor = h.bits.nbits;
        overlays.push_back(std::move(h));
    }
    uint64_t scanPrefix = 0;  // device exclusive scan value at the first symbol of sequence q
    uint64_t totalPayload = 0;
    for (int q = 0; q < nSeq; ++q) {
        SequencePlan& p = plans[q];
        if (!p.golomb && p.maxCodeLen > 32) return fail(MNENC_ERR_INTERNAL, "Huffman code longer than 32 bits");
        Overlay h;
        h.bitPos = cursor;
        if (q > 0) {  // RLE flag (+ RLE length) precedes every symbol stream (CompressedImage.cpp:509-515)
            if (seq.desc[q].src == 2) {
                h.bits.put(1, 1);
                h.bits.put(seq.desc[q].len, 32);
            } else {
                h.bits.put(0, 1);
            }
        }
        h.bits.append(p.header);
        cursor += h.bits.nbits;
        overlays.push_back(std::move(h));
        hTable.seq[q].bitBase = cursor - scanPrefix;  // modular arithmetic: position = bitBase + scan value
        hTable.seq[q].golombM = p.golomb ? p.golombM : 0;
        cursor += p.payloadBits;
        scanPrefix += p.payloadBits;
        totalPayload += p.payloadBits;
        if (!p.golomb) {
            Overlay e;
            e.bitPos = cursor;
            e.bits.put(p.eofCode, p.eofLen);
            cursor += p.eofLen;
            overlays.push_back(std::move(e));
            for (size_t k = 0; k < entryOf[q].size(); ++k)
                hCodes[entryOf[q][k]] = static_cast<unsigned long long>(p.code[k]) | (static_cast<unsigned long long>(p.len[k]) << 32);
        } else {
            for (size_t k = 0; k < entryOf[q].size(); ++k) hCodes[entryOf[q][k]] = 0;
        }
    }
    c->planHuffman = c->planGolomb = c->planRle = 0;
    for (int q = 0; q < nSeq; ++q) {
        (plans[q].golomb ? c->planGolomb : c->planHuffman) += 1;
        if (q > 0 && seq.desc[q].src == 2) c->planRle += 1;
    }
    const uint64_t totalBits = cursor;
    const size_t totalBytes = static_cast<size_t>((totalBits + 7) / 8);
    const uint32_t M = seq.start[nSeq];
    c->ms[4] = static_cast<float>(nowMs() - tHost0);

    // ---- phase 3: bit packing ----
    const size_t outAlloc = (totalBytes + 16 + 3) / 4 * 4;
    CU(c->outWords.ensure(outAlloc));
    CU(cudaMemcpyAsync(c->entryCodes.p, hCodes, static_cast<size_t>(nEntries) * 8, cudaMemcpyHostToDevice, st));
    CU(cudaMemcpyAsync(c->codeTable.p, &hTable, sizeof(CodeTable), cudaMemcpyHostToDevice, st));
    CU(cudaMemsetAsync(c->outWords.p, 0, outAlloc, st));
    b.entryCodes = c->entryCodes.as<unsigned long long>();
    b.codeTable = c->codeTable.as<CodeTable>();
    b.outWords = c->outWords.as<uint32_t>();
    CU(cudaEventRecord(c->ev[5], st));
    c->launches += launchPack(b, M, nSeq, st);
    c->launches += launchResetHistogram(b, nEntries, st);
    c->histDirty = false;
    CU(cudaEventRecord(c->ev[6], st));
    uint8_t* res = resultAcquire(totalBytes);
    if (!res) return fail(MNENC_ERR_INTERNAL, "out of host memory");
    if (cudaError_t e = cudaMemcpyAsync(res, c->outWords.p, totalBytes, cudaMemcpyDeviceToHost, st); e != cudaSuccess) {
        resultRelease(res);
        return fail(MNENC_ERR_CUDA, std::string("D2H of the stream: ") + cudaGetErrorString(e));
    }
    if (cudaError_t e = waitStream(c); e != cudaSuccess) {
        resultRelease(res);
        return fail(MNENC_ERR_CUDA, std::string("cudaStreamSynchronize: ") + cudaGetErrorString(e));
    }
    const double tMerge0 = nowMs();

    // ---- merge host-written bits ----
    mergeOverlays(res, overlays);
    *out = res;
    *outBytes = totalBytes;
    c->nSymbols = M;
    c->nBits = totalPayload;
    float t;
    cudaEventElapsedTime(&t, c->ev[2], c->ev[3]);
    c->ms[2] = t;
    cudaEventElapsedTime(&t, c->ev[3], c->ev[4]);
    c->ms[3] = t;
    cudaEventElapsedTime(&t, c->ev[5], c->ev[6]);
    c->ms[5] = t;
    c->ms[6] = static_cast<float>(nowMs() - tMerge0);
    return MNENC_OK;
}

// Level-synchronous matching pursuit: one round of kernels per greedy iteration (mp_wave.cu). The host queues rounds
// kWaveLookahead iterations ahead of the GPU and stops as soon as a finished round reports no active block-channel.
static int runWavePursuit(mnenc_ctx* c, const uint8_t* rgbDev, size_t width, size_t height, size_t stride, uint32_t nbx, uint32_t nby,
                          uint32_t nbc, const uint8_t* const* framesDev, uint32_t frameBlocks) {
    constexpr int kWaveLookahead = 3;
    if (nbc > 0x07FFFFFFu) return fail(MNENC_ERR_ARG, "image too large: more than 2^27 block-channels");
    const int K = c->K, nbins = 3 * c->nBase;
    const size_t n = nbc;
    CU(c->wRes32.ensure(n * kN * 4));
    CU(c->wRes16.ensure(n * kN * 2));
    CU(c->wState.ensure(n * 16));
    CU(c->wGroups.ensure(n * kMaxK * 2));
    CU(c->wBest.ensure(n * 8));
    CU(c->wAmax.ensure(n * 4));
    CU(c->wNcand.ensure(n * 4));
    CU(c->wCand.ensure(n * kCandSlots * 4));
    CU(c->wNcandDetail.ensure(n * 4));
    CU(c->wCandDetail.ensure(n * kCandSlotsDetail * 4));
    CU(c->wAct0.ensure(n * 4));
    CU(c->wAct1.ensure(n * 4));
    CU(c->wMembers.ensure(waveMembersCapacity(nbc, K) * 4));
    CU(c->wMemberInfo.ensure(waveMembersCapacity(nbc, K) * 8));
    const uint32_t ovfCap = nbc / 2 + 65536;
    CU(c->wOvf.ensure(static_cast<size_t>(ovfCap) * 8));
    const size_t binWords = static_cast<size_t>(4 * nbins + 2), headWords = kMaxK + 4;  // nAct[K + 1] | overflow count | binCount | binCursor | binOffset | binTileStart
    CU(c->wBins.ensure((headWords + binWords) * 4));
    CU(c->hWaveAct.ensure((kMaxK + 2) * 4));
    CU(cudaMemsetAsync(c->wBins.p, 0, (headWords + binWords) * 4, c->stream));
    WaveParams p{};
    p.rgb = rgbDev;
    p.frames = framesDev;
    p.frameBlocks = frameBlocks;
    p.stride = stride;
    p.width = static_cast<int>(width);
    p.height = static_cast<int>(height);
    p.nbx = static_cast<int>(nbx);
    p.nby = static_cast<int>(nby);
    p.K = K;
    p.nBase = c->nBase;
    p.baseUm16 = c->baseUm16.as<uint16_t>();
    p.detUm16 = c->detUm16.as<uint16_t>();
    p.rows32 = c->rows32.as<float>();
    p.detRows = c->detRows.as<uint8_t>();
    p.quant = c->quant.as<float>();
    p.epsRel16 = c->epsRel16Hi;
    p.epsAbs16 = c->epsAbs16;
    p.counts = c->counts.as<uint16_t>();
    p.pairs = c->pairs.as<uint32_t>();
    p.res32 = c->wRes32.as<float>();
    p.res16 = c->wRes16.as<uint16_t>();
    p.state = c->wState.as<uint4>();
    p.groups = c->wGroups.as<uint16_t>();
    p.best = c->wBest.as<unsigned long long>();
    p.amax = c->wAmax.as<uint32_t>();
    p.ncand = c->wNcand.as<uint32_t>();
    p.cand = c->wCand.as<uint32_t>();
    p.ncandDetail = c->wNcandDetail.as<uint32_t>();
    p.candDetail = c->wCandDetail.as<uint32_t>();
    p.act[0] = c->wAct0.as<uint32_t>();
    p.act[1] = c->wAct1.as<uint32_t>();
    uint32_t* words = c->wBins.as<uint32_t>();
    p.nAct = words;
    p.ovfCount = words + kMaxK + 2;
    p.ovf = c->wOvf.as<uint2>();
    p.ovfCap = ovfCap;
    p.memberInfo = c->wMemberInfo.as<uint2>();
    p.binCount = words + headWords;
    p.binCursor = p.binCount + nbins;
    p.binOffset = p.binCursor + nbins;
    p.binTileStart = p.binOffset + nbins + 1;
    p.members = c->wMembers.as<uint32_t>();
    p.stats = reinterpret_cast<unsigned long long*>(c->mpScratch.as<unsigned char>() + 16);
    p.nbc = nbc;
    CU(launchWaveInit(p, c->stream));
    c->launches += 1;
    uint32_t* hAct = c->hWaveAct.as<uint32_t>();
    CU(cudaMemcpyAsync(hAct, p.nAct, 4, cudaMemcpyDeviceToHost, c->stream));
    uint32_t hint = nbc;
    static const bool trace = std::getenv("MNENC_WAVE_TRACE") != nullptr;  // host clock per round (development aid)
    const double tTrace = nowMs();
    for (int r = 0; r < K; ++r) {
        double tWait = 0;
        if (r >= kWaveLookahead) {  // round r - kWaveLookahead has finished or is about to: its survivor count bounds this round's work
            const double t0 = nowMs();
            CU(cudaEventSynchronize(c->evWave[r - kWaveLookahead]));
            tWait = nowMs() - t0;
            hint = hAct[r - kWaveLookahead + 1];
            if (hint == 0) break;
        }
        const double t1 = nowMs();
        CU(launchWaveIteration(p, r, hint, c->numSMs, c->stream, &c->launches));
        CU(cudaMemcpyAsync(hAct + r + 1, p.nAct + r + 1, 4, cudaMemcpyDeviceToHost, c->stream));
        CU(cudaEventRecord(c->evWave[r], c->stream));
        if (trace) std::fprintf(stderr, "wave round %d: at %.3f ms, waited %.3f ms, enqueue %.3f ms, hint %u\n", r, t1 - tTrace, tWait, nowMs() - t1, hint);
    }
    return MNENC_OK;
}

// Block slots per frame when several frames share one pursuit: a multiple of 8, so that every frame's slice of counts / pairs
// starts on a 16-byte boundary.
uint32_t frameBlockSlots(size_t width, size_t height) {
    const uint64_t nb = static_cast<uint64_t>((width + 7) / 8) * ((height + 7) / 8);
    return static_cast<uint32_t>((nb + 7) / 8 * 8);
}

int runMatchingPursuit(mnenc_ctx* c, const uint8_t* rgbDev, size_t width, size_t height, size_t stride, const uint8_t* const* framesDev,
                       uint32_t nFrames) {
    NvtxRange nvtxRange("mnenc:matching_pursuit");
    const uint32_t nbx = static_cast<uint32_t>((width + 7) / 8), nby = static_cast<uint32_t>((height + 7) / 8);
    const uint32_t frameBlocks = framesDev ? frameBlockSlots(width, height) : 0u;
    const uint64_t nbc64 = framesDev ? 3ull * frameBlocks * nFrames : 3ull * nbx * nby;
    c->frameBc = 0;
    if (nbc64 >= 0x7FFFFFFFull) return fail(MNENC_ERR_ARG, "image too large");
    const uint32_t nbc = static_cast<uint32_t>(nbc64);
    CU(c->counts.ensure(static_cast<size_t>(nbc) * 2 + 16));
    CU(c->pairs.ensure(static_cast<size_t>(nbc) * c->K * 4 + 16));
    CU(c->mpScratch.ensure(64));
    CU(c->hStats.ensure(64));
    CU(cudaMemsetAsync(c->mpScratch.p, 0, 64, c->stream));
    CU(cudaEventRecord(c->ev[0], c->stream));
    if (int rc = runWavePursuit(c, rgbDev, width, height, stride, nbx, nby, nbc, framesDev, frameBlocks)) return rc;
    CU(cudaEventRecord(c->ev[1], c->stream));
    CU(cudaMemcpyAsync(c->hStats.p, c->mpScratch.as<unsigned char>() + 16, 48, cudaMemcpyDeviceToHost, c->stream));
    c->lastW = width;
    c->lastH = height;
    c->haveSymbols = framesDev == nullptr;  // mnenc_get_symbols describes ONE image
    return MNENC_OK;
}

void finishMpAccounting(mnenc_ctx* c) {
    float t = 0;
    cudaEventElapsedTime(&t, c->ev[0], c->ev[1]);
    c->ms[1] = t;
    const unsigned long long* s = c->hStats.as<unsigned long long>();
    // F = sum over iterations of 2*64*D_r  +  2*64*count   
    c->mpFlops = 2.0 * kN * static_cast<double>(s[0]) + 2.0 * kN * static_cast<double>(s[1]);
    if (std::getenv("MNENC_WAVE_TRACE"))
        std::fprintf(stderr, "wave candidates: base %llu (no slot: %llu), groups %llu; block-channel iterations %llu\n", s[2], s[3], s[4], s[1]);
}

int checkEncodeArgs(mnenc_ctx* c, const void* rgb, size_t w, size_t h, size_t stride, uint8_t** out, size_t* outBytes) {
    if (!c || !rgb || !out || !outBytes) return fail(MNENC_ERR_ARG, "null argument");
    if (w == 0 || h == 0 || stride < 3 * w) return fail(MNENC_ERR_ARG, "bad image geometry");
    if (w > 0xFFFFFFFFull || h > 0xFFFFFFFFull) return fail(MNENC_ERR_ARG, "image too large");
    if (!c->quantSet) return fail(MNENC_ERR_STATE, "mnenc_set_quant has not been called");
    return MNENC_OK;
}

}  // namespace mnenc

using namespace mnenc;

// ================================================================================================
// C ABI
// ================================================================================================
extern "C" {

const char* mnenc_last_error(void) { return g_error.c_str(); }

void mnenc_free(uint8_t* p) { resultRelease(p); }

int mnenc_quant_tables(int K, int block_size, double bpp_allocation, float* out_3K) try {
    if (!out_3K) return fail(MNENC_ERR_ARG, "null argument");
    if (mnenc::quantTables(K, block_size, bpp_allocation, out_3K) != 0) return fail(MNENC_ERR_ARG, "unsupported K / block size");
    return MNENC_OK;
} catch (...) {
    return mnenc::abiCaught();
}

int mnenc_create(mnenc_t** out, int device, int K, int block_size, const float* base, int n_base, const int32_t* detail_rows,
                 const float* detailY, const float* detailU, const float* detailV) try {
    using namespace mnenc;
    if (!out || !base || !detail_rows || !detailY || !detailU || !detailV) return fail(MNENC_ERR_ARG, "null argument");
    if (block_size != 8) return fail(MNENC_ERR_ARG, "only block_size 8 is supported");
    if (K < 1 || K > kMaxK) return fail(MNENC_ERR_ARG, "K must be in [1,32]");
    if (n_base < 1 || n_base > kRowsPad) return fail(MNENC_ERR_ARG, "n_base must be in [1,512]");
    for (int i = 0; i < n_base; ++i)
        if (detail_rows[i] < 32 || detail_rows[i] > 64) return fail(MNENC_ERR_ARG, "detail groups must have 32..64 rows");
    int ndev = 0;
    CU(cudaGetDeviceCount(&ndev));
    if (device < 0 || device >= ndev) return fail(MNENC_ERR_CUDA, "no such CUDA device");
    CU(cudaSetDevice(device));
    cudaDeviceProp prop;
    CU(cudaGetDeviceProperties(&prop, device));
    if (prop.major != 10) return fail(MNENC_ERR_CUDA, "mnenc is built for sm_100a (B200) only");
    mnenc_ctx* c = new mnenc_ctx;
    c->device = device;
    c->K = K;
    c->nBase = n_base;
    c->numSMs = prop.multiProcessorCount;
    auto bail = [&](int rc) {
        mnenc_destroy(c);
        return rc;
    };
#define CUB(expr)                                                                                              \
    do {                                                                                                       \
        cudaError_t e_ = (expr);                                                                               \
        if (e_ != cudaSuccess) return bail(fail(MNENC_ERR_CUDA, std::string(#expr) + ": " + cudaGetErrorString(e_))); \
    } while (0)
    CUB(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
    for (auto& e : c->ev) CUB(cudaEventCreate(&e));
    CUB(cudaEventCreateWithFlags(&c->evSync, cudaEventDisableTiming));
    for (auto& e : c->evWave) CUB(cudaEventCreateWithFlags(&e, cudaEventDisableTiming | cudaEventBlockingSync));
    {   // fp16 operand images + exact row table of the fp16 tensor-core kernel, and its rigorous pre-filter bound:
        //   |p~ - p_oracle| <= ||d - fp16(d)|| ||r||  (dictionary rounding, Cauchy-Schwarz)
        //                    + ||d|| ||r|| (2^-11 residual rounded to fp16 + 4e-5 tensor-core accumulation + 64 * 2^-24 oracle rounding)
        //                    + 5e-7 (fp16 subnormal residual parts), 15 % margin on the relative part
        std::vector<uint16_t> b16(static_cast<size_t>(kRowsPad) * kN), d16(static_cast<size_t>(3) * n_base * 64 * kN);
        std::vector<float> rows(f16Rows32Count(n_base) * kN);
        const float* det3[3] = {detailY, detailU, detailV};
        double roundNorm = 0.0, maxNorm = 0.0;
        buildF16Images(base, n_base, detail_rows, det3, b16.data(), d16.data(), rows.data(), &roundNorm, &maxNorm);
        // residual rounded to fp16: |d.(r - fp16(r))| <= ||d|| 2^-11 ||r|| (subnormal parts: epsAbs16)
        c->epsRel16Hi = static_cast<float>((roundNorm + maxNorm * (1.0 / 2048.0 + 4e-5 + 64.0 / 16777216.0)) * 1.15);
        c->epsAbs16 = 5e-7f;
        CUB(c->baseUm16.ensure(b16.size() * 2));
        CUB(cudaMemcpy(c->baseUm16.p, b16.data(), b16.size() * 2, cudaMemcpyHostToDevice));
        CUB(c->detUm16.ensure(d16.size() * 2));
        CUB(cudaMemcpy(c->detUm16.p, d16.data(), d16.size() * 2, cudaMemcpyHostToDevice));
        CUB(c->rows32.ensure(rows.size() * 4));
        CUB(cudaMemcpy(c->rows32.p, rows.data(), rows.size() * 4, cudaMemcpyHostToDevice));
    }
    {
        int threads = static_cast<int>(std::thread::hardware_concurrency());
        cpu_set_t set;
        if (sched_getaffinity(0, sizeof(set), &set) == 0) threads = CPU_COUNT(&set);
        if (const char* lw = std::getenv("LOCAL_WORLD_SIZE"))  // one process per GPU (torchrun): share the host cores between the ranks
            if (std::atoi(lw) > 1) threads = std::max(1, threads / std::atoi(lw));
        if (const char* t = std::getenv("MNENC_HOST_THREADS")) threads = std::atoi(t);
        c->hostThreads = std::clamp(threads, 1, 16);
        c->pool = std::make_unique<HostPool>(c->hostThreads);
    }
    std::vector<uint8_t> rows(kRowsPad, 0);
    for (int a = 0; a < n_base; ++a) rows[a] = static_cast<uint8_t>(detail_rows[a]);
    CUB(c->detRows.ensure(kRowsPad));
    CUB(cudaMemcpy(c->detRows.p, rows.data(), kRowsPad, cudaMemcpyHostToDevice));
    CUB(c->quant.ensure(3 * kMaxK * 4));
#undef CUB
    *out = c;
    return MNENC_OK;
} catch (...) {
    return mnenc::abiCaught();
}

void mnenc_destroy(mnenc_t* c) {
    if (!c) return;
    for (mnenc_ctx* peer : c->groupPeers) mnenc_destroy(peer);
    c->groupPeers.clear();
    for (mnenc_ctx* lane : c->lanes) mnenc_destroy(lane);
    c->lanes.clear();
    cudaSetDevice(c->device);
    if (c->stream) cudaStreamSynchronize(c->stream);
    for (auto& e : c->ev)
        if (e) cudaEventDestroy(e);
    if (c->evSync) cudaEventDestroy(c->evSync);
    for (auto& e : c->evChunk)
        if (e) cudaEventDestroy(e);
    if (c->copyStream) cudaStreamDestroy(c->copyStream);
    if (c->stripeResult) resultRelease(c->stripeResult);
    for (auto& e : c->evWave)
        if (e) cudaEventDestroy(e);
    if (c->stream) cudaStreamDestroy(c->stream);
    delete c;
}

int mnenc_set_quant(mnenc_t* c, const float* q) try {
    if (!c || !q) return fail(MNENC_ERR_ARG, "null argument");
    for (int i = 0; i < 3 * c->K; ++i)
        if (!(q[i] >= 1.0f) || q[i] > 65535.0f) return fail(MNENC_ERR_ARG, "quantisers must be in [1,65535]");
    std::memset(c->hQuant, 0, sizeof(c->hQuant));
    for (int ch = 0; ch < 3; ++ch)
        for (int i = 0; i < mnenc::kMaxK; ++i) c->hQuant[ch * mnenc::kMaxK + i] = i < c->K ? q[ch * c->K + i] : 1.0f;
    CU(cudaSetDevice(c->device));
    CU(cudaMemcpyAsync(c->quant.p, c->hQuant, sizeof(c->hQuant), cudaMemcpyHostToDevice, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    c->quantSet = true;
    for (mnenc_ctx* peer : c->groupPeers)
        if (int rc = mnenc_set_quant(peer, q)) return rc;
    return MNENC_OK;
} catch (...) {
    return mnenc::abiCaught();
}

static int encodeDevice(mnenc_t* c, const uint8_t* rgbDev, size_t w, size_t h, size_t stride, uint8_t** out, size_t* outBytes,
                        double t0) {
    if (int rc = runMatchingPursuit(c, rgbDev, w, h, stride)) return rc;
    if (int rc = runBackend(c, w, h, out, outBytes)) return rc;
    finishMpAccounting(c);
    c->ms[7] = static_cast<float>(nowMs() - t0);
    return MNENC_OK;
}

}  // extern "C"

// n frames of one geometry (batch mode): ONE matching pursuit over all their block-channels - small frames alone leave the
// per-iteration kernels launch-bound - then the back end frame by frame on the frames' slices of the symbol arrays. Every stream
// is the one the single-image call writes.
int mnenc::encodeFrames(mnenc_ctx* c, const uint8_t* const* rgb, int onDevice, size_t n, size_t w, size_t h, const size_t* strides,
                        uint8_t** out, size_t* outBytes) {
    if (n == 0) return MNENC_OK;
    CU(cudaSetDevice(c->device));
    const double t0 = nowMs();
    c->launches = 0;
    c->ms[0] = 0;
    CU(c->hFrameTab.ensure(n * sizeof(void*)));
    CU(c->frameTab.ensure(n * sizeof(void*)));
    const uint8_t** tab = c->hFrameTab.as<const uint8_t*>();
    size_t stride = 3 * w;
    if (onDevice) {
        stride = strides[0];
        for (size_t i = 0; i < n; ++i) tab[i] = rgb[i];
    } else {
        const size_t frameBytes = (3 * w * h + 15) / 16 * 16;
        CU(c->rgb.ensure(frameBytes * n + 16));
        for (size_t i = 0; i < n; ++i) {
            uint8_t* d = c->rgb.as<uint8_t>() + frameBytes * i;
            CU(cudaMemcpy2DAsync(d, 3 * w, rgb[i], strides[i], 3 * w, h, cudaMemcpyHostToDevice, c->stream));
            tab[i] = d;
        }
    }
    CU(cudaMemcpyAsync(c->frameTab.p, tab, n * sizeof(void*), cudaMemcpyHostToDevice, c->stream));
    if (int rc = runMatchingPursuit(c, nullptr, w, h, stride, c->frameTab.as<const uint8_t*>(), static_cast<uint32_t>(n))) return rc;
    const uint32_t slots = frameBlockSlots(w, h);
    int rcAll = MNENC_OK;
    for (size_t i = 0; i < n && rcAll == MNENC_OK; ++i) {
        c->frameBc = 3u * slots * static_cast<uint32_t>(i);
        rcAll = runBackend(c, w, h, &out[i], &outBytes[i]);
    }
    c->frameBc = 0;
    if (rcAll) return rcAll;
    finishMpAccounting(c);
    c->ms[7] = static_cast<float>(nowMs() - t0);
    return MNENC_OK;
}

extern "C" {

int mnenc_encode_rgb8_device(mnenc_t* c, const uint8_t* rgb_dev, size_t w, size_t h, size_t stride, uint8_t** out, size_t* outBytes) try {
    if (int rc = checkEncodeArgs(c, rgb_dev, w, h, stride, out, outBytes)) return rc;
    CU(cudaSetDevice(c->device));
    const double t0 = nowMs();
    c->launches = 0;
    c->ms[0] = 0;
    return encodeDevice(c, rgb_dev, w, h, stride, out, outBytes, t0);
} catch (...) {
    return mnenc::abiCaught();
}

int mnenc_encode_rgb8(mnenc_t* c, const uint8_t* rgb, size_t w, size_t h, size_t stride, uint8_t** out, size_t* outBytes) try {
    if (int rc = checkEncodeArgs(c, rgb, w, h, stride, out, outBytes)) return rc;
    if (!c->groupPeers.empty()) return encodeGroup(c, rgb, w, h, stride, out, outBytes);  // mnenc_create_multi: block-column stripes over the GPUs
    CU(cudaSetDevice(c->device));
    const double t0 = nowMs();
    c->launches = 0;
    CU(c->rgb.ensure(3 * w * h + 16));
    // tightly packed on the device; pinned or pageable host memory both work (pinned is the fast path)
    CU(cudaMemcpy2DAsync(c->rgb.p, 3 * w, rgb, stride, 3 * w, h, cudaMemcpyHostToDevice, c->stream));
    c->ms[0] = 0;
    int rc = encodeDevice(c, c->rgb.as<uint8_t>(), w, h, 3 * w, out, outBytes, t0);
    return rc;
} catch (...) {
    return mnenc::abiCaught();
}

int mnenc_get_symbols(mnenc_t* c, uint16_t* counts, uint16_t* pairs) try {
    if (!c || !counts || !pairs) return fail(MNENC_ERR_ARG, "null argument");
    if (!c->haveSymbols) return fail(MNENC_ERR_STATE, "no encode has run on this context");
    CU(cudaSetDevice(c->device));
    const size_t nbc = 3 * ((c->lastW + 7) / 8) * ((c->lastH + 7) / 8);
    CU(cudaMemcpyAsync(counts, c->counts.p, nbc * 2, cudaMemcpyDeviceToHost, c->stream));
    CU(cudaMemcpyAsync(pairs, c->pairs.p, nbc * c->K * 4, cudaMemcpyDeviceToHost, c->stream));  // u32 = {deltaId, intCoeff} little-endian
    CU(cudaStreamSynchronize(c->stream));
    // the kernels write only the `count` pairs of a block-channel (the back end reads no others): 0 beyond, as documented
    for (size_t bc = 0; bc < nbc; ++bc)
        for (int i = counts[bc]; i < c->K; ++i) pairs[(bc * c->K + i) * 2] = pairs[(bc * c->K + i) * 2 + 1] = 0;
    return MNENC_OK;
} catch (...) {
    return mnenc::abiCaught();
}

int mnenc_pack_symbols(mnenc_t* c, size_t w, size_t h, const uint16_t* counts, const uint16_t* pairs, uint8_t** out, size_t* outBytes) try {
    if (!c || !counts || !pairs || !out || !outBytes) return fail(MNENC_ERR_ARG, "null argument");
    if (w == 0 || h == 0) return fail(MNENC_ERR_ARG, "bad image geometry");
    if (!c->quantSet) return fail(MNENC_ERR_STATE, "mnenc_set_quant has not been called");
    CU(cudaSetDevice(c->device));
    const double t0 = nowMs();
    c->launches = 0;
    const size_t nbc = 3 * ((w + 7) / 8) * ((h + 7) / 8);
    for (size_t i = 0; i < nbc; ++i)
        if (counts[i] > c->K) return fail(MNENC_ERR_ARG, "count exceeds K");
    CU(c->counts.ensure(nbc * 2 + 16));
    CU(c->pairs.ensure(nbc * c->K * 4 + 16));
    CU(cudaMemcpyAsync(c->counts.p, counts, nbc * 2, cudaMemcpyHostToDevice, c->stream));
    CU(cudaMemcpyAsync(c->pairs.p, pairs, nbc * c->K * 4, cudaMemcpyHostToDevice, c->stream));
    c->lastW = w;
    c->lastH = h;
    c->haveSymbols = true;
    c->ms[0] = c->ms[1] = 0;
    c->mpFlops = 0;
    int rc = runBackend(c, w, h, out, outBytes);
    c->ms[7] = static_cast<float>(nowMs() - t0);
    return rc;
} catch (...) {
    return mnenc::abiCaught();
}

int mnenc_last_timing(mnenc_t* c, float* ms8) try {
    if (!c || !ms8) return fail(MNENC_ERR_ARG, "null argument");
    std::memcpy(ms8, c->ms, sizeof(c->ms));
    return c->launches;
} catch (...) {
    return mnenc::abiCaught();
}

int mnenc_last_work(mnenc_t* c, double* mp_flops, uint64_t* n_symbols, uint64_t* n_bits) try {
    if (!c) return fail(MNENC_ERR_ARG, "null argument");
    if (mp_flops) *mp_flops = c->mpFlops;
    if (n_symbols) *n_symbols = c->nSymbols;
    if (n_bits) *n_bits = c->nBits;
    return MNENC_OK;
} catch (...) {
    return mnenc::abiCaught();
}

int mnenc_last_plan(mnenc_t* c, int* n_huffman, int* n_golomb, int* n_rle) try {
    if (!c) return fail(MNENC_ERR_ARG, "null argument");
    if (n_huffman) *n_huffman = c->planHuffman;
    if (n_golomb) *n_golomb = c->planGolomb;
    if (n_rle) *n_rle = c->planRle;
    return MNENC_OK;
} catch (...) {
    return mnenc::abiCaught();
}

int mnenc_debug_plan_sequence(const uint32_t* symbols, const uint32_t* freqs, const uint32_t* first_pos, size_t n, uint32_t* golomb_m,
                              uint32_t* codes, uint8_t* lens, uint32_t* eof_code, uint8_t* eof_len, uint8_t* header_bytes,
                              size_t header_cap, uint64_t* header_bits, uint64_t* payload_bits) try {
    if (!golomb_m || !eof_code || !eof_len || !header_bits || !payload_bits || (n && (!symbols || !freqs || !first_pos || !codes || !lens)))
        return fail(MNENC_ERR_ARG, "null argument");
    std::vector<mnenc::SymbolStat> stats(n);
    for (size_t i = 0; i < n; ++i) stats[i] = mnenc::SymbolStat{symbols[i], freqs[i], first_pos[i]};
    mnenc::SequencePlan plan;
    mnenc::planSequence(stats, plan);
    *golomb_m = plan.golomb ? plan.golombM : 0;
    for (size_t i = 0; i < n; ++i) {
        codes[i] = plan.code[i];
        lens[i] = plan.len[i];
    }
    *eof_code = plan.eofCode;
    *eof_len = plan.eofLen;
    *header_bits = plan.header.nbits;
    *payload_bits = plan.payloadBits;
    if (plan.header.bytes.size() > header_cap) return fail(MNENC_ERR_ARG, "header buffer too small");
    if (header_bytes) std::memcpy(header_bytes, plan.header.bytes.data(), plan.header.bytes.size());
    return MNENC_OK;
} catch (...) {
    return mnenc::abiCaught();
}

int mnenc_debug_fail_next_backend(mnenc_t* c) try {
    if (!c) return fail(MNENC_ERR_ARG, "null argument");
    c->failNextBackend = true;
    return MNENC_OK;
} catch (...) {
    return mnenc::abiCaught();
}

int mnenc_debug_stitch_runs(const uint32_t* records, int nranks, int n_streams, int rank, uint32_t* carry_in, uint32_t* extend_out) try {
    if (!records || !carry_in || !extend_out || nranks < 1 || n_streams < 0 || rank < 0 || rank >= nranks)
        return fail(MNENC_ERR_ARG, "bad argument");
    static_assert(sizeof(mnenc::StripeBoundary) == 6 * sizeof(uint32_t), "record layout");
    mnenc::stitchRuns(reinterpret_cast<const mnenc::StripeBoundary*>(records), nranks, n_streams, rank, carry_in, extend_out);
    return MNENC_OK;
} catch (...) {
    return mnenc::abiCaught();
}

uint32_t mnenc_debug_rle_chunk_size(uint32_t carry, uint32_t len, uint64_t total) { return mnenc::rleChunkedSize(carry, len, total); }

}  // extern "C"
