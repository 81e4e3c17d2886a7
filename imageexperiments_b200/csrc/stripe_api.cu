// Stripe protocol: ONE image sharded by block columns over the ranks of a process group (BASELINE.json
// configs[3]). The reference walks blocks column-major (CompressedImage.cpp:592-594), so rank r's symbols are the r-th
// contiguous slice of every stream. Matching pursuit needs no exchange; the entropy back end needs five:
//
//   begin      H2D of the stripe, matching pursuit, raw stream scatter           -> rec1: stream sizes, last raw DC value
//     [all-gather rec1]
//   rle        DC delta chain continued from the previous rank, run detection      -> rec2: boundary run records
//     [all-gather rec2]
//   histogram  run stitching + RLE-or-raw decision on the WHOLE stream (host, identical on every rank), stitched RLE,
//              local histogram with GLOBAL first-occurrence positions              -> device histogram buffers
//     [all-reduce: SUM histogram, MIN first occurrence, MAX largest symbol]  (NCCL over NVLink)
//   plan       identical code tables on every rank (host), payload bits of the rank's slices -> seq_bits
//     [all-gather seq_bits]   (offset scan)
//   pack       every rank packs its slices at their final bit positions of the file -> device words, disjoint bits
//     [reduce SUM to rank 0]  (disjoint bits: the sum is the concatenation of the per-GPU bitstream segments)
//   finish     rank 0: D2H + headers/EOF codes written by the host
//
// The collectives themselves are the caller's (torch.distributed / NCCL in imageexperiments_b200/stripes.py): this file
// is plain C ABI in, plain C ABI out.
#include <algorithm>
#include <cstring>
#include <numeric>

#include "mnenc_ctx.cuh"

namespace mnenc {

struct StripeState {
    int phase = 0;  // last completed phase: 1 begin, 2 rle, 3 histogram, 4 plan, 5 pack
    int rank = 0, nranks = 1;
    size_t W = 0, H = 0;
    uint32_t bx0 = 0, bx1 = 0, nby = 0, nblocks = 0;
    StreamOffsets so{};
    uint32_t N = 0, M = 0, nEntries = 0;
    std::vector<uint32_t> totals;
    BackendBuffers b{};
    StripePlan plan;
    SeqTable seq{};
    std::vector<uint64_t> codedLenGlobal;  // [nSeq] length of the whole coded sequence (RLE size when RLE won)
    std::vector<SequencePlan> plans;
    std::vector<std::vector<uint32_t>> entryOf;
    std::vector<Overlay> overlays;
    size_t totalBytes = 0, outWords = 0;
    uint64_t payloadBits = 0;
    DevBuf recDev, stitchDev, seqBitsDev, scratch;
    PinBuf hRec, hStitch, hSeqBits, hScratch;
};

namespace {

void destroyStripe(StripeState* s) { delete s; }

StripeState* stateOf(mnenc_ctx* c) {
    if (!c->stripe) c->stripe = std::unique_ptr<StripeState, void (*)(StripeState*)>(new StripeState, destroyStripe);
    return c->stripe.get();
}

int needPhase(mnenc_ctx* c, int phase) {
    if (!c) return fail(MNENC_ERR_ARG, "null context");
    if (!c->stripe || c->stripe->phase != phase) return fail(MNENC_ERR_STATE, "stripe protocol called out of order");
    return MNENC_OK;
}

}  // namespace

int stripeExchangeHistogram(mnenc_ctx* c, int localRc) {
    if (!localRc) localRc = needPhase(c, 3);
    return exchangeHistogramSparse(c, c->stripe ? &c->stripe->b : nullptr, 6 * c->K + 1, localRc);
}

}  // namespace mnenc

using namespace mnenc;

extern "C" {

int mnenc_stripe_begin(mnenc_t* c, const uint8_t* rgb_stripe, int on_device, size_t width, size_t height, size_t row_stride_bytes, int rank,
                       int nranks, uint32_t bx0, uint32_t bx1, uint32_t* rec1) try {
    if (!c || !rec1) return fail(MNENC_ERR_ARG, "null argument");
    if (width == 0 || height == 0 || width > 0xFFFFFFFFull || height > 0xFFFFFFFFull) return fail(MNENC_ERR_ARG, "bad image geometry");
    if (!c->quantSet) return fail(MNENC_ERR_STATE, "mnenc_set_quant has not been called");
    if (c->stripeResult) resultRelease(c->stripeResult), c->stripeResult = nullptr;  // left over from a striped encode that failed half way
    const uint32_t nbxAll = static_cast<uint32_t>((width + 7) / 8), nby = static_cast<uint32_t>((height + 7) / 8);
    if (nranks < 1 || rank < 0 || rank >= nranks || bx0 > bx1 || bx1 > nbxAll) return fail(MNENC_ERR_ARG, "bad stripe");
    if (bx1 > bx0 && (!rgb_stripe || row_stride_bytes < 3)) return fail(MNENC_ERR_ARG, "null pixels");
    CU(cudaSetDevice(c->device));
    StripeState* s = stateOf(c);
    s->phase = 0;
    s->rank = rank, s->nranks = nranks, s->W = width, s->H = height, s->bx0 = bx0, s->bx1 = bx1, s->nby = nby;
    s->nblocks = (bx1 - bx0) * nby;
    const int K = c->K;
    const double t0 = nowMs();
    c->launches = 0;
    std::memset(c->ms, 0, sizeof(c->ms));
    c->mpFlops = 0;
    cudaStream_t st = c->stream;

    // ---- pixels of the stripe + matching pursuit ----
    if (s->nblocks) {
        const size_t x0 = static_cast<size_t>(bx0) * 8, wS = std::min<size_t>(static_cast<size_t>(bx1) * 8, width) - x0;
        if (row_stride_bytes < 3 * wS) return fail(MNENC_ERR_ARG, "row stride smaller than the stripe");
        const uint8_t* dev = rgb_stripe;
        size_t stride = row_stride_bytes;
        if (!on_device) {
            CU(c->rgb.ensure(3 * wS * height + 16));
            CU(cudaMemcpy2DAsync(c->rgb.p, 3 * wS, rgb_stripe, row_stride_bytes, 3 * wS, height, cudaMemcpyHostToDevice, st));
            dev = c->rgb.as<uint8_t>();
            stride = 3 * wS;
        }
        if (int rc = runMatchingPursuit(c, dev, wS, height, stride)) return rc;
    } else {
        CU(c->counts.ensure(16));
        CU(c->pairs.ensure(16));
        c->lastW = 0, c->lastH = height, c->haveSymbols = true;
    }

    // ---- stream sizes of the slice ----
    BackendBuffers& b = s->b;
    b = BackendBuffers{};
    b.counts = c->counts.as<uint16_t>();
    b.pairs = c->pairs.as<uint32_t>();
    b.K = K;
    b.nblocks = s->nblocks;
    b.ntiles = (s->nblocks + kTileBlocks - 1) / kTileBlocks;
    CU(c->tileCounts.ensure(static_cast<size_t>(b.ntiles) * 3 * K * 4 + 4));
    CU(c->totals.ensure(3 * K * 4));
    CU(c->dcTmp.ensure(static_cast<size_t>(s->nblocks) * 3 * 2 + 2));
    CU(c->hTotals.ensure(3 * K * 4));
    b.tileCounts = c->tileCounts.as<uint32_t>();
    b.totals = c->totals.as<uint32_t>();
    b.dcTmp = c->dcTmp.as<uint16_t>();
    CU(cudaEventRecord(c->ev[2], st));
    if (b.ntiles) c->launches += launchTileCounts(b, st);
    else CU(cudaMemsetAsync(b.totals, 0, 3 * K * 4, st));
    CU(cudaMemcpyAsync(c->hTotals.p, b.totals, 3 * K * 4, cudaMemcpyDeviceToHost, st));
    CU(waitStream(c));
    if (s->nblocks) finishMpAccounting(c);
    s->totals.assign(c->hTotals.as<uint32_t>(), c->hTotals.as<uint32_t>() + 3 * K);
    uint64_t N64 = 0;
    for (int i = 0; i < 3 * K; ++i) {
        s->so.off[2 * i] = static_cast<uint32_t>(N64);
        s->so.off[2 * i + 1] = static_cast<uint32_t>(N64 + s->totals[i]);
        N64 += 2ull * s->totals[i];
    }
    if (N64 + 3ull * s->nblocks >= 0x7FFFFFF0ull) return fail(MNENC_ERR_ARG, "stripe too large: more than 2^31 symbols");
    const uint32_t N = s->N = static_cast<uint32_t>(N64);
    s->so.off[6 * K] = N;

    // ---- raw scatter; the last raw DC value of each channel goes to the next rank ----
    CU(c->arena.ensure(static_cast<size_t>(N) * 2 + 16));
    CU(c->flags.ensure(static_cast<size_t>(N) + 16));
    CU(c->runIdx.ensure((static_cast<size_t>(N) + 2) * 4));
    CU(c->startPos.ensure((static_cast<size_t>(N) + 2) * 4));
    CU(c->outOff.ensure((static_cast<size_t>(N) + 2) * 4));
    CU(c->rleArena.ensure((static_cast<size_t>(N) + N / 2 + 16 + 6ull * 6 * K) * 2));
    const uint32_t upperM = N + 3 * s->nblocks;
    CU(c->sums32.ensure((static_cast<size_t>(upperM) / 2048 + 4) * 4));
    CU(c->sums64.ensure((static_cast<size_t>(upperM) / 2048 + 4) * 8));
    CU(c->seqTable.ensure(sizeof(SeqTable)));
    b.arena = c->arena.as<uint16_t>();
    b.flags = c->flags.as<uint8_t>();
    b.runIdx = c->runIdx.as<uint32_t>();
    b.startPos = c->startPos.as<uint32_t>();
    b.outOff = c->outOff.as<uint32_t>();
    b.rleArena = c->rleArena.as<uint16_t>();
    b.blockSums32 = c->sums32.as<uint32_t>();
    b.blockSums64 = c->sums64.as<unsigned long long>();
    b.seqTable = c->seqTable.as<SeqTable>();
    CU(s->hScratch.ensure(64));
    uint16_t* hLast = s->hScratch.as<uint16_t>();
    hLast[0] = hLast[1] = hLast[2] = 0;
    if (b.ntiles) {
        c->launches += launchScatterRaw(b, s->so, st);
        for (int ch = 0; ch < 3; ++ch)
            if (const uint32_t n = s->totals[ch * K])
                CU(cudaMemcpyAsync(hLast + ch, b.arena + s->so.off[2 * ch * K + 1] + n - 1, 2, cudaMemcpyDeviceToHost, st));
        CU(waitStream(c));
    }
    for (int i = 0; i < 3 * K; ++i) rec1[i] = s->totals[i];
    for (int ch = 0; ch < 3; ++ch) rec1[3 * K + ch] = hLast[ch];
    c->ms[7] = static_cast<float>(nowMs() - t0);
    s->phase = 1;
    return MNENC_OK;
} catch (...) {
    return mnenc::abiCaught();
}

int mnenc_stripe_rle(mnenc_t* c, const uint32_t* all_rec1, uint32_t* rec2) try {
    if (int rc = needPhase(c, 1)) return rc;
    if (!all_rec1 || !rec2) return fail(MNENC_ERR_ARG, "null argument");
    CU(cudaSetDevice(c->device));
    StripeState* s = c->stripe.get();
    const int K = c->K, nStreams = 6 * K, rl = 3 * K + 3;
    const double t0 = nowMs();
    cudaStream_t st = c->stream;
    DcPrev prev{};
    for (int ch = 0; ch < 3; ++ch)
        for (int q = s->rank - 1; q >= 0; --q)
            if (all_rec1[static_cast<size_t>(q) * rl + ch * K]) {  // nearest earlier rank that coded a block of this channel
                prev.v[ch] = all_rec1[static_cast<size_t>(q) * rl + 3 * K + ch];
                break;
            }
    CU(s->recDev.ensure(static_cast<size_t>(nStreams) * sizeof(StripeBoundary)));
    CU(s->hRec.ensure(static_cast<size_t>(nStreams) * sizeof(StripeBoundary)));
    if (s->nblocks) c->launches += launchDcDelta(s->b, s->so, prev, st);
    c->launches += launchRleRuns(s->b, s->so, s->N, st);
    c->launches += launchStripeRecords(s->b, s->so, s->N, s->recDev.as<StripeBoundary>(), st);
    CU(cudaMemcpyAsync(s->hRec.p, s->recDev.p, static_cast<size_t>(nStreams) * sizeof(StripeBoundary), cudaMemcpyDeviceToHost, st));
    CU(waitStream(c));
    std::memcpy(rec2, s->hRec.p, static_cast<size_t>(nStreams) * sizeof(StripeBoundary));
    c->ms[7] += static_cast<float>(nowMs() - t0);
    s->phase = 2;
    return MNENC_OK;
} catch (...) {
    return mnenc::abiCaught();
}

int mnenc_stripe_histogram(mnenc_t* c, const uint32_t* all_rec2, void** hist_dev, void** first_pos_dev, size_t* hist_elems,
                           void** max_sym_dev, size_t* max_sym_elems) try {
    if (int rc = needPhase(c, 2)) return rc;
    if (!all_rec2 || !hist_dev || !first_pos_dev || !hist_elems || !max_sym_dev || !max_sym_elems) return fail(MNENC_ERR_ARG, "null argument");
    CU(cudaSetDevice(c->device));
    StripeState* s = c->stripe.get();
    const int K = c->K, nStreams = 6 * K, nSeq = nStreams + 1, me = s->rank;
    const double t0 = nowMs();
    cudaStream_t st = c->stream;
    const StripeBoundary* rec = reinterpret_cast<const StripeBoundary*>(all_rec2);
    for (int q = 0; q < nStreams; ++q)
        if (rec[static_cast<size_t>(me) * nStreams + q].n != s->so.off[q + 1] - s->so.off[q]) return fail(MNENC_ERR_ARG, "records do not match this rank's slices");
    planStripes(rec, s->nranks, nStreams, s->plan);

    // ---- stitched RLE of the slices ----
    CU(s->stitchDev.ensure(sizeof(StripeStitch)));
    CU(s->hStitch.ensure(sizeof(StripeStitch)));
    StripeStitch& hs = *s->hStitch.as<StripeStitch>();
    std::memset(&hs, 0, sizeof(hs));
    std::memcpy(hs.off, s->so.off, sizeof(uint32_t) * (nStreams + 1));
    for (int q = 0; q < nStreams; ++q) {
        hs.carry[q] = s->plan.carry[static_cast<size_t>(me) * nStreams + q];
        hs.ext[q] = s->plan.ext[static_cast<size_t>(me) * nStreams + q];
    }
    CU(cudaMemcpyAsync(s->stitchDev.p, &hs, sizeof(hs), cudaMemcpyHostToDevice, st));
    c->launches += launchRleWrite(s->b, s->N, s->stitchDev.as<StripeStitch>(), st);

    // ---- table of coded sequences: local slice + its position in the whole sequence ----
    SeqTable& t = s->seq;
    std::memset(&t, 0, sizeof(t));
    s->codedLenGlobal.assign(nSeq, 0);
    const uint64_t nbxAll = (s->W + 7) / 8;
    t.desc[0] = SeqDesc{3 * s->nblocks, 3 * s->nblocks, 0u, 0u};
    t.posBase[0] = static_cast<uint32_t>(3ull * s->bx0 * s->nby);
    s->codedLenGlobal[0] = 3ull * nbxAll * s->nby;
    uint64_t rleOff = 0, posTotal = s->codedLenGlobal[0];
    for (int q = 0; q < nStreams; ++q) {
        const bool rle = s->plan.useRle[q];
        uint64_t before = 0;
        for (int r = 0; r < me; ++r) before += rle ? s->plan.emitted[static_cast<size_t>(r) * nStreams + q] : rec[static_cast<size_t>(r) * nStreams + q].n;
        const uint32_t emitted = s->plan.emitted[static_cast<size_t>(me) * nStreams + q];
        const uint32_t raw = s->so.off[q + 1] - s->so.off[q];
        t.desc[1 + q] = rle ? SeqDesc{emitted, raw, static_cast<uint32_t>(rleOff), 2u} : SeqDesc{raw, raw, s->so.off[q], 1u};
        t.posBase[1 + q] = static_cast<uint32_t>(before);
        s->codedLenGlobal[1 + q] = rle ? s->plan.rleLen[q] : s->plan.rawLen[q];
        posTotal = std::max(posTotal, s->codedLenGlobal[1 + q]);
        rleOff += emitted;
    }
    if (posTotal >= kNoPos) return fail(MNENC_ERR_ARG, "image too large: a coded sequence has more than 2^31 symbols");
    if (rleOff > static_cast<uint64_t>(s->N) + s->N / 2 + 6ull * nStreams) return fail(MNENC_ERR_INTERNAL, "stitched RLE larger than its arena");
    uint32_t acc = 0;
    for (int q = 0; q < nSeq; ++q) {
        t.start[q] = acc;
        acc += t.desc[q].len;
    }
    t.start[nSeq] = acc;
    s->M = acc;
    CU(c->hSeqTable.ensure(sizeof(SeqTable)));
    std::memcpy(c->hSeqTable.p, &t, sizeof(t));
    CU(cudaMemcpyAsync(s->b.seqTable, c->hSeqTable.p, sizeof(SeqTable), cudaMemcpyHostToDevice, st));

    // ---- local histogram, global first-occurrence positions ----
    if (c->entryCap < (1u << 20)) {  // a failed compaction cannot be redone after the all-reduce: start large
        c->entryCap = 1u << 20;
        CU(c->entries.ensure(static_cast<size_t>(c->entryCap) * sizeof(HistEntry)));
    }
    if (int rc = ensureHistogram(c)) return rc;
    BackendBuffers& b = s->b;
    b.hist = c->hist.as<uint32_t>();
    b.firstPos = c->firstPos.as<uint32_t>();
    b.maxSym = c->maxSym.as<uint32_t>();
    b.entries = c->entries.as<HistEntry>();
    b.nEntries = c->nEntries.as<uint32_t>();
    b.entryCap = c->entryCap;
    CU(cudaEventRecord(c->ev[3], st));
    c->histDirty = true;  // until mnenc_stripe_pack has reset the touched bins
    c->launches += launchHistogramOnly(b, s->M, nSeq, st);
    CU(cudaEventRecord(c->ev[4], st));
    if (c->failNextBackend) {  // test hook (mnenc_debug_fail_next_backend): an ordinary error return on THIS rank only, bins populated
        c->failNextBackend = false;
        CU(waitStream(c));
        return fail(MNENC_ERR_INTERNAL, "injected failure after the histogram (test hook)");
    }
    // consistency: the device's stitched RLE size must be what every rank derived from the records
    CU(s->scratch.ensure(64));
    c->launches += launchReadIndirect(b.outOff, b.runIdx + s->N, s->scratch.as<uint32_t>(), st);
    CU(cudaMemcpyAsync(s->hScratch.p, s->scratch.p, 4, cudaMemcpyDeviceToHost, st));
    CU(waitStream(c));
    if (*s->hScratch.as<uint32_t>() != static_cast<uint32_t>(rleOff)) return fail(MNENC_ERR_INTERNAL, "stitched RLE size differs from the plan");
    *hist_dev = c->hist.p;
    *first_pos_dev = c->firstPos.p;
    *hist_elems = static_cast<size_t>(kMaxSeq) * kHistBins;
    *max_sym_dev = c->maxSym.p;
    *max_sym_elems = kMaxSeq;
    c->ms[7] += static_cast<float>(nowMs() - t0);
    s->phase = 3;
    return MNENC_OK;
} catch (...) {
    return mnenc::abiCaught();
}

int mnenc_stripe_plan(mnenc_t* c, uint64_t* seq_bits) try {
    if (int rc = needPhase(c, 3)) return rc;
    if (!seq_bits) return fail(MNENC_ERR_ARG, "null argument");
    CU(cudaSetDevice(c->device));
    StripeState* s = c->stripe.get();
    const int K = c->K, nSeq = 6 * K + 1;
    const double t0 = nowMs();
    cudaStream_t st = c->stream;
    BackendBuffers& b = s->b;
    // In-library exchanges: a failure of the local steps up to the table builds must not keep this rank out of planSharded's
    // collectives (its peers would wait there forever) - it joins them with its error code instead, and every rank stops.
    const bool dealt = c->stripeInLibrary && stripeRanks(c) > 1;
    std::vector<std::vector<SymbolStat>> stats(nSeq);
    uint32_t nEntries = 0;
    double tHost0 = nowMs();
    const int prepRc = [&]() -> int {
        try {
            c->launches += launchCompact(b, nSeq, st);
            CU(c->hNEntries.ensure(4));
            CU(cudaMemcpyAsync(c->hNEntries.p, b.nEntries, 4, cudaMemcpyDeviceToHost, st));
            CU(waitStream(c));
            nEntries = s->nEntries = *c->hNEntries.as<uint32_t>();
            if (nEntries > c->entryCap) {
                c->histReady = false;  // bins were consumed; rebuild the clean state on the next encode
                return fail(MNENC_ERR_INTERNAL, "more than 2^20 distinct (sequence, symbol) pairs");
            }
            CU(c->hEntries.ensure(static_cast<size_t>(nEntries) * sizeof(HistEntry) + 16));
            CU(cudaMemcpyAsync(c->hEntries.p, b.entries, static_cast<size_t>(nEntries) * sizeof(HistEntry), cudaMemcpyDeviceToHost, st));
            CU(waitStream(c));

            // ---- identical code tables on every rank ----
            tHost0 = nowMs();
            const HistEntry* ent = c->hEntries.as<HistEntry>();
            s->entryOf.assign(nSeq, {});
            for (uint32_t e = 0; e < nEntries; ++e) {
                if (ent[e].seq >= static_cast<uint32_t>(nSeq)) return fail(MNENC_ERR_INTERNAL, "bad histogram entry");
                stats[ent[e].seq].push_back(SymbolStat{ent[e].symbol, ent[e].freq, ent[e].firstPos});
                s->entryOf[ent[e].seq].push_back(e);
            }
            s->plans.assign(nSeq, SequencePlan{});
            return MNENC_OK;
        } catch (...) {
            return abiCaught();
        }
    }();
    if (prepRc && !dealt) return prepRc;
    if (dealt) {  // in-library exchanges only (the caller-driven protocol may run on a context that also owns a communicator, with
                  // other ranks): table builds dealt over the ranks, all-gathered
        if (int rc = planSharded(c, stats, s->plans, prepRc)) return rc;
    } else {
        std::vector<int> byLoad(nSeq);
        std::iota(byLoad.begin(), byLoad.end(), 0);
        std::sort(byLoad.begin(), byLoad.end(), [&](int a, int b2) { return stats[a].size() > stats[b2].size(); });
        c->pool->parallelFor(nSeq, [&](int i) { planSequence(stats[byLoad[i]], s->plans[byLoad[i]]); });
    }
    CU(c->hEntryCodes.ensure(static_cast<size_t>(nEntries) * 8 + 8));
    CU(c->hCodeTable.ensure(sizeof(CodeTable)));
    CU(c->entryCodes.ensure(static_cast<size_t>(nEntries) * 8 + 8));
    CU(c->codeTable.ensure(sizeof(CodeTable)));
    unsigned long long* hCodes = c->hEntryCodes.as<unsigned long long>();
    CodeTable& hTable = *c->hCodeTable.as<CodeTable>();
    std::memset(&hTable, 0, sizeof(hTable));
    for (int q = 0; q < nSeq; ++q) {
        const SequencePlan& p = s->plans[q];
        if (!p.golomb && p.maxCodeLen > 32) return fail(MNENC_ERR_INTERNAL, "Huffman code longer than 32 bits");
        hTable.seq[q].golombM = p.golomb ? p.golombM : 0;
        for (size_t k = 0; k < s->entryOf[q].size(); ++k)
            hCodes[s->entryOf[q][k]] = p.golomb ? 0ull : (static_cast<unsigned long long>(p.code[k]) | (static_cast<unsigned long long>(p.len[k]) << 32));
    }
    c->ms[4] = static_cast<float>(nowMs() - tHost0);
    CU(cudaMemcpyAsync(c->entryCodes.p, hCodes, static_cast<size_t>(nEntries) * 8, cudaMemcpyHostToDevice, st));
    CU(cudaMemcpyAsync(c->codeTable.p, &hTable, sizeof(CodeTable), cudaMemcpyHostToDevice, st));
    b.entryCodes = c->entryCodes.as<unsigned long long>();
    b.codeTable = c->codeTable.as<CodeTable>();

    // ---- payload bits of this rank's slice of every sequence ----
    CU(s->seqBitsDev.ensure(sizeof(unsigned long long) * kMaxSeq));
    CU(s->hSeqBits.ensure(sizeof(unsigned long long) * kMaxSeq));
    c->launches += launchSeqBits(b, s->M, nSeq, s->seqBitsDev.as<unsigned long long>(), st);
    CU(cudaMemcpyAsync(s->hSeqBits.p, s->seqBitsDev.p, sizeof(unsigned long long) * kMaxSeq, cudaMemcpyDeviceToHost, st));
    CU(waitStream(c));
    for (int q = 0; q < nSeq; ++q) seq_bits[q] = s->hSeqBits.as<unsigned long long>()[q];
    c->ms[7] += static_cast<float>(nowMs() - t0);
    s->phase = 4;
    return MNENC_OK;
} catch (...) {
    return mnenc::abiCaught();
}

int mnenc_stripe_pack(mnenc_t* c, const uint64_t* all_seq_bits, void** out_dev, size_t* out_words, size_t* out_bytes) try {
    if (int rc = needPhase(c, 4)) return rc;
    if (!all_seq_bits || !out_dev || !out_words || !out_bytes) return fail(MNENC_ERR_ARG, "null argument");
    CU(cudaSetDevice(c->device));
    StripeState* s = c->stripe.get();
    const int K = c->K, nSeq = 6 * K + 1, me = s->rank;
    const double t0 = nowMs();
    cudaStream_t st = c->stream;
    BackendBuffers& b = s->b;
    CodeTable& hTable = *c->hCodeTable.as<CodeTable>();

    // ---- bit layout of the WHOLE file (identical on every rank); this rank's slices start after the earlier ranks' ----
    s->overlays.clear();
    uint64_t cursor = 0;
    {  // file header (CompressedImage.cpp:472-486)
        Overlay h;
        h.bitPos = 0;
        h.bits.put(0x4D4E3234u, 32);
        h.bits.put(static_cast<uint32_t>(s->W), 32);
        h.bits.put(static_cast<uint32_t>(s->H), 32);
        h.bits.put(static_cast<uint32_t>(K) & 0xFF, 8);
        h.bits.put(8, 8);
        for (int i = 0; i < 3 * K; ++i) h.bits.put(static_cast<uint16_t>(c->hQuant[(i / K) * kMaxK + (i % K)]), 16);
        cursor = h.bits.nbits;
        s->overlays.push_back(std::move(h));
    }
    uint64_t localPrefix = 0, totalPayload = 0;
    for (int q = 0; q < nSeq; ++q) {
        const SequencePlan& p = s->plans[q];
        Overlay h;
        h.bitPos = cursor;
        if (q > 0) {  // RLE flag (+ length of the WHOLE run-length coded stream) (CompressedImage.cpp:509-515)
            if (s->plan.useRle[q - 1]) {
                h.bits.put(1, 1);
                h.bits.put(static_cast<uint32_t>(s->codedLenGlobal[q]), 32);
            } else {
                h.bits.put(0, 1);
            }
        }
        h.bits.append(p.header);
        cursor += h.bits.nbits;
        s->overlays.push_back(std::move(h));
        uint64_t before = 0, all = 0;
        for (int r = 0; r < s->nranks; ++r) {
            const uint64_t v = all_seq_bits[static_cast<size_t>(r) * nSeq + q];
            if (r < me) before += v;
            all += v;
        }
        if (all != p.payloadBits) return fail(MNENC_ERR_INTERNAL, "payload bits of the ranks do not add up to the planned size");
        hTable.seq[q].bitBase = cursor + before - localPrefix;  // modular: position = bitBase + local scan value
        localPrefix += all_seq_bits[static_cast<size_t>(me) * nSeq + q];
        cursor += p.payloadBits;
        totalPayload += p.payloadBits;
        if (!p.golomb) {
            Overlay e;
            e.bitPos = cursor;
            e.bits.put(p.eofCode, p.eofLen);
            cursor += p.eofLen;
            s->overlays.push_back(std::move(e));
        }
    }
    s->totalBytes = static_cast<size_t>((cursor + 7) / 8);
    s->payloadBits = totalPayload;
    const size_t outAlloc = (s->totalBytes + 16 + 3) / 4 * 4;
    s->outWords = outAlloc / 4;
    CU(c->outWords.ensure(outAlloc));
    CU(cudaMemcpyAsync(c->codeTable.p, &hTable, sizeof(CodeTable), cudaMemcpyHostToDevice, st));
    CU(cudaMemsetAsync(c->outWords.p, 0, outAlloc, st));
    b.outWords = c->outWords.as<uint32_t>();
    CU(cudaEventRecord(c->ev[5], st));
    c->launches += launchPack(b, s->M, nSeq, st);
    c->launches += launchResetHistogram(b, s->nEntries, st);
    c->histDirty = false;
    CU(cudaEventRecord(c->ev[6], st));
    CU(waitStream(c));
    *out_dev = c->outWords.p;
    *out_words = s->outWords;
    *out_bytes = s->totalBytes;
    c->nSymbols = s->M;
    c->nBits = totalPayload;
    c->ms[7] += static_cast<float>(nowMs() - t0);
    s->phase = 5;
    return MNENC_OK;
} catch (...) {
    return mnenc::abiCaught();
}

int mnenc_stripe_finish(mnenc_t* c, uint8_t** out, size_t* out_bytes) try {
    if (int rc = needPhase(c, 5)) return rc;
    if (!out || !out_bytes) return fail(MNENC_ERR_ARG, "null argument");
    CU(cudaSetDevice(c->device));
    StripeState* s = c->stripe.get();
    const double t0 = nowMs();
    *out = nullptr;
    *out_bytes = 0;
    if (s->rank == 0) {
        uint8_t* res = c->stripeResult;  // the in-library exchange has already brought the stream to the host, chunk by chunk
        c->stripeResult = nullptr;
        if (!res) {
            res = resultAcquire(s->totalBytes);
            if (!res) return fail(MNENC_ERR_INTERNAL, "out of host memory");
            cudaError_t e = cudaMemcpyAsync(res, c->outWords.p, s->totalBytes, cudaMemcpyDeviceToHost, c->stream);
            if (e == cudaSuccess) e = waitStream(c);
            if (e != cudaSuccess) {
                resultRelease(res);
                return fail(MNENC_ERR_CUDA, std::string("D2H of the stream: ") + cudaGetErrorString(e));
            }
        }
        mergeOverlays(res, s->overlays);
        *out = res;
        *out_bytes = s->totalBytes;
    }
    float t;
    if (cudaEventElapsedTime(&t, c->ev[3], c->ev[4]) == cudaSuccess) c->ms[3] = t;
    if (cudaEventElapsedTime(&t, c->ev[5], c->ev[6]) == cudaSuccess) c->ms[5] = t;
    c->ms[6] = static_cast<float>(nowMs() - t0);
    c->ms[7] += c->ms[6];
    s->phase = 0;
    return MNENC_OK;
} catch (...) {
    return mnenc::abiCaught();
}

}  // extern "C"

size_t mnenc::stripeTotalBytes(const mnenc_ctx* c) { return c->stripe ? c->stripe->totalBytes : 0; }

extern "C" {

int mnenc_debug_plan_stripes(const uint32_t* records, int nranks, int n_streams, uint32_t* emitted, uint8_t* use_rle, uint64_t* raw_len,
                             uint64_t* rle_len) try {
    if (!records || !emitted || !use_rle || !raw_len || !rle_len || nranks < 1 || n_streams < 0) return fail(MNENC_ERR_ARG, "bad argument");
    StripePlan plan;
    planStripes(reinterpret_cast<const StripeBoundary*>(records), nranks, n_streams, plan);
    std::copy(plan.emitted.begin(), plan.emitted.end(), emitted);
    std::copy(plan.useRle.begin(), plan.useRle.end(), use_rle);
    std::copy(plan.rawLen.begin(), plan.rawLen.end(), raw_len);
    std::copy(plan.rleLen.begin(), plan.rleLen.end(), rle_len);
    return MNENC_OK;
} catch (...) {
    return mnenc::abiCaught();
}

}  // extern "C"
