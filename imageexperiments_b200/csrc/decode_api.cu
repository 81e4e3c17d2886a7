// Decoder entry points of the C ABI (include/mnenc.h): compressed::decodeImageFast
// (CompressionLib/src/CompressedImage.cpp:837-889) = readCompressedFast (:709-782) on the host (decode_host.cpp) + the block
// loop on the GPU (decode_kernels.cu). One call:
//
//   host   header -> `lengths` (pinned) ----H2D----> tile counts + scan (GPU, while the host reads on)
//   host   6K streams -> pinned arena, run lengths expanded, DC chains undone ----H2D---->
//   GPU    gather (streams -> per-block pairs), reconstruct (atoms -> YUV -> RGB8) ----D2H----> pooled pinned result
//
// The quantisation tables come from the stream (the context's own tables are not touched), the dictionary from the context;
// K of the stream may differ from the context's K (the dictionary does not depend on it).
#include "decode_kernels.cuh"
#include "mnenc_ctx.cuh"

using namespace mnenc;

namespace {

constexpr size_t kMaxDecodePairsBytes = size_t(8) << 30;

struct DecodeMisc {  // pinned: uploaded tables, downloaded error flag
    float quant[3 * kMaxK];
    unsigned int error;
};

int decodeImageBody(mnenc_ctx* c, const uint8_t* mn, size_t mnBytes, uint8_t* rgbDev, size_t capacity, size_t stride, uint8_t** hostOut,
                    size_t* widthOut, size_t* heightOut);

// Every failure leaves through here: copies from the context's pinned buffers may still be queued on the stream, and the next
// call on the context rewrites (or reallocates) those buffers - so a failed call drains the stream before it returns.
int decodeImage(mnenc_ctx* c, const uint8_t* mn, size_t mnBytes, uint8_t* rgbDev, size_t capacity, size_t stride, uint8_t** hostOut,
                size_t* widthOut, size_t* heightOut) {
    if (!c || !mn || !widthOut || !heightOut) return fail(MNENC_ERR_ARG, "null argument");
    const int rc = decodeImageBody(c, mn, mnBytes, rgbDev, capacity, stride, hostOut, widthOut, heightOut);
    if (rc != MNENC_OK && c->stream) cudaStreamSynchronize(c->stream);
    return rc;
}

int decodeImageBody(mnenc_ctx* c, const uint8_t* mn, size_t mnBytes, uint8_t* rgbDev, size_t capacity, size_t stride, uint8_t** hostOut,
                    size_t* widthOut, size_t* heightOut) {
    NvtxRange nvtxRange("mnenc:decode");
    static const bool envParallel = [] {  // on unless MNENC_DECODE_THREADS=1 (measured on the B200 host: 1240 -> 2108 MP/s per image)
        const char* e = std::getenv("MNENC_DECODE_THREADS");
        return !e || std::atoi(e) != 1;
    }();
    const bool parallelReader = envParallel && !c->inBatch;
    if (hostOut) *hostOut = nullptr;
    *widthOut = *heightOut = 0;
    const double t0 = nowMs();
    CU(cudaSetDevice(c->device));
    MnParser P(mn, mnBytes, parallelReader ? c->pool.get() : nullptr);
    std::string err;
    if (P.readHeader(err)) return fail(MNENC_ERR_STREAM, err);
    const size_t w = P.width, h = P.height, nblocks = P.nblocks, nbc = 3 * nblocks;
    const int K = P.K;
    if (nblocks > 0x7FFFFFFFu / 3) return fail(MNENC_ERR_ARG, "image too large");
    if (rgbDev) {
        if (stride < 3 * w) return fail(MNENC_ERR_ARG, "row stride smaller than 3 * width");
        if (capacity < stride * (h - 1) + 3 * w) return fail(MNENC_ERR_ARG, "device buffer too small for the image in the stream");
    } else {
        stride = 3 * w;
    }
    for (auto& m : c->ms) m = 0;
    c->launches = 0;
    c->haveSymbols = false;  // counts / pairs now hold the DECODED image's symbols

    // `lengths` -> GPU, tile offsets of every (layer, depth) stream; then the streams, while the GPU works on the counts.
    // Single-image calls: the context's host threads stand by as helpers of the parallel reader (decode_host.cpp);
    // MNENC_DECODE_THREADS=1 keeps the reader on the calling thread. Batch lanes always read serially (one stream per lane).
    double t1 = t0, t2 = t0;
    size_t N = 0;
    BackendBuffers b{};
    const int rcParse = P.withTeam([&]() -> int {
        CU(c->hDecCounts.ensure(nbc * 2 + 16));  // + 16: the `lengths` reader stores 8 bytes at a time
        CU(c->hDecMisc.ensure(sizeof(DecodeMisc)));
        if (P.readLengths(c->hDecCounts.as<uint16_t>(), err)) return fail(MNENC_ERR_STREAM, err);
        t1 = nowMs();
        b.K = K, b.nblocks = static_cast<uint32_t>(nblocks), b.ntiles = static_cast<uint32_t>((nblocks + kTileBlocks - 1) / kTileBlocks);
        CU(c->counts.ensure(nbc * 2));
        // [block-channel][K] symbol slots: up to 256 bytes of device memory per block-channel for a stream that may spend one bit
        // on it; the cap keeps a crafted header from demanding more than any real image needs (16384^2 at K = 64 is 3.2 GB)
        if (nbc * static_cast<size_t>(K) * 4 > kMaxDecodePairsBytes) return fail(MNENC_ERR_ARG, "decoded image too large for this build (symbol slots > 8 GiB)");
        CU(c->pairs.ensure(nbc * static_cast<size_t>(K) * 4));
        CU(c->tileCounts.ensure(static_cast<size_t>(b.ntiles) * 3 * K * 4));
        CU(c->totals.ensure(3 * kMaxK * 4));
        CU(c->decQuant.ensure(sizeof(DecodeMisc)));
        b.counts = c->counts.as<uint16_t>(), b.tileCounts = c->tileCounts.as<uint32_t>(), b.totals = c->totals.as<uint32_t>();
        CU(cudaMemcpyAsync(c->counts.p, c->hDecCounts.p, nbc * 2, cudaMemcpyHostToDevice, c->stream));
        c->launches += launchTileCounts(b, c->stream);
        N = P.streamOffset[6 * K];
        CU(c->hDecArena.ensure(N * 2 + 16));
        if (P.readStreams(c->hDecArena.as<uint16_t>(), err)) {
            cudaStreamSynchronize(c->stream);
            return fail(MNENC_ERR_STREAM, err);
        }
        t2 = nowMs();
        return MNENC_OK;
    });
    if (rcParse) return rcParse;
    CU(c->arena.ensure(N * 2 + 16));
    b.arena = c->arena.as<uint16_t>();
    if (N) CU(cudaMemcpyAsync(c->arena.p, c->hDecArena.p, N * 2, cudaMemcpyHostToDevice, c->stream));
    DecodeMisc* misc = c->hDecMisc.as<DecodeMisc>();
    std::memcpy(misc->quant, P.quant, sizeof(misc->quant));
    misc->error = 0;
    CU(cudaMemcpyAsync(c->decQuant.p, misc, sizeof(DecodeMisc), cudaMemcpyHostToDevice, c->stream));
    StreamOffsets so{};
    for (int s = 0; s <= 6 * K; ++s) so.off[s] = static_cast<uint32_t>(P.streamOffset[s]);
    CU(cudaEventRecord(c->ev[0], c->stream));
    c->launches += launchGather(b, so, c->pairs.as<uint32_t>(), c->stream);

    uint8_t* dst = rgbDev;
    if (!dst) {
        CU(c->rgb.ensure(stride * h));
        dst = c->rgb.as<uint8_t>();
    }
    DecodeParams p{};
    p.counts = c->counts.as<uint16_t>(), p.pairs = c->pairs.as<uint32_t>();
    p.quant = c->decQuant.as<float>();
    p.rows32 = c->rows32.as<float>(), p.detRows = c->detRows.as<uint8_t>();
    p.K = K, p.nBase = c->nBase;
    p.nblocks = static_cast<uint32_t>(nblocks), p.nby = static_cast<uint32_t>((h + 7) / 8);
    p.width = static_cast<uint32_t>(w), p.height = static_cast<uint32_t>(h);
    p.rgb = dst, p.stride = stride;
    p.error = reinterpret_cast<unsigned int*>(c->decQuant.as<uint8_t>() + offsetof(DecodeMisc, error));
    c->launches += launchReconstruct(p, c->numSMs, c->stream);
    CU(cudaGetLastError());
    CU(cudaEventRecord(c->ev[1], c->stream));

    uint8_t* result = nullptr;
    if (hostOut) {
        result = resultAcquire(3 * w * h);
        if (!result) return fail(MNENC_ERR_CUDA, "out of pinned host memory for the decoded image");
        CU(cudaMemcpyAsync(result, dst, 3 * w * h, cudaMemcpyDeviceToHost, c->stream));
    }
    cudaError_t e = cudaMemcpyAsync(&misc->error, p.error, sizeof(unsigned int), cudaMemcpyDeviceToHost, c->stream);
    if (e == cudaSuccess) e = waitStream(c);
    if (e != cudaSuccess) {
        resultRelease(result);
        return fail(MNENC_ERR_CUDA, std::string("decode: ") + cudaGetErrorString(e));
    }
    if (misc->error) {
        resultRelease(result);
        return fail(MNENC_ERR_STREAM, "atom index outside its block's dictionary");
    }
    if (hostOut) *hostOut = result;
    *widthOut = w, *heightOut = h;
    c->nSymbols = nbc + N, c->nBits = P.bitsRead;
    c->ms[0] = static_cast<float>(t1 - t0);         // header + `lengths`
    c->ms[1] = static_cast<float>(t2 - t1);         // symbol streams (host), tile counts (GPU) underneath
    c->ms[2] = static_cast<float>(nowMs() - t2);    // uploads, gather, reconstruct, download
    cudaEventElapsedTime(&c->ms[3], c->ev[0], c->ev[1]);  // gather + reconstruct kernels (CUDA events)
    c->ms[7] = static_cast<float>(nowMs() - t0);
    return MNENC_OK;
}

}  // namespace

extern "C" int mnenc_decode_rgb8(mnenc_t* c, const uint8_t* mn, size_t mn_bytes, uint8_t** rgb, size_t* width, size_t* height) try {
    if (!rgb) return fail(MNENC_ERR_ARG, "null argument");
    return decodeImage(c, mn, mn_bytes, nullptr, 0, 0, rgb, width, height);
} catch (...) {
    return mnenc::abiCaught();
}

extern "C" int mnenc_decode_rgb8_device(mnenc_t* c, const uint8_t* mn, size_t mn_bytes, uint8_t* rgb_dev, size_t capacity_bytes,
                                        size_t row_stride_bytes, size_t* width, size_t* height) try {
    if (!rgb_dev) return fail(MNENC_ERR_ARG, "null argument");
    return decodeImage(c, mn, mn_bytes, rgb_dev, capacity_bytes, row_stride_bytes, nullptr, width, height);
} catch (...) {
    return mnenc::abiCaught();
}

extern "C" int mnenc_stream_info(const uint8_t* mn, size_t mn_bytes, size_t* width, size_t* height, int* K, int* block_size, float* quant) try {
    if (!mn) return fail(MNENC_ERR_ARG, "null argument");
    MnParser P(mn, mn_bytes);
    std::string err;
    if (P.readHeader(err, true)) return fail(MNENC_ERR_STREAM, err);
    if (width) *width = P.width;
    if (height) *height = P.height;
    if (K) *K = P.K;
    if (block_size) *block_size = P.blockSize;
    if (quant) std::memcpy(quant, P.quant, sizeof(float) * 3 * P.K);
    return MNENC_OK;
} catch (...) {
    return mnenc::abiCaught();
}

extern "C" int mnenc_debug_parse(const uint8_t* mn, size_t mn_bytes, uint16_t* lengths, size_t lengths_capacity, uint64_t* stream_offsets,
                                 uint16_t* symbols, size_t symbols_capacity, int threads) try {
    if (!mn || !lengths || !stream_offsets) return fail(MNENC_ERR_ARG, "null argument");
    std::unique_ptr<HostPool> pool;
    if (threads > 1) pool = std::make_unique<HostPool>(std::min(threads, 64));
    MnParser P(mn, mn_bytes, pool.get());
    std::string err;
    if (P.readHeader(err)) return fail(MNENC_ERR_STREAM, err);
    if (lengths_capacity < 3 * P.nblocks) return fail(MNENC_ERR_ARG, "lengths buffer too small");
    int code = MNENC_OK;
    P.withTeam([&] {
        if (P.readLengths(lengths, err)) return code = MNENC_ERR_STREAM;
        for (int s = 0; s <= 6 * P.K; ++s) stream_offsets[s] = P.streamOffset[s];
        if (!symbols) return 0;  // sizes only
        if (symbols_capacity < P.streamOffset[6 * P.K]) return err = "symbols buffer too small", code = MNENC_ERR_ARG;
        if (P.readStreams(symbols, err)) return code = MNENC_ERR_STREAM;
        return 0;
    });
    return code == MNENC_OK ? MNENC_OK : fail(code, err);
} catch (...) {
    return mnenc::abiCaught();
}
