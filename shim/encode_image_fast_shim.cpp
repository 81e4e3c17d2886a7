// Drop-in host shim: the reference's own C++ entry point, same signature, backed by the GPU library through the
// C ABI of include/mnenc.h. This is the file a maintainer of mnesbit/ImageExperiments adds to CompressionLib
// (INTEGRATION.md); it needs the reference's headers (Eigen, image.h, MatchingPursuit.h), so it is compiled only
// where the reference tree is present (oracle/Makefile target `shim`, output in oracle/_ref/).
//
//   compressed::encodeImageFast   CompressionLib/inc/CompressedImage.h:67-73, src/CompressedImage.cpp:578-633
//   compressed::decodeImageFast   CompressionLib/inc/CompressedImage.h:75,    src/CompressedImage.cpp:837-889
//
// The dictionary never crosses the reference API as data: encodeImageFast receives three opaque
// DynamicDictionaryFunctionFast callbacks (MatchingPursuit.h:20). They are pure functions of their arguments
// (CompressedImage.cpp:252-289,331-339): dynamic(0,{}) returns BaseDict, dynamic(1,{BasisChoice{c,0}}) returns
// BaseDict followed by DetailBasis[c]. 1 + n_base calls per channel therefore recover exactly the floats the
// linked decoder will use, and the result is cached per process keyed by a cheap probe.
#include <cstring>
#include <memory>
#include <stdexcept>
#include <vector>

#include "CompressionLib/inc/CompressedImage.h"
#include "mnenc.h"

namespace mnenc_shim {

namespace {

struct Cached {
    mnenc_t* ctx = nullptr;
    int K = 0;
    std::vector<float> probe;  // dynamicY(1,{0}) — identifies the dictionary build
    ~Cached() {
        if (ctx) mnenc_destroy(ctx);
    }
};
Cached g_cache;

Eigen::MatrixXf callDynamic(const matching::DynamicDictionaryFunctionFast& f, int K, int prev) {
    std::vector<matching::BasisChoice> choices(K);
    if (prev < 0) return f(0, choices);
    choices[0].deltaId = static_cast<unsigned short>(prev);
    choices[0].intCoeff = 0;
    return f(1, choices);
}

void appendRows(std::vector<float>& out, const Eigen::MatrixXf& m, Eigen::Index firstRow) {
    for (Eigen::Index r = firstRow; r < m.rows(); ++r)
        for (Eigen::Index j = 0; j < m.cols(); ++j) out.push_back(m(r, j));
}

void failWith(const char* what) { throw new std::range_error(std::string(what) + ": " + mnenc_last_error()); }  // reference convention

}  // namespace

namespace {

// GPU context for the dictionary behind the three callbacks (created once per dictionary build and K, then cached).
mnenc_t* contextFor(const size_t K, const size_t blockSize, const matching::DynamicDictionaryFunctionFast& dynamicY,
                    const matching::DynamicDictionaryFunctionFast& dynamicU, const matching::DynamicDictionaryFunctionFast& dynamicV) {
    const int k = static_cast<int>(K);
    // --- dictionary identity probe (one callback call) ---
    const Eigen::MatrixXf probeM = callDynamic(dynamicY, k, 0);
    std::vector<float> probe;
    appendRows(probe, probeM, 0);
    if (!g_cache.ctx || g_cache.K != k || g_cache.probe != probe) {
        if (g_cache.ctx) mnenc_destroy(g_cache.ctx);
        g_cache.ctx = nullptr;
        const Eigen::MatrixXf base = callDynamic(dynamicY, k, -1);
        const int nBase = static_cast<int>(base.rows());
        std::vector<float> baseRows;
        appendRows(baseRows, base, 0);
        std::vector<int32_t> detailRows(nBase);
        std::vector<float> detail[3];
        const matching::DynamicDictionaryFunctionFast* dyn[3] = {&dynamicY, &dynamicU, &dynamicV};
        for (int ch = 0; ch < 3; ++ch)
            for (int a = 0; a < nBase; ++a) {
                const Eigen::MatrixXf full = callDynamic(*dyn[ch], k, a);
                detailRows[a] = static_cast<int32_t>(full.rows() - nBase);
                appendRows(detail[ch], full, nBase);
            }
        // MNENC_DEVICES="0,1,..." / "all": one image is sharded by block columns over those GPUs; otherwise MNENC_DEVICE (default 0)
        if (mnenc_create_auto(&g_cache.ctx, k, static_cast<int>(blockSize), baseRows.data(), nBase, detailRows.data(), detail[0].data(),
                              detail[1].data(), detail[2].data()) != MNENC_OK)
            failWith("mnenc_create_auto");
        g_cache.K = k;
        g_cache.probe = probe;
    }
    return g_cache.ctx;
}

}  // namespace

std::unique_ptr<uint8_t[]> encodeImageFast(const img::image<img::rgb>* imgIn, const size_t K, const size_t blockSize,
                                           const Eigen::VectorXf& quantY, const Eigen::VectorXf& quantU, const Eigen::VectorXf& quantV,
                                           const matching::DynamicDictionaryFunctionFast& dynamicY,
                                           const matching::DynamicDictionaryFunctionFast& dynamicU,
                                           const matching::DynamicDictionaryFunctionFast& dynamicV, size_t& outputByteSize) {
    mnenc_t* ctx = contextFor(K, blockSize, dynamicY, dynamicU, dynamicV);
    std::vector<float> quant(3 * K);
    for (size_t i = 0; i < K; ++i) {
        quant[i] = quantY[i];
        quant[K + i] = quantU[i];
        quant[2 * K + i] = quantV[i];
    }
    if (mnenc_set_quant(ctx, quant.data()) != MNENC_OK) failWith("mnenc_set_quant");
    // img::image<rgb> stores rows contiguously (image.h:67-75): data + y*w
    uint8_t* bytes = nullptr;
    size_t n = 0;
    if (mnenc_encode_rgb8(ctx, reinterpret_cast<const uint8_t*>(imgIn->data), imgIn->width(), imgIn->height(),
                          imgIn->width() * sizeof(img::rgb), &bytes, &n) != MNENC_OK)
        failWith("mnenc_encode_rgb8");
    std::unique_ptr<uint8_t[]> out = std::make_unique<uint8_t[]>(n);
    std::memcpy(out.get(), bytes, n);
    mnenc_free(bytes);
    outputByteSize = n;
    return out;
}

// decodeImageFast (CompressedImage.cpp:837-889). The reference builds its context inside readCompressedFast (:726,
// createCompressionContextFast(K, blockSize, 0.0)); the shim does the same once per (K, blockSize) — the dictionary does not
// depend on the quality argument — and hands its callbacks to contextFor(), so the GPU decodes with the floats of THIS build.
std::unique_ptr<img::image<img::rgb>> decodeImageFast(const uint8_t bytes[], size_t byteSize) {
    size_t w = 0, h = 0;
    int K = 0, blockSize = 0;
    if (mnenc_stream_info(bytes, byteSize, &w, &h, &K, &blockSize, nullptr) != MNENC_OK) failWith("Invalid input data");
    static std::unique_ptr<compressed::CompressionContextFast> refCtx;
    if (!refCtx || refCtx->K != static_cast<size_t>(K) || refCtx->BlockSize != static_cast<size_t>(blockSize))
        refCtx = compressed::createCompressionContextFast(K, blockSize, 0.0);
    mnenc_t* ctx = contextFor(refCtx->K, refCtx->BlockSize, refCtx->Y.Dynamic, refCtx->U.Dynamic, refCtx->V.Dynamic);
    uint8_t* rgb = nullptr;
    if (mnenc_decode_rgb8(ctx, bytes, byteSize, &rgb, &w, &h) != MNENC_OK) failWith("Invalid bitstream");
    auto out = std::make_unique<img::image<img::rgb>>(static_cast<int>(w), static_cast<int>(h), false);
    std::memcpy(out->data, rgb, w * h * sizeof(img::rgb));
    mnenc_free(rgb);
    return out;
}

}  // namespace mnenc_shim

// C entry used by tests/test_shim.py: builds a reference context, encodes through the shim above.
extern "C" int mnenc_shim_selftest(const uint8_t* rgb, size_t w, size_t h, double bpp, uint8_t** out, size_t* outBytes) {
    try {
        auto ctx = compressed::createCompressionContextFast(32, 8, bpp);
        img::image<img::rgb> im(w, h, false);
        std::memcpy(im.data, rgb, w * h * 3);
        size_t n = 0;
        auto bytes = mnenc_shim::encodeImageFast(&im, ctx->K, ctx->BlockSize, ctx->Y.Quant, ctx->U.Quant, ctx->V.Quant, ctx->Y.Dynamic,
                                                 ctx->U.Dynamic, ctx->V.Dynamic, n);
        *out = static_cast<uint8_t*>(std::malloc(n ? n : 1));
        std::memcpy(*out, bytes.get(), n);
        *outBytes = n;
        return 0;
    } catch (std::exception* e) {
        return -1;
    } catch (...) {
        return -2;
    }
}

// C entry used by tests/test_shim.py: decodes through the shim's decodeImageFast.
extern "C" int mnenc_shim_selftest_decode(const uint8_t* mn, size_t n, uint8_t** out, size_t* w, size_t* h) {
    try {
        auto im = mnenc_shim::decodeImageFast(mn, n);
        *w = im->width(), *h = im->height();
        *out = static_cast<uint8_t*>(std::malloc(*w * *h * 3 + 1));
        std::memcpy(*out, im->data, *w * *h * 3);
        return 0;
    } catch (std::exception* e) {
        return -1;
    } catch (...) {
        return -2;
    }
}
