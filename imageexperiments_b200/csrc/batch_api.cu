// Batch mode (BASELINE.json configs[4]: many independent frames per GPU, throughput mode; SURVEY.md §8e "C5"): the
// reference encodes images one after the other (Compression.cpp:117-123 in a loop); every image is an independent `.mn`
// file, so a batch is pipelined over LANES — contexts with their own stream, device buffers, pinned buffers and host
// threads that share one dictionary in HBM. While one lane waits for its host stage (stream sizes, the 193 code tables,
// D2H + header merge) or for its pixels to arrive, another lane's matching-pursuit kernel owns the SMs, and the copy
// engines move the next image's pixels and the previous image's stream. Each image goes through exactly the
// single-image path (mnenc_encode_rgb8*), so every stream is byte-identical to the one-at-a-time result.
#include <atomic>
#include <thread>

#include "mnenc_ctx.cuh"

using namespace mnenc;

namespace {

int makeLane(mnenc_ctx* parent, int threads, mnenc_ctx** out) {
    mnenc_ctx* l = new mnenc_ctx;
    l->device = parent->device, l->K = parent->K, l->nBase = parent->nBase, l->numSMs = parent->numSMs;
    l->epsAbs16 = parent->epsAbs16, l->epsRel16Hi = parent->epsRel16Hi;
    l->detRows.alias(parent->detRows);
    l->baseUm16.alias(parent->baseUm16), l->detUm16.alias(parent->detUm16), l->rows32.alias(parent->rows32);
    l->quant.alias(parent->quant);
    l->hostThreads = threads;
    l->pool = std::make_unique<HostPool>(threads);
    *out = l;
    CU(cudaStreamCreateWithFlags(&l->stream, cudaStreamNonBlocking));
    for (auto& e : l->ev) CU(cudaEventCreate(&e));
    CU(cudaEventCreateWithFlags(&l->evSync, cudaEventDisableTiming));
    for (auto& e : l->evWave) CU(cudaEventCreateWithFlags(&e, cudaEventDisableTiming | cudaEventBlockingSync));
    return MNENC_OK;
}

constexpr int kMaxLanes = 16;

// The context itself + up to `want` - 1 extra lanes (MNENC_LANES / MNENC_DECODE_LANES override, at most kMaxLanes in all), created on first use.
int acquireLanes(mnenc_ctx* c, size_t nImages, std::vector<mnenc_ctx*>& lanes, int want = 6, const char* envName = "MNENC_LANES") {
    if (const char* e = std::getenv(envName)) want = std::atoi(e);
    want = static_cast<int>(std::min<size_t>(static_cast<size_t>(std::clamp(want, 1, kMaxLanes)), nImages));
    while (static_cast<int>(c->lanes.size()) + 1 < want) {
        mnenc_ctx* lane = nullptr;
        const int rc = makeLane(c, std::max(1, c->hostThreads / 2), &lane);
        if (lane) c->lanes.push_back(lane);
        if (rc) return rc;
    }
    lanes.assign(1, c);
    for (int i = 0; i + 1 < want; ++i) lanes.push_back(c->lanes[i]);
    return MNENC_OK;
}

}  // namespace

extern "C" int mnenc_encode_batch_rgb8(mnenc_t* c, size_t n_images, const uint8_t* const* rgb, int on_device, const size_t* width,
                                       const size_t* height, const size_t* row_stride_bytes, uint8_t** out, size_t* out_bytes) try {
    if (!c || (n_images && (!rgb || !width || !height || !row_stride_bytes || !out || !out_bytes))) return fail(MNENC_ERR_ARG, "null argument");
    if (!c->quantSet) return fail(MNENC_ERR_STATE, "mnenc_set_quant has not been called");
    for (size_t i = 0; i < n_images; ++i) out[i] = nullptr, out_bytes[i] = 0;
    if (n_images == 0) return MNENC_OK;
    CU(cudaSetDevice(c->device));
    std::vector<mnenc_ctx*> lanes;
    // Measured on one B200 (16 host cores): 16 MP images 5441 / 5527 / 5580 MP/s with 6 / 8 / 12 lanes, 192 frames of 1080p
    // 3984 / 4210 / 4244 / 4214 MP/s with 6 / 9 / 12 / 16 - small kernels of more images fill the gaps of the big ones. With few
    // host threads per process (one process per GPU on a shared host) the lanes' threads would only crowd each other: 6.
    if (const int rc = acquireLanes(c, n_images, lanes, c->hostThreads >= 12 ? 9 : 6)) return rc;
    const int want = static_cast<int>(lanes.size());
    for (int i = 1; i < want; ++i) {
        std::memcpy(lanes[i]->hQuant, c->hQuant, sizeof(c->hQuant));  // the device table is shared; the header needs the host copy
        lanes[i]->quantSet = true;
    }
    // Frames of one geometry that are small enough to leave the per-iteration kernels of the matching pursuit launch-bound share
    // one pursuit, `group` frames at a time (about the block-channel count of a 16 MP image).
    size_t group = 1;
    {
        bool uniform = n_images > 1;
        for (size_t i = 1; i < n_images && uniform; ++i)
            uniform = width[i] == width[0] && height[i] == height[0] && (!on_device || row_stride_bytes[i] == row_stride_bytes[0]);
        if (uniform && width[0] && height[0] && row_stride_bytes[0] >= 3 * width[0]) {
            for (size_t i = 0; i < n_images && uniform; ++i) uniform = rgb[i] != nullptr && row_stride_bytes[i] >= 3 * width[0];
            const size_t nbcFrame = 3 * static_cast<size_t>(frameBlockSlots(width[0], height[0]));
            if (uniform && nbcFrame < 400000) {
                // up to twice the block-channels of a 16 MP image per group, and a number of groups that fills every lane in every
                // round (64 frames over 6 lanes: 6 groups of 11, not 8 of 8 with two lanes busy in the second round)
                const size_t most = std::clamp<size_t>(1572864 / nbcFrame, 1, 16);
                const size_t rounds = (n_images + lanes.size() * most - 1) / (lanes.size() * most);
                group = std::max<size_t>(1, (n_images + lanes.size() * rounds - 1) / (lanes.size() * rounds));
            }
            if (uniform)
                if (const char* e = std::getenv("MNENC_FRAME_GROUP")) group = std::clamp<size_t>(static_cast<size_t>(std::atoi(e)), 1, 64);
        }
    }
    std::atomic<size_t> next{0};
    std::atomic<int> firstError{MNENC_OK};
    int launches[kMaxLanes] = {};
    double flops[kMaxLanes] = {};
    float mpMs[kMaxLanes] = {};
    auto work = [&](int li) {
        mnenc_ctx* l = lanes[li];
        cudaSetDevice(l->device);
        while (firstError.load(std::memory_order_relaxed) == MNENC_OK) {
            const size_t i = next.fetch_add(group);
            if (i >= n_images) break;
            const size_t g = std::min(group, n_images - i);
            int rc;
            if (g > 1) {
                rc = encodeFrames(l, rgb + i, on_device, g, width[i], height[i], row_stride_bytes + i, out + i, out_bytes + i);
            } else {
                rc = on_device ? mnenc_encode_rgb8_device(l, rgb[i], width[i], height[i], row_stride_bytes[i], &out[i], &out_bytes[i])
                               : mnenc_encode_rgb8(l, rgb[i], width[i], height[i], row_stride_bytes[i], &out[i], &out_bytes[i]);
            }
            if (rc != MNENC_OK) {
                int expected = MNENC_OK;
                if (firstError.compare_exchange_strong(expected, rc)) c->laneError = "image " + std::to_string(i) + ": " + mnenc_last_error();
                break;
            }
            launches[li] += l->launches;
            flops[li] += l->mpFlops;
            mpMs[li] += l->ms[1];
        }
    };
    const double t0 = nowMs();
    std::vector<std::thread> threads;
    for (int li = 1; li < want; ++li) threads.emplace_back(work, li);
    work(0);
    for (auto& t : threads) t.join();
    if (const int rc = firstError.load()) {
        for (size_t i = 0; i < n_images; ++i) {
            resultRelease(out[i]);
            out[i] = nullptr, out_bytes[i] = 0;
        }
        return fail(rc, c->laneError);
    }
    // accounting of the whole batch on the parent context (mnenc_last_timing / mnenc_last_work)
    c->launches = 0, c->mpFlops = 0, c->ms[1] = 0;
    for (int li = 0; li < want; ++li) c->launches += launches[li], c->mpFlops += flops[li], c->ms[1] += mpMs[li];
    c->ms[7] = static_cast<float>(nowMs() - t0);
    return MNENC_OK;
} catch (...) {
    return mnenc::abiCaught();
}

// Decoder twin: n independent `.mn` streams -> n images (decodeImageFast in a loop). The entropy decode of one stream is one
// serial chain on one host thread (decode_host.cpp), so the lanes here buy host parallelism first, copy/kernel overlap second.
extern "C" int mnenc_decode_batch_rgb8(mnenc_t* c, size_t n_images, const uint8_t* const* mn, const size_t* mn_bytes, uint8_t** rgb,
                                       size_t* width, size_t* height) try {
    if (!c || (n_images && (!mn || !mn_bytes || !rgb || !width || !height))) return fail(MNENC_ERR_ARG, "null argument");
    for (size_t i = 0; i < n_images; ++i) rgb[i] = nullptr, width[i] = height[i] = 0;
    if (n_images == 0) return MNENC_OK;
    CU(cudaSetDevice(c->device));
    std::vector<mnenc_ctx*> lanes;
    // the entropy decode of a stream is one serial chain on one host thread: more lanes than the encoder uses (measured on the
    // B200 host: 6 lanes 5.7, 12 lanes ~10 GP/s)
    if (const int rc = acquireLanes(c, n_images, lanes, 12, "MNENC_DECODE_LANES")) return rc;
    const int want = static_cast<int>(lanes.size());
    std::atomic<size_t> next{0};
    std::atomic<int> firstError{MNENC_OK};
    int launches[kMaxLanes] = {};
    auto work = [&](int li) {
        mnenc_ctx* l = lanes[li];
        cudaSetDevice(l->device);
        l->inBatch = true;
        struct Reset {
            mnenc_ctx* l;
            ~Reset() { l->inBatch = false; }
        } reset{l};
        while (firstError.load(std::memory_order_relaxed) == MNENC_OK) {
            const size_t i = next.fetch_add(1);
            if (i >= n_images) break;
            const int rc = mnenc_decode_rgb8(l, mn[i], mn_bytes[i], &rgb[i], &width[i], &height[i]);
            if (rc != MNENC_OK) {
                int expected = MNENC_OK;
                if (firstError.compare_exchange_strong(expected, rc)) c->laneError = "stream " + std::to_string(i) + ": " + mnenc_last_error();
                break;
            }
            launches[li] += l->launches;
        }
    };
    const double t0 = nowMs();
    std::vector<std::thread> threads;
    for (int li = 1; li < want; ++li) threads.emplace_back(work, li);
    work(0);
    for (auto& t : threads) t.join();
    if (const int rc = firstError.load()) {
        for (size_t i = 0; i < n_images; ++i) {
            resultRelease(rgb[i]);
            rgb[i] = nullptr, width[i] = height[i] = 0;
        }
        return fail(rc, c->laneError);
    }
    c->launches = 0;
    for (int li = 0; li < want; ++li) c->launches += launches[li];
    for (auto& m : c->ms) m = 0;
    c->ms[7] = static_cast<float>(nowMs() - t0);
    return MNENC_OK;
} catch (...) {
    return mnenc::abiCaught();
}
