// mnenc_scatter_streams_static: the symbol-stream scatter for callers that must not allocate GPU memory.
//
// Who this is for. Some hosts of this library run in places where cudaMalloc / cudaFree are off limits: a plugin that
// lives inside another program's memory budget (a media server or an inference runtime that hands out every byte of HBM
// itself and treats a foreign allocation as a leak), code that runs between two kernels of a latency-critical loop, where
// an allocator call means an implicit device synchronisation, and processes whose allocator is already exhausted on purpose
// (all of HBM given to one pool at start-up). Every other entry point of include/mnenc.h grows per-context device buffers
// on demand, so none of them suits such a caller. This one does its work with the caller's input and output buffers plus a
// fixed scratch array that is part of the library image itself.
//
// What it computes. The scatter step of compressed::encodeImageFast (CompressionLib/src/CompressedImage.cpp:612-629): the
// per-block matching-pursuit result (counts + BasisChoice pairs, the layout of mnenc_get_symbols) is rearranged into the 6K
// symbol streams that writeCompressed goes on to code. The kernels are the ones the normal encoder runs (tileCountsKernel,
// tileScanKernel, scatterKernel in backend_kernels.cu); what differs is only where their scratch lives. In the normal path
// the per-tile stream offsets sit in a cudaMalloc'ed context buffer (BackendBuffers::tileCounts / totals); here they sit in
// the file-scope __device__ array below.
//
// The scratch array: g_scatterScratch, 1 573 248 bytes (4096 tiles x 96 columns x 4 B, plus 96 column totals). It is a
// static part of the module, so every GPU on which the library's kernels get loaded carries its own copy from module load
// until the process exits, whether or not this entry point is ever called. That is the price of allocating nothing, and the
// reason the limit is modest: at most 2^20 blocks (e.g. 8192 x 8192 pixels); larger inputs are refused, not truncated.
//
// Sharing. There is one array per GPU and every call on that GPU uses it. It still holds the previous call's offsets when
// the next call starts. Nothing in it has to start at zero: tileCountsKernel overwrites all ntiles x 3K entries the later
// kernels read and tileScanKernel overwrites the 3K totals, before anything reads them, so no clearing pass is needed (the
// test runs two different geometries back to back to show that leftovers do not leak through). Two calls that are in
// flight on the same GPU at the same time — two host threads, or two streams — write the same array and corrupt each
// other; callers must serialise them per GPU.
//
// Set-up. In host code the name g_scatterScratch is only a handle for the symbol, not a device pointer; passing it to a
// kernel would make the kernel fault (or, on systems where the GPU can reach host memory, scribble over host memory without
// any error). The first call on each device therefore asks the runtime for the array's device address
// (cudaGetSymbolAddress, with that device current) and caches it per device; that same first call is also when the CUDA
// runtime loads the module onto the device. Make the first call on a device at start-up, outside any stream capture.
#include <mutex>

#include "mnenc_ctx.cuh"

namespace {

using namespace mnenc;

constexpr size_t kStaticMaxBlocks = size_t(1) << 20;  // stated limit of mnenc_scatter_streams_static
constexpr size_t kStaticMaxTiles = kStaticMaxBlocks / kTileBlocks;
constexpr size_t kStaticTotalsAt = kStaticMaxTiles * 3 * kMaxK;  // u32 index of the 3K column totals
constexpr size_t kStaticWords = kStaticTotalsAt + 3 * kMaxK;

static __device__ __align__(256) uint32_t g_scatterScratch[kStaticWords];

constexpr int kMaxDevices = 64;
std::mutex g_setupLock;
uint32_t* g_scratchOnDevice[kMaxDevices] = {};  // device address of g_scatterScratch, per CUDA ordinal (0 = not looked up yet)

// Device address of the scratch array on the CURRENT device; looked up once per device.
int scratchOfCurrentDevice(int dev, uint32_t** out) {
    std::lock_guard<std::mutex> hold(g_setupLock);
    if (!g_scratchOnDevice[dev]) {
        void* p = nullptr;
        CU(cudaGetSymbolAddress(&p, g_scatterScratch));
        if (!p) return fail(MNENC_ERR_CUDA, "cudaGetSymbolAddress returned a null address for the static scratch array");
        g_scratchOnDevice[dev] = static_cast<uint32_t*>(p);
    }
    *out = g_scratchOnDevice[dev];
    return MNENC_OK;
}

// The pointer must be memory the current GPU owns (or managed memory): a host pointer here would otherwise surface as a
// sticky "illegal address" fault that kills the caller's CUDA context.
int requireDevicePointer(const void* p, int dev, const char* what) {
    cudaPointerAttributes a{};
    cudaError_t e = cudaPointerGetAttributes(&a, p);
    if (e != cudaSuccess) {
        cudaGetLastError();
        return fail(MNENC_ERR_ARG, std::string(what) + ": not a pointer CUDA knows (" + cudaGetErrorString(e) + ")");
    }
    if (a.type == cudaMemoryTypeManaged) return MNENC_OK;
    if (a.type != cudaMemoryTypeDevice) return fail(MNENC_ERR_ARG, std::string(what) + " must point to device memory");
    if (a.device != dev) return fail(MNENC_ERR_ARG, std::string(what) + " belongs to another GPU than the current one");
    return MNENC_OK;
}

}  // namespace

extern "C" int mnenc_scatter_streams_static(const uint16_t* counts_dev, const uint16_t* pairs_dev, int K, size_t n_blocks,
                                            uint16_t* streams_dev, size_t streams_capacity, uint64_t* stream_offsets,
                                            void* cuda_stream) try {
    using namespace mnenc;
    if (!stream_offsets) return fail(MNENC_ERR_ARG, "null argument");
    if (K < 1 || K > kMaxK) return fail(MNENC_ERR_ARG, "K must be in [1,32]");
    if (n_blocks > kStaticMaxBlocks)
        return fail(MNENC_ERR_ARG, "mnenc_scatter_streams_static handles at most 1048576 blocks (its scratch is a fixed 1.5 MB array); got " +
                                       std::to_string(n_blocks) + " - use a context-based entry point for larger images");
    const int nStreams = 6 * K;
    if (n_blocks == 0) {
        for (int s = 0; s <= nStreams; ++s) stream_offsets[s] = 0;
        return MNENC_OK;
    }
    if (!counts_dev || !pairs_dev) return fail(MNENC_ERR_ARG, "null argument");
    int dev = -1;
    CU(cudaGetDevice(&dev));
    if (dev < 0 || dev >= kMaxDevices) return fail(MNENC_ERR_ARG, "device ordinal out of range");
    if (int rc = requireDevicePointer(counts_dev, dev, "counts_dev")) return rc;
    if (int rc = requireDevicePointer(pairs_dev, dev, "pairs_dev")) return rc;
    if (streams_dev)
        if (int rc = requireDevicePointer(streams_dev, dev, "streams_dev")) return rc;
    uint32_t* scratch = nullptr;
    if (int rc = scratchOfCurrentDevice(dev, &scratch)) return rc;
    cudaStream_t st = static_cast<cudaStream_t>(cuda_stream);

    BackendBuffers b{};
    b.counts = counts_dev;
    b.pairs = reinterpret_cast<const uint32_t*>(pairs_dev);  // {deltaId, intCoeff} u16 pair == one little-endian u32
    b.K = K;
    b.nblocks = static_cast<uint32_t>(n_blocks);
    b.ntiles = static_cast<uint32_t>((n_blocks + kTileBlocks - 1) / kTileBlocks);
    b.tileCounts = scratch;                  // [ntiles][3K], fully rewritten by tileCountsKernel
    b.totals = scratch + kStaticTotalsAt;    // [3K], fully rewritten by tileScanKernel
    b.arena = streams_dev;

    launchTileCounts(b, st);
    CU(cudaGetLastError());
    uint32_t totals[3 * kMaxK];
    CU(cudaMemcpyAsync(totals, b.totals, sizeof(uint32_t) * 3 * K, cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    // stream 2 * (layer * K + depth) = deltaIds, + 1 = intCoeffs; both have one entry per block with count > depth
    StreamOffsets so{};
    uint64_t at = 0;
    for (int col = 0; col < 3 * K; ++col)
        for (int half = 0; half < 2; ++half) {
            stream_offsets[2 * col + half] = at;
            so.off[2 * col + half] = static_cast<uint32_t>(at);
            at += totals[col];
        }
    stream_offsets[nStreams] = at;
    so.off[nStreams] = static_cast<uint32_t>(at);
    if (at == 0) return MNENC_OK;  // every count is zero: all streams empty
    if (!streams_dev) return fail(MNENC_ERR_ARG, "null argument");
    if (at > streams_capacity)
        return fail(MNENC_ERR_ARG, "streams_dev holds " + std::to_string(streams_capacity) + " symbols, the streams need " +
                                       std::to_string(at) + " (stream_offsets has been filled in; nothing was written)");
    launchScatterRaw(b, so, st);
    CU(cudaGetLastError());
    CU(cudaStreamSynchronize(st));
    return MNENC_OK;
} catch (...) {
    return mnenc::abiCaught();
}
