// Private definition of the encoder context, shared by the C-ABI translation units (mnenc_api.cu: one image on one
// GPU; batch_api.cu: pipelined batches; stripe_api.cu: one image sharded by block columns over the ranks of a process
// group). Not installed; the public boundary is include/mnenc.h.
#pragma once

#include <cuda_runtime.h>
#include <sched.h>
#include <nvtx3/nvToolsExt.h>

#include <algorithm>
#include <chrono>
#include <cstddef>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <string>
#include <vector>

#include "../../include/mnenc.h"
#include "backend_kernels.cuh"
#include "host_pool.h"
#include "mnenc_internal.h"

namespace mnenc {

int fail(int code, const std::string& msg);  // records the thread-local error text and returns `code`
int abiCaught() noexcept;                    // inside a catch (...) of a C-ABI entry point: the exception becomes an error code + text

#define CU(expr)                                                                                                          \
    do {                                                                                                                  \
        cudaError_t e_ = (expr);                                                                                          \
        if (e_ != cudaSuccess) return ::mnenc::fail(MNENC_ERR_CUDA, std::string(#expr) + ": " + cudaGetErrorString(e_)); \
    } while (0)

// NVTX range for the timeline tools (Nsight Systems; header-only NVTX 3: a no-op unless a tool is attached)
struct NvtxRange {
    explicit NvtxRange(const char* name) { nvtxRangePushA(name); }
    ~NvtxRange() { nvtxRangePop(); }
    NvtxRange(const NvtxRange&) = delete;
    NvtxRange& operator=(const NvtxRange&) = delete;
};

inline double nowMs() {
    return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

struct DevBuf {  // grow-only device buffer
    void* p = nullptr;
    size_t cap = 0;
    cudaError_t ensure(size_t bytes) {
        if (bytes <= cap) return cudaSuccess;
        if (!owned) return cudaErrorInvalidValue;
        if (p) cudaFree(p);
        p = nullptr;
        cap = 0;
        size_t want = bytes + bytes / 4 + 256;
        cudaError_t e = cudaMalloc(&p, want);
        if (e == cudaSuccess) cap = want;
        return e;
    }
    template <class T>
    T* as() const { return static_cast<T*>(p); }
    void alias(const DevBuf& other) {  // non-owning view of another context's buffer (lanes share the dictionary)
        if (p && owned) cudaFree(p);
        p = other.p, cap = other.cap, owned = false;
    }
    bool owned = true;
    DevBuf() = default;
    DevBuf(const DevBuf&) = delete;
    DevBuf& operator=(const DevBuf&) = delete;
    ~DevBuf() {
        if (p && owned) cudaFree(p);
    }
};
struct PinBuf {  // grow-only pinned host buffer
    void* p = nullptr;
    size_t cap = 0;
    cudaError_t ensure(size_t bytes) {
        if (bytes <= cap) return cudaSuccess;
        if (p) cudaFreeHost(p);
        p = nullptr;
        cap = 0;
        size_t want = bytes + bytes / 4 + 256;
        cudaError_t e = cudaMallocHost(&p, want);
        if (e == cudaSuccess) cap = want;
        return e;
    }
    template <class T>
    T* as() const { return static_cast<T*>(p); }
    PinBuf() = default;
    PinBuf(const PinBuf&) = delete;
    PinBuf& operator=(const PinBuf&) = delete;
    ~PinBuf() {
        if (p) cudaFreeHost(p);
    }
};

struct Overlay {  // host-written bits (headers, EOF codes) OR-ed into the packed payload
    uint64_t bitPos;
    BitString bits;
};
// ORs the overlays into the byte stream `res` (MSB-first bit positions).
void mergeOverlays(uint8_t* res, const std::vector<Overlay>& overlays);

// Result buffers (the `.mn` stream handed to the caller) come from a process-wide pool of PINNED host buffers, recycled by
// mnenc_free(): the D2H copy of the packed stream lands directly in the caller's buffer — no staging copy, no page faults
// of a fresh malloc (for a 56 MB stream those cost more than the whole back end).
uint8_t* resultAcquire(size_t bytes);  // nullptr when out of memory
void resultRelease(uint8_t* p);        // pool buffer -> back to the pool; anything else -> free()

// State of one image sharded over ranks (stripe_api.cu), between the phases of the stripe protocol.
struct StripeState;
// NCCL communicator + staging buffers of the in-library stripe exchanges (stripe_comm.cu).
struct StripeComm;
void destroyStripeComm(StripeComm* sc);

}  // namespace mnenc

struct mnenc_ctx {
    int device = 0, K = 0, nBase = 0, numSMs = 0;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev[8] = {};
    cudaEvent_t evSync = nullptr;  // waitStream()
    bool quantSet = false;
    float hQuant[3 * mnenc::kMaxK] = {};
    // dictionary
    mnenc::DevBuf detRows, quant, baseUm16, detUm16, rows32;
    float epsAbs16 = 0.0f, epsRel16Hi = 0.0f;  // pre-filter bound: eps = epsRel16Hi * ||r|| + epsAbs16 (mnenc_create)
    // image + symbols
    mnenc::DevBuf rgb, counts, pairs, mpScratch;
    // level-synchronous matching pursuit (mp_wave.cu): per block-channel state, work lists, (channel, atom) bins
    mnenc::DevBuf wRes32, wRes16, wState, wGroups, wBest, wAmax, wNcand, wCand, wNcandDetail, wCandDetail, wAct0, wAct1, wBins, wMembers, wMemberInfo, wOvf;
    mnenc::PinBuf hWaveAct;              // active block-channels per iteration, read back while later iterations are queued
    cudaEvent_t evWave[mnenc::kMaxK] = {};
    size_t lastW = 0, lastH = 0;
    bool haveSymbols = false;
    // back end
    mnenc::DevBuf tileCounts, totals, arena, dcTmp, flags, runIdx, startPos, outOff, rleArena, sums32, sums64, seqTable, hist, firstPos,
        maxSym, entries, nEntries, entryCodes, codeTable, outWords;
    uint32_t entryCap = 0;
    bool histReady = false;
    bool histDirty = false;  // the histogram holds counts of an encode that has not reset them yet
    mnenc::PinBuf hTotals, hSeqTable, hEntries, hNEntries, hEntryCodes, hCodeTable, hStats;
    std::unique_ptr<mnenc::HostPool> pool;  // code-table builds of one image run on these threads
    // decoder (decode_api.cu): parsed `lengths` / streams in pinned memory, the stream's quantisers + error flag on the device
    mnenc::PinBuf hDecCounts, hDecArena, hDecMisc;
    mnenc::DevBuf decQuant;
    // accounting
    float ms[8] = {};
    int launches = 0;
    double mpFlops = 0;
    uint64_t nSymbols = 0, nBits = 0;
    int planHuffman = 0, planGolomb = 0, planRle = 0;  // coded sequences by mode, symbol streams that were run-length coded
    // stripe protocol
    std::unique_ptr<mnenc::StripeState, void (*)(mnenc::StripeState*)> stripe{nullptr, nullptr};
    std::unique_ptr<mnenc::StripeComm, void (*)(mnenc::StripeComm*)> comm{nullptr, mnenc::destroyStripeComm};
    bool stripeInLibrary = false;      // set by the in-library striped encode for the length of the call: mnenc_stripe_plan deals the table builds over the communicator
    uint8_t* stripeResult = nullptr;   // striped encode, rank 0: result buffer already filled by the chunked reduce + D2H (stripe_comm.cu)
    cudaStream_t copyStream = nullptr; // second stream of that pipeline, created on first use
    cudaEvent_t evChunk[4] = {};
    bool failNextBackend = false;      // test hook: the next back end returns an error right after its histogram launch
    uint32_t frameBc = 0;      // block-channel offset of the frame the back end works on (frames that shared one pursuit)
    mnenc::DevBuf frameTab;    // device pointers of the frames of a group
    mnenc::PinBuf hFrameTab;
    std::vector<mnenc_ctx*> groupPeers;  // mnenc_create_multi: the contexts of the other GPUs (this one is rank 0)
    // batch mode (batch_api.cu): extra lanes = contexts with their own stream and buffers that alias this dictionary
    std::vector<mnenc_ctx*> lanes;
    int hostThreads = 1;
    bool inBatch = false;  // set on every lane for the duration of a batch call: per-image host parallelism off, the lanes are the parallelism
    std::string laneError;
};

namespace mnenc {
// Host wait for the context's stream that YIELDS the core while the GPU works. With several lanes per GPU and one process
// per GPU there are more waiting threads than cores; cudaStreamSynchronize spins, and the spinning waiters starved the
// threads that build the code tables (8 replicas on a 32-core box scaled 0.89 before this).
inline cudaError_t waitStream(mnenc_ctx* c) {
    if (!c->evSync) return cudaStreamSynchronize(c->stream);
    cudaError_t e = cudaEventRecord(c->evSync, c->stream);
    if (e != cudaSuccess) return e;
    for (int spins = 0; (e = cudaEventQuery(c->evSync)) == cudaErrorNotReady; ++spins)
        if (spins >= 32) sched_yield();
    return e;
}

// Pieces of the single-GPU encode that the other translation units reuse (mnenc_api.cu).
int ensureHistogram(mnenc_ctx* c);
int runMatchingPursuit(mnenc_ctx* c, const uint8_t* rgbDev, size_t width, size_t height, size_t stride,
                       const uint8_t* const* framesDev = nullptr, uint32_t nFrames = 1);
uint32_t frameBlockSlots(size_t width, size_t height);
int encodeFrames(mnenc_ctx* c, const uint8_t* const* rgb, int onDevice, size_t n, size_t w, size_t h, const size_t* strides, uint8_t** out,
                 size_t* outBytes);
void finishMpAccounting(mnenc_ctx* c);
int checkEncodeArgs(mnenc_ctx* c, const void* rgb, size_t w, size_t h, size_t stride, uint8_t** out, size_t* outBytes);
// stripe_comm.cu / stripe_api.cu: exchanges of the stripe protocol inside the library
// localRc: result of this rank's local work since the previous exchange. It travels with the exchange; when any rank reports a
// failure EVERY rank returns an error after the same collective instead of leaving its peers inside the next one.
int stripeExchangeHistogram(mnenc_ctx* c, int localRc);
int exchangeHistogramSparse(mnenc_ctx* c, BackendBuffers* b, int nSeq, int localRc);
int planSharded(mnenc_ctx* c, const std::vector<std::vector<SymbolStat>>& stats, std::vector<SequencePlan>& plans, int localRc);
int stripeRanks(const mnenc_ctx* c);  // ranks of the context's communicator (1 without one)
size_t stripeTotalBytes(const mnenc_ctx* c);
int encodeGroup(mnenc_ctx* c, const uint8_t* rgb, size_t width, size_t height, size_t stride, uint8_t** out, size_t* outBytes);
}  // namespace mnenc
