// NCCL, resolved at run time (dlopen of libnccl.so.2): the library has no link-time dependency on it, a process that already
// carries an NCCL (torch's bundled copy has the same soname) shares that instance, and single-GPU users never load it.
// Only the stripe path (stripe_comm.cu: ONE image sharded by block columns over the GPUs) talks to NCCL.
#pragma once

#include <cuda_runtime.h>
#include <nccl.h>  // types and enums only

#include <string>

namespace mnenc {

struct NcclApi {
    ncclResult_t (*getUniqueId)(ncclUniqueId*);
    ncclResult_t (*commInitRank)(ncclComm_t*, int, ncclUniqueId, int);
    ncclResult_t (*commInitAll)(ncclComm_t*, int, const int*);
    ncclResult_t (*commDestroy)(ncclComm_t);
    ncclResult_t (*allGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t);
    ncclResult_t (*allReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t);
    ncclResult_t (*reduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, int, ncclComm_t, cudaStream_t);
    ncclResult_t (*send)(const void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t);
    ncclResult_t (*recv)(void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t);
    ncclResult_t (*groupStart)();
    ncclResult_t (*groupEnd)();
    const char* (*getErrorString)(ncclResult_t);
};

// nullptr (+ text in *err) when libnccl.so.2 cannot be loaded or lacks a symbol.
const NcclApi* ncclApi(std::string* err);

}  // namespace mnenc
