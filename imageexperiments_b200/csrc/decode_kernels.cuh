// Device side of the decoder — declarations (decode_kernels.cu).
#pragma once

#include <cuda_runtime.h>

#include <cstdint>

#include "backend_kernels.cuh"

namespace mnenc {

struct DecodeParams {
    const uint16_t* counts;   // [3 * nblocks] the `lengths` stream
    const uint32_t* pairs;    // [3 * nblocks][K] deltaId | intCoeff << 16, DC chains undone
    const float* quant;       // [3][K] from the stream header
    const float* rows32;      // [512 + 3 * nBase * 64][64] dictionary rows (WaveParams::rows32)
    const uint8_t* detRows;   // [nBase] rows of each detail group
    int K, nBase;
    uint32_t nblocks, nby;    // block index = bx * nby + by
    uint32_t width, height;
    uint8_t* rgb;             // packed RGB8, `stride` bytes per row
    size_t stride;
    unsigned int* error;      // set to 1 when an atom index points outside its block's dictionary
};

// arena streams -> dense pairs; needs b.tileCounts from launchTileCounts on the same counts.
int launchGather(const BackendBuffers& b, const StreamOffsets& so, uint32_t* pairsOut, cudaStream_t st);
int launchReconstruct(const DecodeParams& p, int numSMs, cudaStream_t st);

}  // namespace mnenc
