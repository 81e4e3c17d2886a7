// Host: operand images of the imported dictionary (mnenc_create). The fp16 images feed the tensor-core pre-filter
// (tcgen05.mma kind::f16), the exact FP32 row table feeds the oracle-exact re-score and the residual update.
#include <cuda_fp16.h>

#include <algorithm>
#include <cmath>

#include "mnenc_internal.h"

namespace mnenc {

// element index in a canonical K-major (no swizzle) UMMA image of fp16: 8-row groups of 1 KB, 8-element K chunks of 128 B
static inline size_t umOff16(int row, int k) { return static_cast<size_t>(row >> 3) * 512 + (k >> 3) * 64 + (row & 7) * 8 + (k & 7); }

void buildF16Images(const float* baseRowMajor, int nBase, const int32_t* detailRows, const float* const detail[3], uint16_t* baseUm16,
                    uint16_t* detUm16, float* rows32, double* maxRoundingNorm, double* maxNorm) {
    double worst = 0.0, worstNorm = 0.0;
    auto put = [&](uint16_t* image, int row, const float* src, float* exactRow) {
        double err = 0.0, nrm = 0.0;
        for (int k = 0; k < kN; ++k) {
            const __half h = __float2half_rn(src[k]);
            image[umOff16(row, k)] = __half_as_ushort(h);
            const double e = static_cast<double>(src[k]) - static_cast<double>(__half2float(h));
            err += e * e;
            nrm += static_cast<double>(src[k]) * src[k];
            exactRow[k] = src[k];
        }
        worst = std::max(worst, std::sqrt(err));
        worstNorm = std::max(worstNorm, std::sqrt(nrm));
    };
    std::fill(baseUm16, baseUm16 + static_cast<size_t>(kRowsPad) * kN, uint16_t(0));
    std::fill(rows32, rows32 + (static_cast<size_t>(kRowsPad) + static_cast<size_t>(3) * nBase * 64) * kN, 0.0f);
    std::fill(detUm16, detUm16 + static_cast<size_t>(3) * nBase * 64 * kN, uint16_t(0));
    for (int a = 0; a < nBase; ++a) put(baseUm16, a, baseRowMajor + static_cast<size_t>(a) * kN, rows32 + static_cast<size_t>(a) * kN);
    for (int ch = 0; ch < 3; ++ch) {
        size_t rowOff = 0;
        for (int a = 0; a < nBase; ++a) {
            const size_t g = static_cast<size_t>(ch) * nBase + a;
            for (int r = 0; r < detailRows[a]; ++r)
                put(detUm16 + g * 64 * kN, r, detail[ch] + (rowOff + r) * kN, rows32 + (static_cast<size_t>(kRowsPad) + g * 64 + r) * kN);
            rowOff += detailRows[a];
        }
    }
    *maxRoundingNorm = worst;
    *maxNorm = worstNorm;
}

size_t f16Rows32Count(int nBase) { return static_cast<size_t>(kRowsPad) + static_cast<size_t>(3) * nBase * 64; }

}  // namespace mnenc
