// Stage geometry of the TMA rings (2048 row entries per stage) and the 32-bucket exponent histogram
// constants of the row-pair kernel.
#pragma once

#include "split_kernels.cuh"

namespace duodec {

constexpr int kStageElems = 2048;                 // row entries per ring stage: 8 per worker thread
constexpr int kStageBytes = kStageElems * 4;      // sized for fp32 rows; bf16 rows fill half of it
constexpr int kHB = 32;                           // histogram buckets
constexpr float kHTmin = -27.0f, kHTmax = 4.0f;   // bucket b holds round(t) == b - 27 (ends clamped)
constexpr float kHMagic = 12582912.0f;            // 1.5 * 2^23: t + magic rounds t to an integer in the low bits

struct MinMax {
  float mx, ngmn;  // max and max of the negated values
};
struct OpMinMax {
  __device__ __forceinline__ MinMax operator()(MinMax a, MinMax b) const {
    return MinMax{fmaxf(a.mx, b.mx), fmaxf(a.ngmn, b.ngmn)};
  }
};

__device__ __forceinline__ int hist_bits(float t) {
  return __float_as_int(fminf(fmaxf(t, kHTmin), kHTmax) + kHMagic);
}

}  // namespace duodec
