// Element i of the [128, 128] bf16 weight operand of the sliding-operand conv1 (csrc/conv1_sliding.cu):
// row m = (column parity, channel), column k = (MMA step, half, pixel-in-pair p, input channel c).
#pragma once
#include <cuda_bf16.h>

namespace tadnn {

// w: Torch7 conv1 weight [64, 75] fp32 (k = (c*5+kh)*5+kw), bias [64]
__device__ __forceinline__ float conv1s_weight_value(const float* __restrict__ w, const float* __restrict__ bias, int i) {
  const int m = i >> 7, k = i & 127;
  const int par = m >> 6, ch = m & 63;
  const int st = k >> 4, half = (k >> 3) & 1, e = k & 7;
  int kh, j;
  if (st < 5) { kh = st; j = half; }
  else if (st == 5) { kh = half; j = 2; }
  else if (st == 6) { kh = 2 + half; j = 2; }
  else { kh = 4; j = half ? -1 : 2; }
  float v = 0.f;
  if (j >= 0) {
    const int p = e >> 2, c = e & 3;
    const int kw = 2 * j + p - par;
    if (kw >= 0 && kw <= 4) {
      if (c < 3) v = w[ch * 75 + (c * 5 + kh) * 5 + kw];
      else if (kh == 0 && kw == 0) v = bias[ch];
    }
  }
  return v;
}

}  // namespace tadnn
