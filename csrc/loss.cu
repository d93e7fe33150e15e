// tadnn native core — fused mean-squared-error criterion for wide bf16 outputs.
//
// Reference: nn.MSECriterion under StochasticGradient (datamodule.lua:139-141: criterion:forward, then
// criterion:updateGradInput).  On a [4096, 16384] model-parallel output the stock sequence
// (cast to fp32, subtract, square, mean, and the mirror image in backward) moves about 2 GB through HBM per
// step; this kernel reads the output and the target once and writes the loss and the bf16 gradient once.
//
//   loss  = scale_loss * sum_i (y_i - t_i)^2          (scale_loss = 1/n for sizeAverage)
//   grad_i = scale_grad * (y_i - t_i)                 (scale_grad = 2/n)
//
// y: bf16, t: bf16 or fp32, grad: bf16 (may alias nothing), loss: one fp32 (accumulated with atomicAdd; the
// entry point zeroes it first).
#include "common.cuh"

namespace tadnn {
namespace {

template <bool T_F32>
__global__ void __launch_bounds__(256) mse_loss_grad_kernel(const __nv_bfloat16* __restrict__ y, const void* __restrict__ t_,
                                                            __nv_bfloat16* __restrict__ grad, float* __restrict__ loss, long long n,
                                                            float scale_loss, float scale_grad) {
  const long long nvec = n >> 3;
  float acc = 0.f;
  for (long long i = blockIdx.x * 256ll + threadIdx.x; i < nvec; i += 256ll * gridDim.x) {
    const uint4 yv = *reinterpret_cast<const uint4*>(y + i * 8);
    float yf[8], tf[8];
    float2 f;
    f = unpack_bf16x2(yv.x); yf[0] = f.x; yf[1] = f.y;
    f = unpack_bf16x2(yv.y); yf[2] = f.x; yf[3] = f.y;
    f = unpack_bf16x2(yv.z); yf[4] = f.x; yf[5] = f.y;
    f = unpack_bf16x2(yv.w); yf[6] = f.x; yf[7] = f.y;
    if (T_F32) {
      const float4 a = *reinterpret_cast<const float4*>(static_cast<const float*>(t_) + i * 8);
      const float4 b = *reinterpret_cast<const float4*>(static_cast<const float*>(t_) + i * 8 + 4);
      tf[0] = a.x; tf[1] = a.y; tf[2] = a.z; tf[3] = a.w; tf[4] = b.x; tf[5] = b.y; tf[6] = b.z; tf[7] = b.w;
    } else {
      const uint4 tv = *reinterpret_cast<const uint4*>(static_cast<const __nv_bfloat16*>(t_) + i * 8);
      f = unpack_bf16x2(tv.x); tf[0] = f.x; tf[1] = f.y;
      f = unpack_bf16x2(tv.y); tf[2] = f.x; tf[3] = f.y;
      f = unpack_bf16x2(tv.z); tf[4] = f.x; tf[5] = f.y;
      f = unpack_bf16x2(tv.w); tf[6] = f.x; tf[7] = f.y;
    }
    float d[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      d[e] = yf[e] - tf[e];
      acc = fmaf(d[e], d[e], acc);
    }
    if (grad != nullptr) {
      uint4 g;
      g.x = pack_bf16x2(d[0] * scale_grad, d[1] * scale_grad);
      g.y = pack_bf16x2(d[2] * scale_grad, d[3] * scale_grad);
      g.z = pack_bf16x2(d[4] * scale_grad, d[5] * scale_grad);
      g.w = pack_bf16x2(d[6] * scale_grad, d[7] * scale_grad);
      *reinterpret_cast<uint4*>(grad + i * 8) = g;
    }
  }
  // tail (n not a multiple of 8): the first threads of block 0, one element each
  if (blockIdx.x == 0) {
    const long long i = (nvec << 3) + threadIdx.x;
    if (threadIdx.x < 8 && i < n) {
      const float tv = T_F32 ? static_cast<const float*>(t_)[i] : bf16_to_f32(static_cast<const __nv_bfloat16*>(t_)[i]);
      const float dd = bf16_to_f32(y[i]) - tv;
      acc = fmaf(dd, dd, acc);
      if (grad != nullptr) grad[i] = __float2bfloat16_rn(dd * scale_grad);
    }
  }
  __shared__ float red[8];
  acc = warp_sum(acc);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    float s = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) s += red[w];
    atomicAdd(loss, s * scale_loss);
  }
}

}  // namespace
}  // namespace tadnn

using namespace tadnn;

extern "C" int tadnn_mse_loss_grad(const void* y, const void* target, int target_fp32, void* grad, float* loss, long long n,
                                   float scale_loss, float scale_grad, tadnn_stream_t stream_) {
  TADNN_REQUIRE(y && target && loss && n > 0, "mse_loss_grad: bad arguments");
  const uintptr_t al = reinterpret_cast<uintptr_t>(y) | reinterpret_cast<uintptr_t>(grad) | reinterpret_cast<uintptr_t>(target);
  TADNN_REQUIRE((al & 15) == 0, "mse_loss_grad: y, target and grad must be 16-byte aligned");
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  TADNN_CHECK_CUDA(cudaMemsetAsync(loss, 0, sizeof(float), stream));
  long long blocks = ((n >> 3) + 255) / 256;
  if (blocks > 8ll * sm_count()) blocks = 8ll * sm_count();
  if (blocks < 1) blocks = 1;
  if (target_fp32)
    mse_loss_grad_kernel<true><<<static_cast<int>(blocks), 256, 0, stream>>>(static_cast<const __nv_bfloat16*>(y), target,
                                                                             static_cast<__nv_bfloat16*>(grad), loss, n, scale_loss, scale_grad);
  else
    mse_loss_grad_kernel<false><<<static_cast<int>(blocks), 256, 0, stream>>>(static_cast<const __nv_bfloat16*>(y), target,
                                                                              static_cast<__nv_bfloat16*>(grad), loss, n, scale_loss, scale_grad);
  TADNN_LAUNCH_CHECK("mse_loss_grad");
  return 0;
}
