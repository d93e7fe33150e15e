// tadnn native core — dataset preprocessing of the CIFAR example on the GPU.
//
// Reference: cifar_example/sgd-torchad_nn-cifar.lua:182-216 — image.rgb2yuv per image, global
// standardisation of U and V with the training shard's statistics, and
// nn.SpatialContrastiveNormalization(1, image.gaussian1D(7)) on the Y channel.  The reference does
// this on the CPU one image at a time; here the whole shard is one pass over HBM:
//
//   preprocess_uv_stats_kernel : sum / sum-of-squares of U and V over the shard (fp64 atomics)
//   preprocess_yuv_lcn_kernel  : one CTA per image; RGB is read once, Y and the centred Y live in
//                                shared memory, the two 7x7 neighbourhood averages are computed from
//                                there with Torch7's arithmetic (border coefficient = the estimator
//                                applied to ones; divisor = sqrt(local mean of squares) / coefficient,
//                                replaced by `thresval` where it does not exceed `threshold`), U and V
//                                are standardised with the statistics passed in, and the three planes
//                                are written once (fp32 or bf16).
//
// Off the training hot path (runs once per dataset), so it is bandwidth-shaped, not tensor-core work.
#include "common.cuh"

namespace tadnn {
namespace {

constexpr int kMaxTaps = 15;

struct PreTaps {
  float k[kMaxTaps];  // normalised 1-D taps: outer(k, k) sums to 1
  int n;
};

__device__ __forceinline__ void rgb_to_yuv(float r, float g, float b, float& y, float& u, float& v) {
  y = 0.299f * r + 0.587f * g + 0.114f * b;
  u = -0.14713f * r - 0.28886f * g + 0.436f * b;
  v = 0.615f * r - 0.51499f * g - 0.10001f * b;
}

// stats[0..3] = sum u, sum u^2, sum v, sum v^2
__global__ void __launch_bounds__(256) preprocess_uv_stats_kernel(const float* __restrict__ rgb, double* __restrict__ stats,
                                                                  long long n_img, int hw) {
  double su = 0, suu = 0, sv = 0, svv = 0;
  const long long total = n_img * hw;
  for (long long i = blockIdx.x * 256ll + threadIdx.x; i < total; i += 256ll * gridDim.x) {
    const long long img = i / hw;
    const int p = static_cast<int>(i - img * hw);
    const float* base = rgb + img * 3 * hw + p;
    float y, u, v;
    rgb_to_yuv(base[0], base[hw], base[2 * hw], y, u, v);
    su += u; suu += static_cast<double>(u) * u;
    sv += v; svv += static_cast<double>(v) * v;
  }
  __shared__ double red[4][8];
  double vals[4] = {su, suu, sv, svv};
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    double x = vals[j];
    for (int o = 16; o > 0; o >>= 1) x += __shfl_xor_sync(0xffffffffu, x, o);
    if ((threadIdx.x & 31) == 0) red[j][threadIdx.x >> 5] = x;
  }
  __syncthreads();
  if (threadIdx.x < 4) {
    double x = 0;
    for (int w = 0; w < 8; ++w) x += red[threadIdx.x][w];
    atomicAdd(stats + threadIdx.x, x);
  }
}

template <bool OUT_BF16>
__global__ void __launch_bounds__(256) preprocess_yuv_lcn_kernel(const float* __restrict__ rgb, void* __restrict__ out_, int H, int W,
                                                                 PreTaps taps, float mean_u, float inv_std_u, float mean_v,
                                                                 float inv_std_v, float threshold, float thresval) {
  extern __shared__ float sm[];
  const int hw = H * W;
  float* ys = sm;            // Y
  float* cs = sm + hw;       // centred Y
  float* c1h = sm + 2 * hw;  // 1-D border coefficients along H
  float* c1w = c1h + H;      // ... and along W
  const int pad = taps.n / 2;
  const long long img = blockIdx.x;
  const float* src = rgb + img * 3 * hw;

  for (int p = threadIdx.x; p < hw; p += 256) {
    float y, u, v;
    rgb_to_yuv(src[p], src[hw + p], src[2 * hw + p], y, u, v);
    ys[p] = y;
    u = (u - mean_u) * inv_std_u;
    v = (v - mean_v) * inv_std_v;
    if (OUT_BF16) {
      __nv_bfloat16* o = static_cast<__nv_bfloat16*>(out_) + img * 3 * hw;
      o[hw + p] = __float2bfloat16(u);
      o[2 * hw + p] = __float2bfloat16(v);
    } else {
      float* o = static_cast<float*>(out_) + img * 3 * hw;
      o[hw + p] = u;
      o[2 * hw + p] = v;
    }
  }
  for (int i = threadIdx.x; i < H + W; i += 256) {
    const int n = i < H ? H : W, c = i < H ? i : i - H;
    float s = 0.f;
    for (int t = 0; t < taps.n; ++t) {
      const int q = c + t - pad;
      if (q >= 0 && q < n) s += taps.k[t];
    }
    c1h[i] = s;
  }
  __syncthreads();

  // subtractive step: centred = y - (k2 * y) / coef
  for (int p = threadIdx.x; p < hw; p += 256) {
    const int i = p / W, j = p - i * W;
    float acc = 0.f;
    for (int a = 0; a < taps.n; ++a) {
      const int ii = i + a - pad;
      if (ii < 0 || ii >= H) continue;
      float row = 0.f;
      for (int b = 0; b < taps.n; ++b) {
        const int jj = j + b - pad;
        if (jj >= 0 && jj < W) row += taps.k[b] * ys[ii * W + jj];
      }
      acc += taps.k[a] * row;
    }
    cs[p] = ys[p] - acc / (c1h[i] * c1w[j]);
  }
  __syncthreads();

  // divisive step (Torch7 nn.SpatialDivisiveNormalization): std = sqrt(k2 * centred^2) / coef -- the square root
  // is divided by the border coefficient -- and nn.Threshold(threshold, thresval) element-wise on the divisor
  for (int p = threadIdx.x; p < hw; p += 256) {
    const int i = p / W, j = p - i * W;
    float acc = 0.f;
    for (int a = 0; a < taps.n; ++a) {
      const int ii = i + a - pad;
      if (ii < 0 || ii >= H) continue;
      float row = 0.f;
      for (int b = 0; b < taps.n; ++b) {
        const int jj = j + b - pad;
        if (jj >= 0 && jj < W) {
          const float c = cs[ii * W + jj];
          row += taps.k[b] * c * c;
        }
      }
      acc += taps.k[a] * row;
    }
    const float sd = sqrtf(fmaxf(acc, 0.f)) / (c1h[i] * c1w[j]);
    const float y = cs[p] / (sd > threshold ? sd : thresval);
    if (OUT_BF16) static_cast<__nv_bfloat16*>(out_)[img * 3 * hw + p] = __float2bfloat16(y);
    else static_cast<float*>(out_)[img * 3 * hw + p] = y;
  }
}

// uint8 -> bf16 with a per-channel affine map (y = x * scale[c] + shift[c]); 16 pixels per thread.
// This is what lets the training loop ship the dataset over PCIe as what CIFAR is on disk (one byte per
// value, 3 KB per image) and widen it on the device, instead of copying bf16/fp32 tensors every step.
struct U8Affine {
  float scale[4], shift[4];
};
__global__ void __launch_bounds__(256) u8_to_bf16_kernel(const uint4* __restrict__ x, uint4* __restrict__ y, long long nvec, int plane16,
                                                         int C, U8Affine m) {
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < nvec;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int c = static_cast<int>((i / plane16) % C);
    const float sc = m.scale[c], sh = m.shift[c];
    const uint4 v = __ldg(x + i);
    const uint32_t w[4] = {v.x, v.y, v.z, v.w};
    uint32_t o[8];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const float f0 = static_cast<float>(w[k] & 0xffu) * sc + sh, f1 = static_cast<float>((w[k] >> 8) & 0xffu) * sc + sh;
      const float f2 = static_cast<float>((w[k] >> 16) & 0xffu) * sc + sh, f3 = static_cast<float>(w[k] >> 24) * sc + sh;
      o[2 * k] = pack_bf16x2(f0, f1);
      o[2 * k + 1] = pack_bf16x2(f2, f3);
    }
    y[2 * i] = make_uint4(o[0], o[1], o[2], o[3]);
    y[2 * i + 1] = make_uint4(o[4], o[5], o[6], o[7]);
  }
}
__global__ void u8_to_bf16_scalar_kernel(const unsigned char* __restrict__ x, __nv_bfloat16* __restrict__ y, long long n, int plane, int C,
                                         U8Affine m) {
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < n;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int c = static_cast<int>((i / plane) % C);
    y[i] = __float2bfloat16_rn(static_cast<float>(x[i]) * m.scale[c] + m.shift[c]);
  }
}

}  // namespace
}  // namespace tadnn

using namespace tadnn;

extern "C" int tadnn_preprocess_uv_stats(const float* rgb, double* stats4, long long n_img, int H, int W, tadnn_stream_t stream_) {
  TADNN_REQUIRE(rgb && stats4 && n_img > 0 && H > 0 && W > 0, "preprocess_uv_stats: bad arguments");
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  TADNN_CHECK_CUDA(cudaMemsetAsync(stats4, 0, 4 * sizeof(double), stream));
  long long blocks = (n_img * H * W + 255) / 256;
  if (blocks > 8ll * sm_count()) blocks = 8ll * sm_count();
  preprocess_uv_stats_kernel<<<static_cast<int>(blocks), 256, 0, stream>>>(rgb, stats4, n_img, H * W);
  TADNN_LAUNCH_CHECK("preprocess_uv_stats");
  return 0;
}

extern "C" int tadnn_preprocess_yuv_lcn(const float* rgb, void* out, int out_bf16, long long n_img, int H, int W, const float* taps1d,
                                        int n_taps, float mean_u, float std_u, float mean_v, float std_v, float threshold,
                                        float thresval, tadnn_stream_t stream_) {
  TADNN_REQUIRE(rgb && out && taps1d && n_img > 0, "preprocess_yuv_lcn: bad arguments");
  TADNN_REQUIRE(n_taps >= 1 && n_taps <= kMaxTaps && (n_taps & 1), "preprocess_yuv_lcn: the neighbourhood needs an odd tap count <= %d (got %d)",
                kMaxTaps, n_taps);
  TADNN_REQUIRE(H > 0 && W > 0 && static_cast<long long>(H) * W <= 5120, "preprocess_yuv_lcn: image %dx%d exceeds the 5120-pixel shared-memory plan", H, W);
  TADNN_REQUIRE(std_u != 0.f && std_v != 0.f, "preprocess_yuv_lcn: zero standard deviation");
  PreTaps taps;
  double s = 0;
  for (int i = 0; i < n_taps; ++i) s += taps1d[i];
  TADNN_REQUIRE(s != 0, "preprocess_yuv_lcn: taps sum to zero");
  for (int i = 0; i < kMaxTaps; ++i) taps.k[i] = i < n_taps ? static_cast<float>(taps1d[i] / s) : 0.f;
  taps.n = n_taps;
  const size_t smem = (2ull * H * W + H + W) * sizeof(float);
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (out_bf16) {
    preprocess_yuv_lcn_kernel<true><<<static_cast<unsigned>(n_img), 256, smem, stream>>>(rgb, out, H, W, taps, mean_u, 1.f / std_u, mean_v,
                                                                                          1.f / std_v, threshold, thresval);
  } else {
    preprocess_yuv_lcn_kernel<false><<<static_cast<unsigned>(n_img), 256, smem, stream>>>(rgb, out, H, W, taps, mean_u, 1.f / std_u, mean_v,
                                                                                           1.f / std_v, threshold, thresval);
  }
  TADNN_LAUNCH_CHECK("preprocess_yuv_lcn");
  return 0;
}

extern "C" int tadnn_u8_to_bf16(const void* x, void* y, long long n, int plane, int C, const float* host_scale, const float* host_shift,
                                tadnn_stream_t stream_) {
  TADNN_REQUIRE(x && y && n > 0 && plane > 0 && C >= 1 && C <= 4, "u8_to_bf16: bad arguments (n=%lld plane=%d C=%d; C <= 4)", n, plane, C);
  U8Affine m;
  for (int c = 0; c < 4; ++c) {
    m.scale[c] = (host_scale && c < C) ? host_scale[c] : 1.f;
    m.shift[c] = (host_shift && c < C) ? host_shift[c] : 0.f;
  }
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  const bool vec = (plane % 16 == 0) && (n % 16 == 0) && ((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(y)) & 15) == 0;
  if (vec) {
    const long long nvec = n / 16;
    long long blocks = (nvec + 255) / 256;
    if (blocks > 8ll * sm_count()) blocks = 8ll * sm_count();
    u8_to_bf16_kernel<<<static_cast<int>(blocks), 256, 0, stream>>>(static_cast<const uint4*>(x), static_cast<uint4*>(y), nvec, plane / 16, C, m);
  } else {
    long long blocks = (n + 255) / 256;
    if (blocks > 8ll * sm_count()) blocks = 8ll * sm_count();
    u8_to_bf16_scalar_kernel<<<static_cast<int>(blocks), 256, 0, stream>>>(static_cast<const unsigned char*>(x), static_cast<__nv_bfloat16*>(y), n, plane, C, m);
  }
  TADNN_LAUNCH_CHECK("u8_to_bf16");
  return 0;
}
