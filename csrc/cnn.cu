// CIFAR-CNN kernels around the tcgen05 GEMM (config 3; K15-K20 of SURVEY §2.5).
//
// The reference's example model (cifar_example/sgd-torchad_nn-cifar.lua:88-118) is
//   conv5x5(3->64)+ReLU+maxpool3/3, conv5x5(64->64)+ReLU+maxpool3/3, View(64), Dropout(.5),
//   Linear(64,100)+ReLU, Linear(100,10), LogSoftMax, CrossEntropyCriterion
// executed by cunn as im2col + cuBLAS SGEMM + pointwise kernels, fp32, one sample at a time.
// Here activations are bf16 NHWC, convolutions are im2col + tadnn_gemm_bf16 (bias+ReLU in the
// GEMM epilogue), and everything after the second pooling (dropout, both FC layers,
// log-softmax, NLL loss and all of their gradients) is ONE fused kernel per direction.
#include "common.cuh"
#include "sm100.cuh"

#include <mutex>

namespace tadnn {
namespace {

__device__ __forceinline__ uint32_t hash32(uint32_t x) {
  x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16;
  return x;
}
// Bernoulli(keep = 1-p) from a counter-based hash: recomputed in backward, no mask tensor.
__device__ __forceinline__ float dropout_scale(uint32_t seed, uint32_t idx, float p) {
  if (p <= 0.f) return 1.f;
  const float u = static_cast<float>(hash32(seed ^ hash32(idx)) >> 8) * (1.0f / 16777216.0f);
  return u >= p ? 1.0f / (1.0f - p) : 0.0f;
}

// ---- im2col ---------------------------------------------------------------
// col[row, (kh*KW+kw)*C + c], row = (b*OH + oy)*OW + ox, valid (unpadded) convolution, stride 1.
// NHWC source, C % 8 == 0: one 16-byte copy per (row, kh, kw, c8).
__global__ void im2col_nhwc_kernel(const uint4* __restrict__ x, uint4* __restrict__ col, int B, int H, int W, int C8,
                                   int KH, int KW, int OH, int OW, int Kpad8) {
  const long long total = static_cast<long long>(B) * OH * OW * KH * KW * C8;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int c8 = static_cast<int>(i % C8);
    long long t = i / C8;
    const int kw = static_cast<int>(t % KW); t /= KW;
    const int kh = static_cast<int>(t % KH); t /= KH;
    const long long row = t;
    const int ox = static_cast<int>(row % OW);
    const int oy = static_cast<int>((row / OW) % OH);
    const int b = static_cast<int>(row / (static_cast<long long>(OW) * OH));
    const uint4 v = x[((static_cast<long long>(b) * H + oy + kh) * W + ox + kw) * C8 + c8];
    col[row * Kpad8 + (kh * KW + kw) * C8 + c8] = v;
  }
}
// NCHW source with few channels (the 3-channel input image): each thread builds 8 consecutive
// k entries (one 16-byte store); K is zero-padded to Kpad.  Shapes are template constants for the
// CIFAR case so that every div/mod folds into multiply-shift.
template <int C, int H, int W, int KH, int KW>
__global__ void __launch_bounds__(256) im2col_nchw_fixed_kernel(const __nv_bfloat16* __restrict__ x, uint4* __restrict__ col,
                                                                 int B) {
  constexpr int OH = H - KH + 1, OW = W - KW + 1, K = KH * KW * C, Kpad8 = (K + 7) / 8;
  // one block = one image: stage it in shared memory, then emit its OH*OW rows
  __shared__ __nv_bfloat16 img[C * H * W];
  for (int b = blockIdx.x; b < B; b += gridDim.x) {
    const uint4* src = reinterpret_cast<const uint4*>(x + static_cast<long long>(b) * C * H * W);
    for (int i = threadIdx.x; i < C * H * W / 8; i += blockDim.x) reinterpret_cast<uint4*>(img)[i] = src[i];
    __syncthreads();
    uint4* dst = col + static_cast<long long>(b) * OH * OW * Kpad8;
    for (int i = threadIdx.x; i < OH * OW * Kpad8; i += blockDim.x) {
      const int k8 = i % Kpad8, pix = i / Kpad8;
      const int ox = pix % OW, oy = pix / OW;
      __align__(16) __nv_bfloat16 v[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const int k = k8 * 8 + j;
        const int c = k % C, kw = (k / C) % KW, kh = k / (C * KW);
        v[j] = (k < K) ? img[(c * H + oy + kh) * W + ox + kw] : __float2bfloat16_rn(0.f);
      }
      dst[i] = *reinterpret_cast<const uint4*>(v);
    }
    __syncthreads();
  }
}
__global__ void im2col_nchw_kernel(const __nv_bfloat16* __restrict__ x, uint4* __restrict__ col, int B, int C, int H,
                                   int W, int KH, int KW, int OH, int OW, int Kpad8) {
  const long long total = static_cast<long long>(B) * OH * OW * Kpad8;
  const int K = KH * KW * C;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int k8 = static_cast<int>(i % Kpad8);
    const long long row = i / Kpad8;
    const int ox = static_cast<int>(row % OW);
    const int oy = static_cast<int>((row / OW) % OH);
    const int b = static_cast<int>(row / (static_cast<long long>(OW) * OH));
    __align__(16) __nv_bfloat16 v[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int k = k8 * 8 + j;
      if (k < K) {
        const int c = k % C, kw = (k / C) % KW, kh = k / (C * KW);
        v[j] = x[((static_cast<long long>(b) * C + c) * H + oy + kh) * W + ox + kw];
      } else {
        v[j] = __float2bfloat16_rn(0.f);
      }
    }
    col[i] = *reinterpret_cast<const uint4*>(v);
  }
}

// ---- im2col with the pooled window first ---------------------------------------------------------
// Same patch layout as im2col_nhwc, but the output rows are ordered so that the win x win output
// pixels that the following max-pool actually reads come first ([B*win*win] rows, pixel-major per
// image), followed by all remaining pixels.  The backward pass only needs the first block: the
// gradient of every other conv output is exactly zero.
__global__ void im2col_nhwc_winfirst_kernel(const uint4* __restrict__ x, uint4* __restrict__ col, int B, int H, int W, int C8,
                                            int KH, int KW, int OH, int OW, int Kpad8, int win, int only_window) {
  const int npix = only_window ? win * win : OH * OW;
  const long long total = static_cast<long long>(B) * npix * KH * KW * C8;
  const int nwin = win * win, nrest = OH * OW - nwin;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int c8 = static_cast<int>(i % C8);
    long long t = i / C8;
    const int kw = static_cast<int>(t % KW); t /= KW;
    const int kh = static_cast<int>(t % KH); t /= KH;
    const int pix = static_cast<int>(t % npix);
    const int b = static_cast<int>(t / npix);
    int oy, ox;
    long long row;
    if (pix < nwin) {
      oy = pix / win; ox = pix % win;
      row = static_cast<long long>(b) * nwin + pix;
    } else {
      // enumerate the pixels outside the window in raster order
      int r = pix - nwin;                       // 0 .. nrest-1
      const int per_top_row = OW - win;         // pixels right of the window in each of the first `win` rows
      if (r < win * per_top_row) { oy = r / per_top_row; ox = win + r % per_top_row; }
      else { r -= win * per_top_row; oy = win + r / OW; ox = r % OW; }
      row = static_cast<long long>(B) * nwin + static_cast<long long>(b) * nrest + (pix - nwin);
    }
    const uint4 v = x[((static_cast<long long>(b) * H + oy + kh) * W + ox + kw) * C8 + c8];
    col[row * Kpad8 + (kh * KW + kw) * C8 + c8] = v;
  }
}

// Compile-time specialisation for the example's second conv (9x9x64 -> 5x5, 5x5 taps, window 3):
// every div/mod folds to a multiply-shift.
template <int H, int W, int C8, int KH, int KW, int WIN>
__global__ void __launch_bounds__(256) im2col_nhwc_winfirst_fixed_kernel(const uint4* __restrict__ x, uint4* __restrict__ col, int B,
                                                                          int only_window) {
  constexpr int OH = H - KH + 1, OW = W - KW + 1, NWIN = WIN * WIN, NREST = OH * OW - NWIN, TOP = OW - WIN;
  constexpr int KP8 = KH * KW * C8;
  const int npix = only_window ? NWIN : OH * OW;
  const long long total = static_cast<long long>(B) * npix * KP8;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int k8 = static_cast<int>(i % KP8);           // (tap, c8)
    const long long t = i / KP8;
    const int pix = static_cast<int>(t % npix);
    const int b = static_cast<int>(t / npix);
    const int c8 = k8 % C8, tap = k8 / C8, kw = tap % KW, kh = tap / KW;
    int oy, ox;
    long long row;
    if (pix < NWIN) {
      oy = pix / WIN; ox = pix % WIN;
      row = static_cast<long long>(b) * NWIN + pix;
    } else {
      int r = pix - NWIN;
      if (r < WIN * TOP) { oy = r / TOP; ox = WIN + r % TOP; }
      else { r -= WIN * TOP; oy = WIN + r / OW; ox = r % OW; }
      row = static_cast<long long>(B) * NWIN + static_cast<long long>(b) * NREST + (pix - NWIN);
    }
    col[row * KP8 + k8] = x[((static_cast<long long>(b) * H + oy + kh) * W + ox + kw) * C8 + c8];
  }
}

// ---- Toeplitz expansion of the packed conv weights for the transposed convolution ----------------
// The gradient of the conv output is non-zero only inside the win x win pooled window, so
//   dX[b, y, x, c] = sum_{(oy,ox) in window} sum_co dY[b,oy,ox,co] * Wp[co, ((y-oy)*KW + (x-ox))*C + c]
// is ONE dense GEMM  dX[B, (FH*FW*C)] = dYw[B, (win*win*O)] * T^T  with FH = win+KH-1 and
//   T[(y*FW+x)*C + c, (oy*win+ox)*O + co] = Wp[co, ((y-oy)*KW + (x-ox))*C + c]   (0 outside the kernel).
__global__ void conv_toeplitz_pack_kernel(const __nv_bfloat16* __restrict__ wp, __nv_bfloat16* __restrict__ T, int O, int C, int KH,
                                          int KW, int win) {
  const int FH = win + KH - 1, FW = win + KW - 1;
  const int Kt = win * win * O;
  const long long total = static_cast<long long>(FH) * FW * C * Kt;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int k = static_cast<int>(i % Kt);
    const int n = static_cast<int>(i / Kt);
    const int co = k % O, pos = k / O, oy = pos / win, ox = pos % win;
    const int c = n % C, pix = n / C, y = pix / FW, x = pix % FW;
    const int kh = y - oy, kw = x - ox;
    __nv_bfloat16 v = __float2bfloat16_rn(0.f);
    if (kh >= 0 && kh < KH && kw >= 0 && kw < KW) v = wp[static_cast<long long>(co) * (KH * KW * C) + (kh * KW + kw) * C + c];
    T[i] = v;
  }
}

// dst[B, H, W, C] = src[B, FH, FW, C] in the top-left corner, zero elsewhere (16-byte vectors)
__global__ void expand_corner_nhwc_kernel(const uint4* __restrict__ src, uint4* __restrict__ dst, int B, int H, int W, int FH, int FW,
                                          int C8) {
  const long long total = static_cast<long long>(B) * H * W * C8;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int c8 = static_cast<int>(i % C8);
    long long t = i / C8;
    const int x = static_cast<int>(t % W); t /= W;
    const int y = static_cast<int>(t % H);
    const int b = static_cast<int>(t / H);
    uint4 v = make_uint4(0, 0, 0, 0);
    if (y < FH && x < FW) v = src[((static_cast<long long>(b) * FH + y) * FW + x) * C8 + c8];
    dst[i] = v;
  }
}

// ---- col2im (gather form): dx[b,y,x,c] = sum_{kh,kw} dcol[(b,y-kh,x-kw), (kh*KW+kw)*C + c] ----
__global__ void col2im_nhwc_kernel(const uint4* __restrict__ dcol, uint4* __restrict__ dx, int B, int H, int W, int C8,
                                   int KH, int KW, int OH, int OW, int Kpad8) {
  const long long total = static_cast<long long>(B) * H * W * C8;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int c8 = static_cast<int>(i % C8);
    long long t = i / C8;
    const int xx = static_cast<int>(t % W); t /= W;
    const int yy = static_cast<int>(t % H);
    const int b = static_cast<int>(t / H);
    float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    for (int kh = 0; kh < KH; ++kh) {
      const int oy = yy - kh;
      if (oy < 0 || oy >= OH) continue;
      for (int kw = 0; kw < KW; ++kw) {
        const int ox = xx - kw;
        if (ox < 0 || ox >= OW) continue;
        const long long row = (static_cast<long long>(b) * OH + oy) * OW + ox;
        const uint4 v = dcol[row * Kpad8 + (kh * KW + kw) * C8 + c8];
        float2 f;
        f = unpack_bf16x2(v.x); acc[0] += f.x; acc[1] += f.y;
        f = unpack_bf16x2(v.y); acc[2] += f.x; acc[3] += f.y;
        f = unpack_bf16x2(v.z); acc[4] += f.x; acc[5] += f.y;
        f = unpack_bf16x2(v.w); acc[6] += f.x; acc[7] += f.y;
      }
    }
    dx[i] = make_uint4(pack_bf16x2(acc[0], acc[1]), pack_bf16x2(acc[2], acc[3]), pack_bf16x2(acc[4], acc[5]),
                       pack_bf16x2(acc[6], acc[7]));
  }
}

// ---- max pooling k x k, stride k, NHWC, floor mode; argmax kept as a byte -----------
__global__ void maxpool_fwd_nhwc_kernel(const uint4* __restrict__ x, uint4* __restrict__ y, uint2* __restrict__ idx, int B,
                                        int H, int W, int C8, int PK, int PH, int PW) {
  const long long total = static_cast<long long>(B) * PH * PW * C8;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int c8 = static_cast<int>(i % C8);
    long long t = i / C8;
    const int px = static_cast<int>(t % PW); t /= PW;
    const int py = static_cast<int>(t % PH);
    const int b = static_cast<int>(t / PH);
    float best[8];
    unsigned char bi[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) { best[j] = -INFINITY; bi[j] = 0; }
    for (int dy = 0; dy < PK; ++dy)
      for (int dx = 0; dx < PK; ++dx) {
        const uint4 v = x[((static_cast<long long>(b) * H + py * PK + dy) * W + px * PK + dx) * C8 + c8];
        float f[8];
        float2 q;
        q = unpack_bf16x2(v.x); f[0] = q.x; f[1] = q.y;
        q = unpack_bf16x2(v.y); f[2] = q.x; f[3] = q.y;
        q = unpack_bf16x2(v.z); f[4] = q.x; f[5] = q.y;
        q = unpack_bf16x2(v.w); f[6] = q.x; f[7] = q.y;
#pragma unroll
        for (int j = 0; j < 8; ++j)
          if (f[j] > best[j]) { best[j] = f[j]; bi[j] = static_cast<unsigned char>(dy * PK + dx); }
      }
    y[i] = make_uint4(pack_bf16x2(best[0], best[1]), pack_bf16x2(best[2], best[3]), pack_bf16x2(best[4], best[5]),
                      pack_bf16x2(best[6], best[7]));
    uint2 o;
    o.x = bi[0] | (bi[1] << 8) | (bi[2] << 16) | (bi[3] << 24);
    o.y = bi[4] | (bi[5] << 8) | (bi[6] << 16) | (bi[7] << 24);
    idx[i] = o;
  }
}

// dx[b,y,x,c] = dy[b,y/k,x/k,c] if (y,x) is the argmax of its window and the pooled value is > 0
// (ReLU backward folded in: the pooled value is the max of ReLU outputs), else 0.
__global__ void maxpool_relu_bwd_nhwc_kernel(const uint4* __restrict__ dy, const uint4* __restrict__ y,
                                             const uint2* __restrict__ idx, uint4* __restrict__ dx, int B, int H, int W,
                                             int C8, int PK, int PH, int PW, int relu) {
  const long long total = static_cast<long long>(B) * H * W * C8;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int c8 = static_cast<int>(i % C8);
    long long t = i / C8;
    const int xx = static_cast<int>(t % W); t /= W;
    const int yy = static_cast<int>(t % H);
    const int b = static_cast<int>(t / H);
    const int py = yy / PK, px = xx / PK;
    uint4 o = make_uint4(0, 0, 0, 0);
    if (py < PH && px < PW) {
      const long long pi = ((static_cast<long long>(b) * PH + py) * PW + px) * C8 + c8;
      const uint2 id = idx[pi];
      const uint4 g = dy[pi];
      const uint4 pv = y[pi];
      const unsigned me = static_cast<unsigned>((yy - py * PK) * PK + (xx - px * PK));
      const uint32_t gw[4] = {g.x, g.y, g.z, g.w};
      const uint32_t pw[4] = {pv.x, pv.y, pv.z, pv.w};
      uint32_t ow[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const unsigned i0 = ((j < 2 ? id.x : id.y) >> (16 * (j & 1))) & 0xffu;
        const unsigned i1 = ((j < 2 ? id.x : id.y) >> (16 * (j & 1) + 8)) & 0xffu;
        const float2 gv = unpack_bf16x2(gw[j]);
        const float2 yv = unpack_bf16x2(pw[j]);
        ow[j] = pack_bf16x2((i0 == me && (!relu || yv.x > 0.f)) ? gv.x : 0.f, (i1 == me && (!relu || yv.y > 0.f)) ? gv.y : 0.f);
      }
      o = make_uint4(ow[0], ow[1], ow[2], ow[3]);
    }
    dx[i] = o;
  }
}

// ---- column sum (bias gradient): out[c] += sum_r x[r, c]; x bf16 [R, C], C <= 256 ------------
__global__ void colsum_bf16_kernel(const __nv_bfloat16* __restrict__ x, float* __restrict__ out, long long R, int C, int ld) {
  // blockDim = (64, 4): threadIdx.x walks columns, threadIdx.y rows
  __shared__ float part[4][64];
  for (int c0 = 0; c0 < C; c0 += 64) {
    const int c = c0 + threadIdx.x;
    float acc = 0.f;
    if (c < C)
      for (long long r = blockIdx.x * 4ll + threadIdx.y; r < R; r += gridDim.x * 4ll)
        acc += __bfloat162float(x[r * ld + c]);
    part[threadIdx.y][threadIdx.x] = acc;
    __syncthreads();
    if (threadIdx.y == 0 && c < C)
      atomicAdd(out + c, part[0][threadIdx.x] + part[1][threadIdx.x] + part[2][threadIdx.x] + part[3][threadIdx.x]);
    __syncthreads();
  }
}

// Wide/tall variant: 16-byte loads, 2-D grid (64-column chunk x row split), one atomic per column
// per block.  blockDim = (8, 32): x = which 8-column vector of the chunk, y = row lane.
__global__ void __launch_bounds__(256) colsum_bf16_vec_kernel(const __nv_bfloat16* __restrict__ x, float* __restrict__ out,
                                                               long long R, int C, int ld) {
  __shared__ float part[32][65];
  const int c = blockIdx.x * 64 + threadIdx.x * 8;
  float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  if (c < C) {
    const long long rows_per = (R + gridDim.y - 1) / gridDim.y;
    const long long r0 = blockIdx.y * rows_per, r1 = min(R, r0 + rows_per);
    for (long long r = r0 + threadIdx.y; r < r1; r += 32) {
      const uint4 v = *reinterpret_cast<const uint4*>(x + r * ld + c);
      float2 f;
      f = unpack_bf16x2(v.x); acc[0] += f.x; acc[1] += f.y;
      f = unpack_bf16x2(v.y); acc[2] += f.x; acc[3] += f.y;
      f = unpack_bf16x2(v.z); acc[4] += f.x; acc[5] += f.y;
      f = unpack_bf16x2(v.w); acc[6] += f.x; acc[7] += f.y;
    }
  }
#pragma unroll
  for (int e = 0; e < 8; ++e) part[threadIdx.y][threadIdx.x * 8 + e] = acc[e];
  __syncthreads();
  const int t = threadIdx.y * 8 + threadIdx.x;
  if (t < 64 && blockIdx.x * 64 + t < C) {
    float sum = 0.f;
#pragma unroll 8
    for (int y = 0; y < 32; ++y) sum += part[y][t];
    atomicAdd(out + blockIdx.x * 64 + t, sum);
  }
}

// Bias gradient of conv+ReLU+maxpool taken at the POOLED resolution: the un-pooled gradient is
// zero except at each window's arg-max, so  sum_pixels d(conv out)[., c] = sum_windows dy[., c] * (y > 0).
// Reads k*k times less data than a column sum over the un-pooled gradient.
__global__ void colsum_relu_mask_kernel(const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ y,
                                        float* __restrict__ out, long long R, int C) {
  __shared__ float part[4][64];
  for (int c0 = 0; c0 < C; c0 += 64) {
    const int c = c0 + threadIdx.x;
    float acc = 0.f;
    if (c < C)
      for (long long r = blockIdx.x * 4ll + threadIdx.y; r < R; r += gridDim.x * 4ll)
        if (__bfloat162float(y[r * C + c]) > 0.f) acc += __bfloat162float(dy[r * C + c]);
    part[threadIdx.y][threadIdx.x] = acc;
    __syncthreads();
    if (threadIdx.y == 0 && c < C)
      atomicAdd(out + c, part[0][threadIdx.x] + part[1][threadIdx.x] + part[2][threadIdx.x] + part[3][threadIdx.x]);
    __syncthreads();
  }
}

// Vectorised, corner-restricted variant: dy is [B, FH, FW, C] (compact), y is [B, H, W, C].
// blockDim = (8, 32): x = 8-channel vector of a 64-channel chunk, y = row lane; grid.x = C/64 chunks.
__global__ void __launch_bounds__(256) colsum_relu_mask_corner_kernel(const __nv_bfloat16* __restrict__ dy,
                                                                       const __nv_bfloat16* __restrict__ y, float* __restrict__ out,
                                                                       int B, int H, int W, int FH, int FW, int C) {
  __shared__ float part[32][65];
  const int c = blockIdx.x * 64 + threadIdx.x * 8;
  float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  const long long R = static_cast<long long>(B) * FH * FW;
  if (c < C) {
    const long long rows_per = (R + gridDim.y - 1) / gridDim.y;
    const long long r0 = blockIdx.y * rows_per, r1 = min(R, r0 + rows_per);
    for (long long r = r0 + threadIdx.y; r < r1; r += 32) {
      const int fx = static_cast<int>(r % FW), fy = static_cast<int>((r / FW) % FH);
      const long long b = r / (static_cast<long long>(FW) * FH);
      const uint4 dv = *reinterpret_cast<const uint4*>(dy + r * C + c);
      const uint4 yv = *reinterpret_cast<const uint4*>(y + ((b * H + fy) * W + fx) * C + c);
      const uint32_t dw[4] = {dv.x, dv.y, dv.z, dv.w}, yw[4] = {yv.x, yv.y, yv.z, yv.w};
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const float2 d2 = unpack_bf16x2(dw[e]);
        if (yw[e] & 0xffffu) acc[2 * e] += d2.x;          // y is a ReLU output: > 0  <=>  bits != 0
        if (yw[e] & 0xffff0000u) acc[2 * e + 1] += d2.y;
      }
    }
  }
#pragma unroll
  for (int e = 0; e < 8; ++e) part[threadIdx.y][threadIdx.x * 8 + e] = acc[e];
  __syncthreads();
  const int t = threadIdx.y * 8 + threadIdx.x;
  if (t < 64 && blockIdx.x * 64 + t < C) {
    float sum = 0.f;
#pragma unroll 8
    for (int yy = 0; yy < 32; ++yy) sum += part[yy][t];
    atomicAdd(out + blockIdx.x * 64 + t, sum);
  }
}

// ---- conv weight layout/dtype conversion -----------------------------------------------------
// Torch7 SpatialConvolutionMM keeps weight [O, C*KH*KW] with k = (c*KH + kh)*KW + kw.  The GEMMs
// want bf16 [O, Kpad] with k' = (kh*KW + kw)*C + c (matches the NHWC im2col), zero padded.
__global__ void conv_weight_pack_kernel(const float* __restrict__ w, __nv_bfloat16* __restrict__ wp, int O, int C, int KH,
                                        int KW, int Kpad) {
  const int total = O * Kpad;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const int o = i / Kpad, k = i % Kpad;
    float v = 0.f;
    if (k < C * KH * KW) {
      const int c = k % C, kw = (k / C) % KW, kh = k / (C * KW);
      v = w[o * (C * KH * KW) + (c * KH + kh) * KW + kw];
    }
    wp[i] = __float2bfloat16_rn(v);
  }
}
// grad[O, C*KH*KW] (Torch7 order) += scale * dwp[O, Kpad] (packed order)
__global__ void conv_wgrad_unpack_kernel(const float* __restrict__ dwp, float* __restrict__ grad, int O, int C, int KH,
                                         int KW, int Kpad, float scale) {
  const int K = C * KH * KW;
  const int total = O * K;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const int o = i / K, kk = i % K;
    const int kw = kk % KW, kh = (kk / KW) % KH, c = kk / (KW * KH);
    grad[i] += scale * dwp[o * Kpad + (kh * KW + kw) * C + c];
  }
}

// ---- fused classifier head --------------------------------------------------------------------
// p2 [B, D] bf16  -> dropout(p) -> fc1 (H x D) + ReLU -> fc2 (NC x H) -> log-softmax -> NLL(target)
// D = 64, H = 100, NC = 10 for the reference model; limits: D <= 64, H <= 128, NC <= 32.
constexpr int kHeadD = 64, kHeadH = 128, kHeadC = 32;

struct HeadArgs {
  const __nv_bfloat16* p2;   // [B, D]
  const float *w1, *b1, *w2, *b2;  // fc1 [H, D], [H]; fc2 [NC, H], [NC]  (Torch7 nn.Linear layout)
  const long long* target;   // [B], 0-based
  int B, D, H, NC;
  float p;                   // dropout probability (0 in eval)
  uint32_t seed;
  const uint32_t* seed_dev;  // optional device word mixed into the seed (lets a captured CUDA graph draw a new mask per replay)
  float* hbuf;               // [B, H] fp32 post-ReLU hidden (saved for backward)
  float* logp;               // [B, NC] fp32
  float* loss_partial;       // scratch: [gridDim.x] per-block partial sums  (+1 counter word after it)
  unsigned int* counter;     // scratch: last-block-done counter, must be 0 on entry; left at 0
  float* loss;               // [1] mean loss
  // backward only
  float *gw1, *gb1, *gw2, *gb2;  // fp32 gradient accumulators (+=)
  __nv_bfloat16* dp2;            // [B, D] (optional)
  float grad_scale;              // dLoss (normally 1): gradients are d(mean loss) * grad_scale
  const float* grad_scale_dev;   // optional device scalar folded into grad_scale
  // backward, optional: route dp2 through the preceding ReLU + 3x3 max-pool in the same pass
  const unsigned char* pool_idx; // [B, D] arg-max position (0..8) of the pooling window
  __nv_bfloat16* dpre;           // [B*9, D]: row s*9+pos gets dp2[s] where pos is the arg-max and p2 > 0, else 0
  float* gbias_pre;              // [D] += sum_s dp2[s] * (p2[s] > 0)   (bias gradient of the layer before the pool)
};

// Both head kernels work on chunks of 32 samples with register-tiled inner products (4 samples x 4
// hidden units per thread in the forward; 4 x 2 and 4 x 8 tiles in the backward): two 16-byte
// shared-memory loads feed 16 FMAs, where the first version (one warp per sample) fed one.
constexpr int kHeadS = 32;
__host__ __device__ constexpr int head_hp(int H) { return (H + 3) & ~3; }

// Stage `n` elements into shared memory with the global loads of 8 elements per thread in flight at once
// (written as load-all-then-store-all: in the plain one-element loop every iteration waited for its own L2
// round trip, and those serial round trips were most of the head kernels' time).
template <typename Load>
__device__ __forceinline__ void stage8(float* dst, int n, int t, Load load) {
#pragma unroll 1
  for (int i0 = t; i0 < n; i0 += 256 * 8) {
    float v[8];
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      const int i = i0 + u * 256;
      v[u] = i < n ? load(i) : 0.f;
    }
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      const int i = i0 + u * 256;
      if (i < n) dst[i] = v[u];
    }
  }
}

// global -> shared without a register round trip (Ampere-style cp.async): every copy of a staging phase is in flight at
// once, so the phase costs one L2 latency instead of one per batch of register loads
__device__ __forceinline__ void cp_async4(float* dst, const float* src) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(static_cast<uint32_t>(__cvta_generic_to_shared(dst))), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async16(float* dst, const float* src) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(static_cast<uint32_t>(__cvta_generic_to_shared(dst))), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }

// FIXED: the example network's sizes (D = 64, H = 100, NC = 10) as compile-time constants: the index arithmetic of the
// staging loops (divisions by H) was most of the 8 K instructions each warp of these kernels executed.
template <bool FIXED>
__global__ void __launch_bounds__(256) head_fwd_kernel(HeadArgs a_in) {
  HeadArgs a = a_in;
  if (FIXED) { a.D = 64; a.H = 100; a.NC = 10; }
  extern __shared__ __align__(16) float sm[];
  sm100::pdl_wait();
  sm100::pdl_launch_dependents();
  const int HP = head_hp(a.H);
  const uint32_t seed = a.seed ^ (a.seed_dev != nullptr ? __ldg(a.seed_dev) * 0x9E3779B9u : 0u);
  float* w1t = sm;                          // [64][HP]   w1t[i*HP + j] = w1[j][i]
  float* w2s = w1t + kHeadD * HP;           // [NC][HP]
  float* b1 = w2s + a.NC * HP;              // [HP]
  float* b2 = b1 + HP;                      // [32]
  float* sdt = b2 + kHeadC;                 // [64][32]   dropped input, sample-minor
  float* sh = sdt + kHeadD * kHeadS;        // [32][HP]
  float* slog = sh + kHeadS * HP;           // [32][32]
  __shared__ float block_loss;
  __shared__ bool is_last;
  const int t = threadIdx.x;
  if (FIXED) {   // HP == H, D == 64: no padding to write; the transposed fc1 weights go element by element
    for (int i = t; i < kHeadD * 100; i += 256) {
      const int j = i >> 6, d = i & 63;          // coalesced global reads, conflict-free transposed writes
      cp_async4(w1t + d * 100 + j, a.w1 + i);
    }
    for (int i = t; i < 10 * 100 / 4; i += 256) cp_async16(w2s + 4 * i, a.w2 + 4 * i);
  } else {
  stage8(w1t, kHeadD * HP, t, [&](int i) {
    const int d = i / HP, j = i - d * HP;
    return (d < a.D && j < a.H) ? __ldg(a.w1 + j * a.D + d) : 0.f;
  });
  stage8(w2s, a.NC * HP, t, [&](int i) {
    const int c = i / HP, j = i - c * HP;
    return j < a.H ? __ldg(a.w2 + c * a.H + j) : 0.f;
  });
  }
  for (int i = t; i < HP; i += 256) b1[i] = i < a.H ? __ldg(a.b1 + i) : 0.f;
  if (t < kHeadC) b2[t] = t < a.NC ? __ldg(a.b2 + t) : 0.f;
  const int nJ = HP >> 2;
  const int jq = t % nJ, sq = t / nJ;       // phase-1 tile: hidden units 4*jq.., samples 4*sq..
  float my_loss = 0.f;
  const int chunks = (a.B + kHeadS - 1) / kHeadS;
  for (int ch = blockIdx.x; ch < chunks; ch += gridDim.x) {
    const int base = ch * kHeadS;
    if (FIXED) cp_async_wait_all();   // (the weights; nothing is pending after the first chunk)
    __syncthreads();
    {
      float v[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const int i = t + u * 256, s = i >> 6, d = i & 63;
        v[u] = (base + s < a.B && d < a.D) ? __bfloat162float(a.p2[static_cast<long long>(base + s) * a.D + d]) : 0.f;
      }
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const int i = t + u * 256, s = i >> 6, d = i & 63;
        sdt[d * kHeadS + s] = v[u] * dropout_scale(seed, static_cast<uint32_t>(base + s) * a.D + d, a.p);
      }
    }
    __syncthreads();
    if (sq < kHeadS / 4) {
      float acc[4][4];
#pragma unroll
      for (int x = 0; x < 4; ++x)
#pragma unroll
        for (int y = 0; y < 4; ++y) acc[x][y] = 0.f;
#pragma unroll 8
      for (int i = 0; i < kHeadD; ++i) {
        const float4 dv = *reinterpret_cast<const float4*>(sdt + i * kHeadS + 4 * sq);
        const float4 wv = *reinterpret_cast<const float4*>(w1t + i * HP + 4 * jq);
        const float ds[4] = {dv.x, dv.y, dv.z, dv.w}, ws[4] = {wv.x, wv.y, wv.z, wv.w};
#pragma unroll
        for (int x = 0; x < 4; ++x)
#pragma unroll
          for (int y = 0; y < 4; ++y) acc[x][y] = fmaf(ds[x], ws[y], acc[x][y]);
      }
#pragma unroll
      for (int x = 0; x < 4; ++x) {
        const int s = 4 * sq + x;
        float4 o;
        o.x = fmaxf(acc[x][0] + b1[4 * jq], 0.f); o.y = fmaxf(acc[x][1] + b1[4 * jq + 1], 0.f);
        o.z = fmaxf(acc[x][2] + b1[4 * jq + 2], 0.f); o.w = fmaxf(acc[x][3] + b1[4 * jq + 3], 0.f);
        *reinterpret_cast<float4*>(sh + s * HP + 4 * jq) = o;
      }
    }
    __syncthreads();
    for (int i = t; i < kHeadS * a.H; i += 256) {   // hidden -> global (coalesced)
      const int s = i / a.H, j = i - s * a.H;
      if (base + s < a.B) a.hbuf[static_cast<long long>(base + s) * a.H + j] = sh[s * HP + j];
    }
    for (int it = t; it < kHeadS * a.NC; it += 256) {
      const int s = it / a.NC, c = it - s * a.NC;
      const float* hr = sh + s * HP;
      const float* wr = w2s + c * HP;
      // four independent partial sums (HP is a multiple of 4): a single 100-long dependent FMA chain was ~2.5 us of latency
      float a0 = b2[c], a1 = 0.f, a2 = 0.f, a3 = 0.f;
      for (int j = 0; j < HP; j += 4) {
        const float4 hv = *reinterpret_cast<const float4*>(hr + j);
        const float4 wv = *reinterpret_cast<const float4*>(wr + j);
        a0 = fmaf(hv.x, wv.x, a0); a1 = fmaf(hv.y, wv.y, a1); a2 = fmaf(hv.z, wv.z, a2); a3 = fmaf(hv.w, wv.w, a3);
      }
      slog[s * kHeadC + c] = (a0 + a1) + (a2 + a3);
    }
    __syncthreads();
    if (t < kHeadS && base + t < a.B) {
      const float* lg = slog + t * kHeadC;
      float mx = -INFINITY;
      for (int c = 0; c < a.NC; ++c) mx = fmaxf(mx, lg[c]);
      float se = 0.f;
      for (int c = 0; c < a.NC; ++c) se += __expf(lg[c] - mx);
      const float lse = mx + __logf(se);
      const int tg = static_cast<int>(a.target[base + t]);
      float* lp = a.logp + static_cast<long long>(base + t) * a.NC;
      for (int c = 0; c < a.NC; ++c) lp[c] = lg[c] - lse;
      if (tg >= 0 && tg < a.NC) my_loss -= lg[tg] - lse;
    }
  }
  // deterministic mean: per-block partials in scratch, summed in block order by the last block
  if (t < 32) {
    const float tot = warp_sum(my_loss);   // fixed shuffle tree: bit-reproducible
    if (t == 0) block_loss = tot;
  }
  __syncthreads();
  if (t == 0) {
    a.loss_partial[blockIdx.x] = block_loss;
    __threadfence();
    const unsigned done = atomicAdd(a.counter, 1u);
    is_last = (done == gridDim.x - 1);
  }
  __syncthreads();
  if (is_last) {
    // fixed assignment of partials to threads + fixed reduction tree: deterministic, and 128 L2 round trips
    // are taken in parallel instead of one after the other (that serial loop was most of this kernel's time)
    __threadfence();
    float v = 0.f;
    for (unsigned b = t; b < gridDim.x; b += 256) v += reinterpret_cast<volatile float*>(a.loss_partial)[b];
    v = warp_sum(v);
    __shared__ float red[8];
    if ((t & 31) == 0) red[t >> 5] = v;
    __syncthreads();
    if (t == 0) {
      float tot = 0.f;
#pragma unroll
      for (int w = 0; w < 8; ++w) tot += red[w];
      *a.loss = tot / static_cast<float>(a.B);
      *a.counter = 0u;  // ready for the next launch
    }
  }
}

template <bool FIXED>
__global__ void __launch_bounds__(256) head_bwd_kernel(HeadArgs a_in) {
  HeadArgs a = a_in;
  if (FIXED) { a.D = 64; a.H = 100; a.NC = 10; }
  extern __shared__ __align__(16) float sm[];
  sm100::pdl_wait();
  sm100::pdl_launch_dependents();
  const int HP = head_hp(a.H);
  const uint32_t seed = a.seed ^ (a.seed_dev != nullptr ? __ldg(a.seed_dev) * 0x9E3779B9u : 0u);
  float* w1s = sm;                          // [HP][64]  as stored (row j), zero padded
  float* w2s = w1s + HP * kHeadD;           // [NC][HP]
  float* sd = w2s + a.NC * HP;              // [32][64]  dropped input
  float* sdh = sd + kHeadS * kHeadD;        // [32][HP]  dL/d(pre-ReLU hidden)
  float* sdht = sdh + kHeadS * HP;          // [HP][32]  same, transposed
  float* sh = sdht + HP * kHeadS;           // [32][HP]  hidden
  float* sdl = sh + kHeadS * HP;            // [32][32]  dL/dlogits
  float* sgb = sdl + kHeadS * kHeadC;       // [8][64]   bias-gradient partials of the layer before the pool
  const int t = threadIdx.x;
  if (FIXED) {   // both weight matrices are stored exactly as staged: 16-byte asynchronous copies, waited for with the first chunk
    for (int i = t; i < 100 * kHeadD / 4; i += 256) cp_async16(w1s + 4 * i, a.w1 + 4 * i);
    for (int i = t; i < 10 * 100 / 4; i += 256) cp_async16(w2s + 4 * i, a.w2 + 4 * i);
  } else {
  stage8(w1s, HP * kHeadD, t, [&](int i) {
    const int j = i >> 6, d = i & 63;
    return (j < a.H && d < a.D) ? __ldg(a.w1 + j * a.D + d) : 0.f;
  });
  stage8(w2s, a.NC * HP, t, [&](int i) {
    const int c = i / HP, j = i - c * HP;
    return j < a.H ? __ldg(a.w2 + c * a.H + j) : 0.f;
  });
  }
  const float inv_b = a.grad_scale * (a.grad_scale_dev != nullptr ? *a.grad_scale_dev : 1.0f) / static_cast<float>(a.B);
  const int nJ = HP >> 2;
  // register accumulators
  float acc1[4][8];                          // gw1 tile: hidden 4*jq.. x input 8*iq..   (threads < nJ*8)
#pragma unroll
  for (int x = 0; x < 4; ++x)
#pragma unroll
    for (int y = 0; y < 8; ++y) acc1[x][y] = 0.f;
  constexpr int kW2PerThread = (kHeadC * kHeadH + 255) / 256;  // 16
  float acc2[kW2PerThread];
#pragma unroll
  for (int k = 0; k < kW2PerThread; ++k) acc2[k] = 0.f;
  float accb = 0.f;                          // threads 0..H-1: gb1[t]; threads 128..128+NC-1: gb2[t-128]
  float accp[2] = {0.f, 0.f};                // bias gradient before the pool, channels 2*ip, 2*ip+1
  const int jq = t >> 3, iq = t & 7;         // gw1 tile
  const int ip = t & 31, sg = t >> 5;        // dgrad tile: channels 2*ip.., samples 4*sg..
  const int n2 = a.NC * HP;
  const int chunks = (a.B + kHeadS - 1) / kHeadS;
  for (int ch = blockIdx.x; ch < chunks; ch += gridDim.x) {
    const int base = ch * kHeadS;
    __syncthreads();
    // ---- A: stage the chunk
    const bool async_chunk = FIXED && base + kHeadS <= a.B;   // a full chunk of the fixed-size network: hidden and logp are contiguous
    if (async_chunk) {
      const float* hsrc = a.hbuf + static_cast<long long>(base) * 100;
      for (int i = t; i < kHeadS * 100 / 4; i += 256) cp_async16(sh + 4 * i, hsrc + 4 * i);
      const float* lsrc = a.logp + static_cast<long long>(base) * 10;
      for (int i = t; i < kHeadS * 10; i += 256) {
        const int s = i / 10, c = i - s * 10;
        cp_async4(sdl + s * kHeadC + c, lsrc + i);
      }
    }
    stage8(sd, kHeadS * kHeadD, t, [&](int i) {
      const int s = i >> 6, d = i & 63;
      return (base + s < a.B && d < a.D) ? __bfloat162float(a.p2[static_cast<long long>(base + s) * a.D + d]) : 0.f;
    });
    if (!async_chunk) {
      stage8(sh, kHeadS * HP, t, [&](int i) {
        const int s = i / HP, j = i - s * HP;
        return (base + s < a.B && j < a.H) ? __ldg(a.hbuf + static_cast<long long>(base + s) * a.H + j) : 0.f;
      });
      stage8(sdl, kHeadS * kHeadC, t, [&](int i) {
        const int s = i >> 5, c = i & 31;
        return (base + s < a.B && c < a.NC) ? __ldg(a.logp + static_cast<long long>(base + s) * a.NC + c) : 1.f;
      });
    }
    if (FIXED) cp_async_wait_all();   // the chunk's copies and, the first time, the weights
    __syncthreads();
    // dropout on the staged input; dL/dlogits from the staged log-probabilities (own elements only: no sync needed in between)
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      const int i = t + u * 256, s = i >> 6, d = i & 63;
      sd[i] *= dropout_scale(seed, static_cast<uint32_t>(base + s) * a.D + d, a.p);
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int i = t + u * 256, s = i >> 5, c = i & 31;
      float v = 0.f;
      if (base + s < a.B && c < a.NC) {
        const int tg = static_cast<int>(__ldg(a.target + base + s));
        v = (__expf(sdl[i]) - (c == tg ? 1.f : 0.f)) * inv_b;
      }
      sdl[i] = v;
    }
    __syncthreads();
    // ---- B: dL/d(pre-ReLU hidden), both layouts
    for (int i = t; i < kHeadS * HP; i += 256) {
      const int s = i / HP, j = i - s * HP;
      float g = 0.f;
      for (int c = 0; c < a.NC; ++c) g = fmaf(sdl[s * kHeadC + c], w2s[c * HP + j], g);
      g = sh[i] > 0.f ? g : 0.f;
      sdh[i] = g;
      sdht[j * kHeadS + s] = g;
    }
    __syncthreads();
    // ---- C: gradient w.r.t. the input (4 samples x 2 channels per thread), dropout, optional ReLU+pool routing
    {
      float acc[4][2] = {{0.f, 0.f}, {0.f, 0.f}, {0.f, 0.f}, {0.f, 0.f}};
      for (int j = 0; j < a.H; ++j) {
        const float2 w = *reinterpret_cast<const float2*>(w1s + j * kHeadD + 2 * ip);
        const float4 g = *reinterpret_cast<const float4*>(sdht + j * kHeadS + 4 * sg);
        acc[0][0] = fmaf(g.x, w.x, acc[0][0]); acc[0][1] = fmaf(g.x, w.y, acc[0][1]);
        acc[1][0] = fmaf(g.y, w.x, acc[1][0]); acc[1][1] = fmaf(g.y, w.y, acc[1][1]);
        acc[2][0] = fmaf(g.z, w.x, acc[2][0]); acc[2][1] = fmaf(g.z, w.y, acc[2][1]);
        acc[3][0] = fmaf(g.w, w.x, acc[3][0]); acc[3][1] = fmaf(g.w, w.y, acc[3][1]);
      }
      if (2 * ip < a.D) {
        const bool route = a.dpre != nullptr && (a.D & 1) == 0;
        uint32_t pvs[4] = {0, 0, 0, 0};
        uint32_t ivs[4] = {0, 0, 0, 0};
        if (route) {   // all four samples' pooled values and arg-max bytes in flight together
#pragma unroll
          for (int x = 0; x < 4; ++x) {
            const long long s = base + 4 * sg + x;
            if (s < a.B) {
              pvs[x] = *reinterpret_cast<const uint32_t*>(a.p2 + s * a.D + 2 * ip);
              ivs[x] = *reinterpret_cast<const uint16_t*>(a.pool_idx + s * a.D + 2 * ip);
            }
          }
        }
#pragma unroll
        for (int x = 0; x < 4; ++x) {
          const long long s = base + 4 * sg + x;
          if (s < a.B) {
            const uint32_t e = static_cast<uint32_t>(s) * a.D + 2 * ip;
            const float g0 = acc[x][0] * dropout_scale(seed, e, a.p);
            const float g1 = (2 * ip + 1 < a.D) ? acc[x][1] * dropout_scale(seed, e + 1, a.p) : 0.f;
            const uint32_t gp = pack_bf16x2(g0, g1);
            if ((a.D & 1) == 0) {
              if (a.dp2 != nullptr) *reinterpret_cast<uint32_t*>(a.dp2 + s * a.D + 2 * ip) = gp;
              if (route) {
                const uint32_t pv = pvs[x], iv = ivs[x];
                const uint32_t live = ((pv & 0xffffu) ? 0xffffu : 0u) | ((pv & 0xffff0000u) ? 0xffff0000u : 0u);  // ReLU output > 0
                const uint32_t gl = gp & live;
                const float2 gf = unpack_bf16x2(gl);
                accp[0] += gf.x;
                accp[1] += gf.y;
                const uint32_t p0 = iv & 0xffu, p1 = iv >> 8;
#pragma unroll
                for (uint32_t pos = 0; pos < 9; ++pos) {
                  const uint32_t m = (p0 == pos ? 0xffffu : 0u) | (p1 == pos ? 0xffff0000u : 0u);
                  *reinterpret_cast<uint32_t*>(a.dpre + (s * 9 + pos) * a.D + 2 * ip) = gl & m;
                }
              }
            } else if (a.dp2 != nullptr) {
              a.dp2[s * a.D + 2 * ip] = __float2bfloat16_rn(g0);
              if (2 * ip + 1 < a.D) a.dp2[s * a.D + 2 * ip + 1] = __float2bfloat16_rn(g1);
            }
          }
        }
      }
    }
    // ---- D: fc1 weight gradient tile (4 hidden x 8 inputs per thread)
    if (jq < nJ) {
      for (int s = 0; s < kHeadS; ++s) {
        const float4 g = *reinterpret_cast<const float4*>(sdh + s * HP + 4 * jq);
        const float4 d0 = *reinterpret_cast<const float4*>(sd + s * kHeadD + 8 * iq);
        const float4 d1 = *reinterpret_cast<const float4*>(sd + s * kHeadD + 8 * iq + 4);
        const float gs[4] = {g.x, g.y, g.z, g.w};
        const float ds[8] = {d0.x, d0.y, d0.z, d0.w, d1.x, d1.y, d1.z, d1.w};
#pragma unroll
        for (int x = 0; x < 4; ++x)
#pragma unroll
          for (int y = 0; y < 8; ++y) acc1[x][y] = fmaf(gs[x], ds[y], acc1[x][y]);
      }
    }
    // ---- E: fc2 weight gradient, F: bias gradients
#pragma unroll
    for (int k = 0; k < kW2PerThread; ++k) {
      const int e = t + k * 256;
      if (e < n2) {
        const int c = e / HP, j = e - c * HP;
        float acc = acc2[k];
        for (int s = 0; s < kHeadS; ++s) acc = fmaf(sdl[s * kHeadC + c], sh[s * HP + j], acc);
        acc2[k] = acc;
      }
    }
    if (t < a.H) {
      for (int s = 0; s < kHeadS; ++s) accb += sdh[s * HP + t];
    } else if (t >= 128 && t < 128 + a.NC) {
      for (int s = 0; s < kHeadS; ++s) accb += sdl[s * kHeadC + (t - 128)];
    }
  }
  // ---- flush: the fc1 tile goes through shared memory so that a warp's 32 atomics hit 32 consecutive
  // floats (4 sectors) instead of 32 different sectors
  __syncthreads();
  float* stage = w1s;   // [HP][64]; the weights are no longer needed
  if (jq < nJ) {
#pragma unroll
    for (int x = 0; x < 4; ++x) {
      *reinterpret_cast<float4*>(stage + (4 * jq + x) * kHeadD + 8 * iq) = make_float4(acc1[x][0], acc1[x][1], acc1[x][2], acc1[x][3]);
      *reinterpret_cast<float4*>(stage + (4 * jq + x) * kHeadD + 8 * iq + 4) = make_float4(acc1[x][4], acc1[x][5], acc1[x][6], acc1[x][7]);
    }
  }
  __syncthreads();
  for (int e = t; e < a.H * kHeadD; e += 256) {
    const int j = e >> 6, d = e & 63;
    if (d < a.D) atomicAdd(a.gw1 + j * a.D + d, stage[e]);
  }
#pragma unroll
  for (int k = 0; k < kW2PerThread; ++k) {
    const int e = t + k * 256;
    if (e < n2) {
      const int c = e / HP, j = e - c * HP;
      if (j < a.H) atomicAdd(a.gw2 + c * a.H + j, acc2[k]);
    }
  }
  if (t < a.H) atomicAdd(a.gb1 + t, accb);
  else if (t >= 128 && t < 128 + a.NC) atomicAdd(a.gb2 + (t - 128), accb);
  if (a.gbias_pre != nullptr) {
    __syncthreads();
    sgb[sg * kHeadD + 2 * ip] = accp[0];
    sgb[sg * kHeadD + 2 * ip + 1] = accp[1];
    __syncthreads();
    if (t < a.D) {
      float v = 0.f;
#pragma unroll
      for (int g = 0; g < 8; ++g) v += sgb[g * kHeadD + t];
      atomicAdd(a.gbias_pre + t, v);
    }
  }
}

// ---- optimiser / misc elementwise -------------------------------------------------------------
__global__ void sgd_kernel(float* __restrict__ p, float* __restrict__ g, long long n, float lr, float wd, int zero_grad) {
  sm100::pdl_wait();
  sm100::pdl_launch_dependents();
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < n;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const float w = p[i];
    p[i] = w - lr * (g[i] + wd * w);
    if (zero_grad) g[i] = 0.f;
  }
}
// SGD step that also refreshes the bf16 operand copy of the weights (what the GEMMs read): one pass instead of an
// optimiser pass plus a cast pass per forward.
__global__ void sgd_shadow_kernel(float4* __restrict__ p, const float4* __restrict__ g, uint2* __restrict__ shadow, long long n4, float lr,
                                  float wd) {
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < n4;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    float4 w = p[i];
    const float4 d = g[i];
    w.x -= lr * (d.x + wd * w.x); w.y -= lr * (d.y + wd * w.y); w.z -= lr * (d.z + wd * w.z); w.w -= lr * (d.w + wd * w.w);
    p[i] = w;
    shadow[i] = make_uint2(pack_bf16x2(w.x, w.y), pack_bf16x2(w.z, w.w));
  }
}
__global__ void cast_f32_to_bf16_kernel(const float* __restrict__ x, __nv_bfloat16* __restrict__ y, long long n) {
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < n;
       i += static_cast<long long>(gridDim.x) * blockDim.x)
    y[i] = __float2bfloat16_rn(x[i]);
}

// fp32 [rows, cols] (row pitch ldx) -> bf16 (row pitch ldy >= cols, typically cols rounded up to 8 so that the
// result is a legal TMA operand whatever the layer width)
__global__ void cast_f32_to_bf16_2d_kernel(const float* __restrict__ x, long long ldx, __nv_bfloat16* __restrict__ y, long long ldy,
                                           long long rows, int cols) {
  const long long total = rows * cols;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const long long r = i / cols;
    const int c = static_cast<int>(i - r * cols);
    y[r * ldy + c] = __float2bfloat16_rn(x[r * ldx + c]);
  }
}

int grid_for(long long total, int threads) {
  long long b = (total + threads - 1) / threads;
  const long long cap = static_cast<long long>(sm_count()) * 16;
  if (b > cap) b = cap;
  if (b < 1) b = 1;
  return static_cast<int>(b);
}

size_t head_fwd_smem(int D, int H, int NC) {
  (void)D;
  const size_t HP = head_hp(H);
  return sizeof(float) * (kHeadD * HP + NC * HP + HP + kHeadC + kHeadD * kHeadS + kHeadS * HP + kHeadS * kHeadC);
}
size_t head_bwd_smem(int D, int H, int NC) {
  (void)D;
  const size_t HP = head_hp(H);
  return sizeof(float) * (HP * kHeadD + NC * HP + kHeadS * kHeadD + 3 * kHeadS * HP + kHeadS * kHeadC + 8 * kHeadD);
}

}  // namespace
}  // namespace tadnn

using namespace tadnn;

extern "C" int tadnn_im2col(const void* x, void* col, int B, int C, int H, int W, int KH, int KW, int Kpad, int nchw, tadnn_stream_t stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  const int OH = H - KH + 1, OW = W - KW + 1;
  TADNN_REQUIRE(B > 0 && OH > 0 && OW > 0, "im2col: bad shape");
  TADNN_REQUIRE(Kpad % 8 == 0 && Kpad >= C * KH * KW, "im2col: Kpad=%d must be a multiple of 8 and >= %d", Kpad, C * KH * KW);
  if (nchw && C == 3 && H == 32 && W == 32 && KH == 5 && KW == 5 && Kpad == 80 &&
      (reinterpret_cast<uintptr_t>(x) & 15) == 0) {
    const int blocks = B < sm_count() * 8 ? B : sm_count() * 8;
    im2col_nchw_fixed_kernel<3, 32, 32, 5, 5><<<blocks, 256, 0, stream>>>(static_cast<const __nv_bfloat16*>(x), static_cast<uint4*>(col), B);
    TADNN_LAUNCH_CHECK("im2col_nchw");
  } else if (nchw) {
    const long long total = static_cast<long long>(B) * OH * OW * (Kpad / 8);
    im2col_nchw_kernel<<<grid_for(total, 256), 256, 0, stream>>>(static_cast<const __nv_bfloat16*>(x), static_cast<uint4*>(col), B, C, H, W, KH, KW, OH, OW, Kpad / 8);
    TADNN_LAUNCH_CHECK("im2col_nchw");
  } else {
    TADNN_REQUIRE(C % 8 == 0 && Kpad == C * KH * KW, "im2col (NHWC): C must be a multiple of 8 and Kpad == C*KH*KW");
    const long long total = static_cast<long long>(B) * OH * OW * KH * KW * (C / 8);
    im2col_nhwc_kernel<<<grid_for(total, 256), 256, 0, stream>>>(static_cast<const uint4*>(x), static_cast<uint4*>(col), B, H, W, C / 8, KH, KW, OH, OW, Kpad / 8);
    TADNN_LAUNCH_CHECK("im2col_nhwc");
  }
  return 0;
}

extern "C" int tadnn_im2col_winfirst(const void* x, void* col, int B, int C, int H, int W, int KH, int KW, int win, int only_window,
                                     tadnn_stream_t stream_) {
  const int OH = H - KH + 1, OW = W - KW + 1;
  TADNN_REQUIRE(B > 0 && OH > 0 && OW > 0 && C % 8 == 0, "im2col_winfirst: bad shape (C must be a multiple of 8)");
  TADNN_REQUIRE(win >= 1 && win <= OH && win <= OW, "im2col_winfirst: window %d does not fit %dx%d", win, OH, OW);
  const long long total = static_cast<long long>(B) * (only_window ? win * win : OH * OW) * KH * KW * (C / 8);
  if (C == 64 && H == 9 && W == 9 && KH == 5 && KW == 5 && win == 3) {
    im2col_nhwc_winfirst_fixed_kernel<9, 9, 8, 5, 5, 3><<<grid_for(total, 256), 256, 0, static_cast<cudaStream_t>(stream_)>>>(
        static_cast<const uint4*>(x), static_cast<uint4*>(col), B, only_window);
    TADNN_LAUNCH_CHECK("im2col_nhwc_winfirst");
    return 0;
  }
  im2col_nhwc_winfirst_kernel<<<grid_for(total, 256), 256, 0, static_cast<cudaStream_t>(stream_)>>>(
      static_cast<const uint4*>(x), static_cast<uint4*>(col), B, H, W, C / 8, KH, KW, OH, OW, C * KH * KW / 8, win, only_window);
  TADNN_LAUNCH_CHECK("im2col_nhwc_winfirst");
  return 0;
}

extern "C" int tadnn_conv_toeplitz_pack(const void* wp, void* T, int O, int C, int KH, int KW, int win, tadnn_stream_t stream_) {
  const long long total = static_cast<long long>(win + KH - 1) * (win + KW - 1) * C * win * win * O;
  conv_toeplitz_pack_kernel<<<grid_for(total, 256), 256, 0, static_cast<cudaStream_t>(stream_)>>>(
      static_cast<const __nv_bfloat16*>(wp), static_cast<__nv_bfloat16*>(T), O, C, KH, KW, win);
  TADNN_LAUNCH_CHECK("conv_toeplitz_pack");
  return 0;
}

extern "C" int tadnn_expand_corner(const void* src, void* dst, int B, int H, int W, int FH, int FW, int C, tadnn_stream_t stream_) {
  TADNN_REQUIRE(C % 8 == 0 && FH <= H && FW <= W, "expand_corner: bad shape");
  const long long total = static_cast<long long>(B) * H * W * (C / 8);
  expand_corner_nhwc_kernel<<<grid_for(total, 256), 256, 0, static_cast<cudaStream_t>(stream_)>>>(
      static_cast<const uint4*>(src), static_cast<uint4*>(dst), B, H, W, FH, FW, C / 8);
  TADNN_LAUNCH_CHECK("expand_corner_nhwc");
  return 0;
}

extern "C" int tadnn_col2im(const void* dcol, void* dx, int B, int C, int H, int W, int KH, int KW, tadnn_stream_t stream_) {
  const int OH = H - KH + 1, OW = W - KW + 1;
  TADNN_REQUIRE(B > 0 && OH > 0 && OW > 0 && C % 8 == 0, "col2im: bad shape (C must be a multiple of 8)");
  const long long total = static_cast<long long>(B) * H * W * (C / 8);
  col2im_nhwc_kernel<<<grid_for(total, 256), 256, 0, static_cast<cudaStream_t>(stream_)>>>(static_cast<const uint4*>(dcol), static_cast<uint4*>(dx), B, H, W, C / 8, KH, KW, OH, OW, C * KH * KW / 8);
  TADNN_LAUNCH_CHECK("col2im_nhwc");
  return 0;
}

extern "C" int tadnn_maxpool_fwd(const void* x, void* y, void* idx, int B, int H, int W, int C, int k, tadnn_stream_t stream_) {
  TADNN_REQUIRE(B > 0 && C % 8 == 0 && k >= 1 && k <= 15 && H >= k && W >= k, "maxpool_fwd: bad shape");
  const int PH = H / k, PW = W / k;
  const long long total = static_cast<long long>(B) * PH * PW * (C / 8);
  maxpool_fwd_nhwc_kernel<<<grid_for(total, 256), 256, 0, static_cast<cudaStream_t>(stream_)>>>(static_cast<const uint4*>(x), static_cast<uint4*>(y), static_cast<uint2*>(idx), B, H, W, C / 8, k, PH, PW);
  TADNN_LAUNCH_CHECK("maxpool_fwd_nhwc");
  return 0;
}

static int maxpool_bwd_impl(const void* dy, const void* y, const void* idx, void* dx, int B, int H, int W, int C, int k, int relu, tadnn_stream_t stream_);
extern "C" int tadnn_maxpool_relu_bwd(const void* dy, const void* y, const void* idx, void* dx, int B, int H, int W, int C, int k, tadnn_stream_t stream_) {
  return maxpool_bwd_impl(dy, y, idx, dx, B, H, W, C, k, 1, stream_);
}
extern "C" int tadnn_maxpool_bwd(const void* dy, const void* y, const void* idx, void* dx, int B, int H, int W, int C, int k, tadnn_stream_t stream_) {
  return maxpool_bwd_impl(dy, y, idx, dx, B, H, W, C, k, 0, stream_);
}
static int maxpool_bwd_impl(const void* dy, const void* y, const void* idx, void* dx, int B, int H, int W, int C, int k, int relu, tadnn_stream_t stream_) {
  TADNN_REQUIRE(B > 0 && C % 8 == 0 && k >= 1 && k <= 15 && H >= k && W >= k, "maxpool_relu_bwd: bad shape");
  const int PH = H / k, PW = W / k;
  const long long total = static_cast<long long>(B) * H * W * (C / 8);
  maxpool_relu_bwd_nhwc_kernel<<<grid_for(total, 256), 256, 0, static_cast<cudaStream_t>(stream_)>>>(static_cast<const uint4*>(dy), static_cast<const uint4*>(y), static_cast<const uint2*>(idx), static_cast<uint4*>(dx), B, H, W, C / 8, k, PH, PW, relu);
  TADNN_LAUNCH_CHECK(relu ? "maxpool_relu_bwd_nhwc" : "maxpool_bwd_nhwc");
  return 0;
}

extern "C" int tadnn_colsum_bf16(const void* x, float* out, long long R, int C, int ld, tadnn_stream_t stream_) {
  TADNN_REQUIRE(R > 0 && C > 0, "colsum: empty");
  if (C % 8 == 0 && ld % 8 == 0 && (reinterpret_cast<uintptr_t>(x) & 15) == 0) {
    const int cx = (C + 63) / 64;
    long long ry = (4ll * sm_count() + cx - 1) / cx;
    if (ry > (R + 255) / 256) ry = (R + 255) / 256;
    if (ry < 1) ry = 1;
    colsum_bf16_vec_kernel<<<dim3(cx, static_cast<unsigned>(ry)), dim3(8, 32), 0, static_cast<cudaStream_t>(stream_)>>>(static_cast<const __nv_bfloat16*>(x), out, R, C, ld);
    TADNN_LAUNCH_CHECK("colsum_bf16");
    return 0;
  }
  long long blocks = (R + 63) / 64;
  if (blocks > sm_count() * 4) blocks = sm_count() * 4;
  colsum_bf16_kernel<<<static_cast<int>(blocks), dim3(64, 4), 0, static_cast<cudaStream_t>(stream_)>>>(static_cast<const __nv_bfloat16*>(x), out, R, C, ld);
  TADNN_LAUNCH_CHECK("colsum_bf16");
  return 0;
}

extern "C" int tadnn_colsum_relu_mask(const void* dy, const void* y, float* out, long long R, int C, tadnn_stream_t stream_) {
  TADNN_REQUIRE(R > 0 && C > 0, "colsum_relu_mask: empty");
  long long blocks = (R + 63) / 64;
  if (blocks > sm_count() * 4) blocks = sm_count() * 4;
  colsum_relu_mask_kernel<<<static_cast<int>(blocks), dim3(64, 4), 0, static_cast<cudaStream_t>(stream_)>>>(static_cast<const __nv_bfloat16*>(dy), static_cast<const __nv_bfloat16*>(y), out, R, C);
  TADNN_LAUNCH_CHECK("colsum_relu_mask");
  return 0;
}

extern "C" int tadnn_colsum_relu_mask_corner(const void* dy, const void* y, float* out, int B, int H, int W, int FH, int FW, int C,
                                             tadnn_stream_t stream_) {
  TADNN_REQUIRE(B > 0 && C % 8 == 0 && FH >= 1 && FW >= 1 && FH <= H && FW <= W, "colsum_relu_mask_corner: bad shape");
  const int cx = (C + 63) / 64;
  const long long R = static_cast<long long>(B) * FH * FW;
  long long ry = (4ll * sm_count() + cx - 1) / cx;
  if (ry > (R + 255) / 256) ry = (R + 255) / 256;
  if (ry < 1) ry = 1;
  colsum_relu_mask_corner_kernel<<<dim3(cx, static_cast<unsigned>(ry)), dim3(8, 32), 0, static_cast<cudaStream_t>(stream_)>>>(
      static_cast<const __nv_bfloat16*>(dy), static_cast<const __nv_bfloat16*>(y), out, B, H, W, FH, FW, C);
  TADNN_LAUNCH_CHECK("colsum_relu_mask_corner");
  return 0;
}

extern "C" int tadnn_conv_weight_pack(const float* w, void* wp, int O, int C, int KH, int KW, int Kpad, tadnn_stream_t stream_) {
  TADNN_REQUIRE(Kpad >= C * KH * KW, "conv_weight_pack: Kpad too small");
  conv_weight_pack_kernel<<<grid_for(static_cast<long long>(O) * Kpad, 256), 256, 0, static_cast<cudaStream_t>(stream_)>>>(w, static_cast<__nv_bfloat16*>(wp), O, C, KH, KW, Kpad);
  TADNN_LAUNCH_CHECK("conv_weight_pack");
  return 0;
}

extern "C" int tadnn_conv_wgrad_unpack(const float* dwp, float* grad, int O, int C, int KH, int KW, int Kpad, float scale, tadnn_stream_t stream_) {
  conv_wgrad_unpack_kernel<<<grid_for(static_cast<long long>(O) * C * KH * KW, 256), 256, 0, static_cast<cudaStream_t>(stream_)>>>(dwp, grad, O, C, KH, KW, Kpad, scale);
  TADNN_LAUNCH_CHECK("conv_wgrad_unpack");
  return 0;
}

static int head_check(const tadnn_head_args* h, const char* what) {
  TADNN_REQUIRE(h != nullptr, "%s: null args", what);
  TADNN_REQUIRE(h->B > 0 && h->D > 0 && h->D <= kHeadD && h->H > 0 && h->H <= kHeadH && h->NC > 0 && h->NC <= kHeadC,
                "%s: unsupported head shape D=%d (<=%d) H=%d (<=%d) NC=%d (<=%d)", what, h->D, kHeadD, h->H, kHeadH, h->NC, kHeadC);
  return 0;
}

static HeadArgs to_dev(const tadnn_head_args* h) {
  HeadArgs a;
  memset(&a, 0, sizeof(a));
  a.p2 = static_cast<const __nv_bfloat16*>(h->p2);
  a.w1 = h->w1; a.b1 = h->b1; a.w2 = h->w2; a.b2 = h->b2;
  a.target = h->target;
  a.B = h->B; a.D = h->D; a.H = h->H; a.NC = h->NC;
  a.p = h->dropout_p; a.seed = h->seed;
  a.seed_dev = h->seed_dev;
  a.hbuf = h->hidden; a.logp = h->logp; a.loss = h->loss;
  a.gw1 = h->gw1; a.gb1 = h->gb1; a.gw2 = h->gw2; a.gb2 = h->gb2;
  a.dp2 = static_cast<__nv_bfloat16*>(h->dp2);
  a.grad_scale = h->grad_scale;
  a.grad_scale_dev = h->grad_scale_dev;
  a.pool_idx = static_cast<const unsigned char*>(h->pool_idx);
  a.dpre = static_cast<__nv_bfloat16*>(h->dpre);
  a.gbias_pre = h->gbias_pre;
  return a;
}

extern "C" int tadnn_head_workspace_bytes(int B) {
  (void)B;
  return static_cast<int>(sizeof(float) * 1024 + 256);
}

extern "C" int tadnn_head_forward(const tadnn_head_args* h, void* workspace, tadnn_stream_t stream_) {
  if (int rc = head_check(h, "head_forward")) return rc;
  TADNN_REQUIRE(workspace != nullptr, "head_forward: workspace is required (tadnn_head_workspace_bytes)");
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  HeadArgs a = to_dev(h);
  a.loss_partial = static_cast<float*>(workspace);
  a.counter = reinterpret_cast<unsigned int*>(static_cast<uint8_t*>(workspace) + sizeof(float) * 1024);
  int blocks = (h->B + kHeadS - 1) / kHeadS;
  const int cap = sm_count() * 2 < 1024 ? sm_count() * 2 : 1024;
  if (blocks > cap) blocks = cap;
  const size_t smem = head_fwd_smem(h->D, h->H, h->NC);
  static std::atomic<unsigned long long> attr{0};
  static std::atomic<unsigned long long> attr_fixed{0};
  if (int rc = ensure_dyn_smem(head_fwd_kernel<false>, 96 * 1024, attr)) return rc;
  if (int rc = ensure_dyn_smem(head_fwd_kernel<true>, 96 * 1024, attr_fixed)) return rc;
  const bool fixed = h->D == 64 && h->H == 100 && h->NC == 10 && ((reinterpret_cast<uintptr_t>(h->w1) | reinterpret_cast<uintptr_t>(h->w2)) & 15) == 0;
  TADNN_REQUIRE(smem <= 96 * 1024, "head_forward: shared memory %zu too large", smem);
  if (fixed) TADNN_CHECK_CUDA(launch_pdl(head_fwd_kernel<true>, dim3(blocks), dim3(256), smem, stream, a));
  else TADNN_CHECK_CUDA(launch_pdl(head_fwd_kernel<false>, dim3(blocks), dim3(256), smem, stream, a));
  TADNN_LAUNCH_CHECK("head_fwd");
  return 0;
}

// ---- allocation-free variant of the classifier-head forward -------------------------------------
// Some callers of this library are not in a position to hand us a workspace, because they cannot
// allocate device memory where the call happens: a serving or control loop that must never hit
// cudaMalloc (it synchronises the whole device and its latency is unbounded), a host program that
// has already given every byte of device memory to its own pool, or a plain-C client of this ABI
// that has no allocator wired up at all.  For them the scratch that tadnn_head_forward takes as a
// pointer lives here instead, in a fixed array inside the library image.  Every GPU that loads
// libtadnn.so carries these 2 KB until the process exits.
static __device__ __align__(256) unsigned char tadnn_head_noalloc_scratch[2048];

namespace {
constexpr int kNoAllocMaxBlocks = 256;                      // partial-sum slots that fit in the array
constexpr int kNoAllocMaxBatch = 65536;                    // samples per call this entry point accepts
constexpr size_t kNoAllocCounterOffset = 1024;              // 256 floats of partials, then the counter
std::mutex g_noalloc_mu;
void* g_noalloc_ptr[64] = {nullptr};                        // device address of the array, per device
}  // namespace

extern "C" int tadnn_head_forward_noalloc(const tadnn_head_args* h, tadnn_stream_t stream_) {
  if (int rc = head_check(h, "head_forward_noalloc")) return rc;
  TADNN_REQUIRE(h->B <= kNoAllocMaxBatch,
                "head_forward_noalloc: batch %d exceeds the %d samples the built-in scratch is sized for; "
                "use tadnn_head_forward with a caller-provided workspace", h->B, kNoAllocMaxBatch);
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  int dev = 0;
  TADNN_CHECK_CUDA(cudaGetDevice(&dev));
  TADNN_REQUIRE(dev >= 0 && dev < 64, "head_forward_noalloc: device ordinal %d out of range", dev);
  const size_t smem = head_fwd_smem(h->D, h->H, h->NC);
  TADNN_REQUIRE(smem <= 96 * 1024, "head_forward_noalloc: shared memory %zu too large", smem);
  void* scratch = nullptr;
  {
    // One-time set-up per device, with that device current.  In host code the array's name is only
    // a stand-in; the address a kernel can use has to be asked for.
    std::lock_guard<std::mutex> lk(g_noalloc_mu);
    scratch = g_noalloc_ptr[dev];
    if (scratch == nullptr) {
      TADNN_CHECK_CUDA(cudaGetSymbolAddress(&scratch, tadnn_head_noalloc_scratch));
      TADNN_CHECK_CUDA(cudaFuncSetAttribute(head_fwd_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024));
      g_noalloc_ptr[dev] = scratch;
    }
  }
  HeadArgs a = to_dev(h);
  a.loss_partial = static_cast<float*>(scratch);
  a.counter = reinterpret_cast<unsigned int*>(static_cast<unsigned char*>(scratch) + kNoAllocCounterOffset);
  // The array is shared by every call and still holds whatever the previous call left behind.
  // Only the completion counter has to start at zero; the partial sums are overwritten before use.
  TADNN_CHECK_CUDA(cudaMemsetAsync(a.counter, 0, sizeof(unsigned int), stream));
  int blocks = (h->B + kHeadS - 1) / kHeadS;
  if (blocks > kNoAllocMaxBlocks) blocks = kNoAllocMaxBlocks;
  TADNN_CHECK_CUDA(launch_pdl(head_fwd_kernel<false>, dim3(blocks), dim3(256), smem, stream, a));
  TADNN_LAUNCH_CHECK("head_fwd_noalloc");
  return 0;
}

extern "C" int tadnn_head_backward(const tadnn_head_args* h, tadnn_stream_t stream_) {
  if (int rc = head_check(h, "head_backward")) return rc;
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  HeadArgs a = to_dev(h);
  TADNN_REQUIRE(h->dpre == nullptr || (h->pool_idx != nullptr && h->D % 2 == 0), "head_backward: dpre needs pool_idx and an even D");
  int blocks = (h->B + kHeadS - 1) / kHeadS;
  if (blocks > sm_count()) blocks = sm_count();
  const size_t smem = head_bwd_smem(h->D, h->H, h->NC);
  static std::atomic<unsigned long long> attr{0};
  static std::atomic<unsigned long long> attr_fixed{0};
  if (int rc = ensure_dyn_smem(head_bwd_kernel<false>, 160 * 1024, attr)) return rc;
  if (int rc = ensure_dyn_smem(head_bwd_kernel<true>, 160 * 1024, attr_fixed)) return rc;
  TADNN_REQUIRE(smem <= 160 * 1024, "head_backward: shared memory %zu too large", smem);
  if (h->D == 64 && h->H == 100 && h->NC == 10 &&
      ((reinterpret_cast<uintptr_t>(h->w1) | reinterpret_cast<uintptr_t>(h->w2) | reinterpret_cast<uintptr_t>(h->hidden)) & 15) == 0)
    TADNN_CHECK_CUDA(launch_pdl(head_bwd_kernel<true>, dim3(blocks), dim3(256), smem, stream, a));
  else TADNN_CHECK_CUDA(launch_pdl(head_bwd_kernel<false>, dim3(blocks), dim3(256), smem, stream, a));
  TADNN_LAUNCH_CHECK("head_bwd");
  return 0;
}

extern "C" int tadnn_sgd_step(float* p, float* g, long long n, float lr, float weight_decay, int zero_grad, tadnn_stream_t stream_) {
  if (n <= 0) return 0;
  TADNN_CHECK_CUDA(launch_pdl(sgd_kernel, dim3(grid_for(n, 256)), dim3(256), 0, static_cast<cudaStream_t>(stream_), p, g, n, lr, weight_decay, zero_grad));
  TADNN_LAUNCH_CHECK("sgd_step");
  return 0;
}

extern "C" int tadnn_cast_f32_bf16(const float* x, void* y, long long n, tadnn_stream_t stream_) {
  if (n <= 0) return 0;
  cast_f32_to_bf16_kernel<<<grid_for(n, 256), 256, 0, static_cast<cudaStream_t>(stream_)>>>(x, static_cast<__nv_bfloat16*>(y), n);
  TADNN_LAUNCH_CHECK("cast_f32_bf16");
  return 0;
}

extern "C" int tadnn_cast_f32_bf16_2d(const float* x, long long ldx, void* y, long long ldy, long long rows, int cols, tadnn_stream_t stream_) {
  if (rows <= 0 || cols <= 0) return 0;
  TADNN_REQUIRE(ldx >= cols && ldy >= cols, "cast_f32_bf16_2d: pitches must be >= cols");
  cast_f32_to_bf16_2d_kernel<<<grid_for(rows * cols, 256), 256, 0, static_cast<cudaStream_t>(stream_)>>>(x, ldx, static_cast<__nv_bfloat16*>(y), ldy, rows, cols);
  TADNN_LAUNCH_CHECK("cast_f32_bf16");
  return 0;
}

extern "C" int tadnn_sgd_step_shadow(float* p, const float* g, void* shadow_bf16, long long n, float lr, float weight_decay, tadnn_stream_t stream_) {
  if (n <= 0) return 0;
  TADNN_REQUIRE(n % 4 == 0 && ((reinterpret_cast<uintptr_t>(p) | reinterpret_cast<uintptr_t>(g)) & 15) == 0 && (reinterpret_cast<uintptr_t>(shadow_bf16) & 7) == 0,
                "sgd_step_shadow: n must be a multiple of 4 and the buffers 16-byte (shadow: 8-byte) aligned");
  sgd_shadow_kernel<<<grid_for(n / 4, 256), 256, 0, static_cast<cudaStream_t>(stream_)>>>(reinterpret_cast<float4*>(p), reinterpret_cast<const float4*>(g),
                                                                                         static_cast<uint2*>(shadow_bf16), n / 4, lr, weight_decay);
  TADNN_LAUNCH_CHECK("sgd_step_shadow");
  return 0;
}
