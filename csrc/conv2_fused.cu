// Second convolution stage of the example CNN as implicit GEMMs on tcgen05, operands fetched by TMA
// straight from the NHWC activation (no im2col matrix in HBM):
//
//   forward : p2, idx2 = maxpool3x3/3( relu( conv5x5(p1; W2) + b2 ) )       [B,9,9,64] -> [B,64]
//   wgrad   : dW2[o, tap, c] += sum_{n, window pixels} d(conv out)[n, pix, o] * p1[n, pix + tap, c]
//
// Replaces, for cifar_example/sgd-torchad_nn-cifar.lua:96-98, the cunn calls SpatialConvolutionMM
// (im2col + SGEMM), ReLU and SpatialMaxPooling, and SpatialConvolutionMM.accGradParameters.
//
// The trick: with p1 stored NHWC [B, 9, 9, 64], the rows of an im2col tile for ONE
// filter tap (kh, kw) are a 4-D box of the activation: (64 channels, 5 x, 5 y, 5 images) at coordinate
// (0, kw, kh, n0) is 125 rows of 128 bytes in (n, y, x) order - exactly a K-major SWIZZLE_128B operand
// tile (rows 125..127 of the 128-row MMA are don't-care).  25 taps = 25 boxes per output tile, each a
// K = 64 slice; the weights are the matching [64 o, 64 c] slice of the packed filter.  The same boxes
// restricted to the pooling window, (64, 3, 3, 7 images) = 63 rows, are the MN-major operand of the
// weight gradient, whose reduction dimension is (image, window pixel).
#include "common.cuh"
#include "sm100.cuh"
#include "conv1s_pack.cuh"

namespace tadnn {
namespace {

using namespace sm100;

constexpr int kC = 64, kO = 64, kKH = 5, kKW = 5, kTaps = 25, kH = 9, kW = 9, kOH = 5, kOW = 5;
constexpr int kWin = 3;                         // pool2 window (floor mode: one window per image)
constexpr int kThreads = 256;

// ------------------------------------------------------------------------------------------------
// forward
// ------------------------------------------------------------------------------------------------
constexpr int kFImgs = 5;                       // images per tile: 5 x 25 pixels = 125 rows
constexpr int kFABytes = 128 * 128;             // A slot (TMA writes 125 rows = 16000 bytes of it)
constexpr int kFABox = kFImgs * kOH * kOW * 128;
constexpr int kFBBytes = kO * 128;              // one tap of the filter: [64 o][64 c] bf16
constexpr int kFStage = kFABytes + kFBBytes;    // 24 KB
constexpr int kFStages = 6;
constexpr int kFPoolRows = kFImgs * kWin * kWin;  // 45 staged rows per tile
constexpr int kFPoolBytes = 6144;               // 45 x 128 B rounded up
constexpr int kFSmem = kFStages * kFStage + kFPoolBytes + 1024;

struct FwdArgs {
  const float* bias;          // [64]
  __nv_bfloat16* p2;          // [B, 64]
  unsigned char* idx2;        // [B, 64] arg-max position inside the 3x3 window (dy*3+dx)
  __nv_bfloat16* a2;          // optional [B, 5, 5, 64] un-pooled activation (nullptr: not stored)
  int B;
  int num_tiles;
};

__global__ void __launch_bounds__(kThreads, 1)
conv2_fwd_kernel(const __grid_constant__ CUtensorMap tma_x, const __grid_constant__ CUtensorMap tma_w, const FwdArgs a) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
  uint8_t* pool = smem + kFStages * kFStage;
  __shared__ __align__(8) uint64_t full_bar[kFStages], empty_bar[kFStages], tmem_full[2], tmem_empty[2];
  __shared__ uint32_t tmem_base_smem;
  __shared__ float s_bias[kO];

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tma_x);
    tma_prefetch_desc(&tma_w);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < kFStages; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(&tmem_full[s], 1);
      mbar_init(&tmem_empty[s], 128);
    }
    fence_mbar_init();
  }
  if (warp == 2) tmem_alloc<128>(&tmem_base_smem);
  // rows 125..127 of every A slot are never written by TMA; give them defined (zero) contents once
  for (int i = threadIdx.x; i < kFStages * 3 * 8; i += blockDim.x) {
    const int s = i / 24, rem = i % 24;
    *reinterpret_cast<uint4*>(smem + s * kFStage + (125 + rem / 8) * 128 + (rem % 8) * 16) = make_uint4(0, 0, 0, 0);
  }
  fence_proxy_async_smem();
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  const uint32_t tmem_base = tmem_base_smem;
  pdl_wait();                 // everything above touched only this CTA's shared / tensor memory
  pdl_launch_dependents();
  if (threadIdx.x < kO) s_bias[threadIdx.x] = a.bias[threadIdx.x];
  __syncthreads();

  if (warp == 0) {
    // ===================== TMA producer: 25 (activation box, filter slice) pairs per tile =====================
    if (elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      for (int t = blockIdx.x; t < a.num_tiles; t += gridDim.x) {
        const int n0 = t * kFImgs;
        for (int tap = 0; tap < kTaps; ++tap) {
          const int kh = tap / kKW, kw = tap - kh * kKW;
          mbar_wait(&empty_bar[stage], phase ^ 1);
          uint8_t* sa = smem + stage * kFStage;
          mbar_arrive_expect_tx(&full_bar[stage], kFABox + kFBBytes);
          tma_load_4d(sa, &tma_x, &full_bar[stage], 0, kw, kh, n0);
          tma_load_2d(sa + kFABytes, &tma_w, &full_bar[stage], tap * kC, 0);
          if (++stage == kFStages) { stage = 0; phase ^= 1; }
        }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    if (elect_one()) {
      constexpr uint32_t idesc = make_idesc_bf16(128, kO, 0, 0);
      int stage = 0, acc = 0;
      uint32_t phase = 0, acc_phase = 0;
      for (int t = blockIdx.x; t < a.num_tiles; t += gridDim.x) {
        mbar_wait(&tmem_empty[acc], acc_phase ^ 1);
        tc_fence_after_sync();
        const uint32_t d = tmem_base + static_cast<uint32_t>(acc * kO);
        for (int tap = 0; tap < kTaps; ++tap) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after_sync();
          const uint32_t a_addr = smem_u32(smem + stage * kFStage);
          const uint32_t b_addr = a_addr + kFABytes;
#pragma unroll
          for (int k = 0; k < 4; ++k)
            umma_f16(d, make_smem_desc_sw128(a_addr + k * 32, 16, 1024), make_smem_desc_sw128(b_addr + k * 32, 16, 1024), idesc,
                     (tap > 0 || k > 0) ? 1u : 0u);
          umma_commit(&empty_bar[stage]);
          if (++stage == kFStages) { stage = 0; phase ^= 1; }
        }
        umma_commit(&tmem_full[acc]);
        if (++acc == 2) { acc = 0; acc_phase ^= 1; }
      }
    }
    __syncwarp();
  } else if (warp >= 4) {
    // ===================== epilogue: bias + ReLU, 3x3 max-pool with first-arg-max =====================
    const int wg = warp & 3;
    const int row = wg * 32 + lane;               // accumulator row = (image, oy, ox) of the tile
    const int img = row / 25, pix = row - img * 25, oy = pix / 5, ox = pix - oy * 5;
    const bool in_tile = row < kFImgs * 25;
    const bool in_win = in_tile && oy < kWin && ox < kWin;
    const int et = threadIdx.x - 128;
    int acc = 0;
    uint32_t acc_phase = 0;
    for (int t = blockIdx.x; t < a.num_tiles; t += gridDim.x) {
      const int n0 = t * kFImgs;
      mbar_wait(&tmem_full[acc], acc_phase);
      tc_fence_after_sync();
      uint32_t r0[32], r1[32];
      const uint32_t taddr = tmem_base + (static_cast<uint32_t>(wg * 32) << 16) + static_cast<uint32_t>(acc * kO);
      tmem_ld_32x32b_x32(taddr, r0);
      tmem_ld_32x32b_x32(taddr + 32, r1);
      tmem_ld_wait();
      tc_fence_before_sync();
      mbar_arrive(&tmem_empty[acc]);
      if (++acc == 2) { acc = 0; acc_phase ^= 1; }
      const bool store_a2 = a.a2 != nullptr && in_tile && n0 + img < a.B;
      if (in_win || store_a2) {
        uint4* prow = reinterpret_cast<uint4*>(pool + (img * 9 + oy * kWin + ox) * 128);
        uint4* grow = store_a2 ? reinterpret_cast<uint4*>(a.a2 + (static_cast<long long>(n0) * 25 + row) * kO) : nullptr;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          float v[8];
#pragma unroll
          for (int e = 0; e < 8; ++e) {
            const int c = j * 8 + e;
            v[e] = fmaxf(__uint_as_float(c < 32 ? r0[c] : r1[c - 32]) + s_bias[c], 0.f);
          }
          const uint4 u = make_uint4(pack_bf16x2(v[0], v[1]), pack_bf16x2(v[2], v[3]), pack_bf16x2(v[4], v[5]), pack_bf16x2(v[6], v[7]));
          if (in_win) prow[(j + row) & 7] = u;   // chunk position rotated by the row: spreads the banks
          if (store_a2) grow[j] = u;
        }
      }
      asm volatile("bar.sync 1, 128;" ::: "memory");
      // 5 images x 32 channel pairs
      for (int o = et; o < kFImgs * 32; o += 128) {
        const int im = o >> 5, cp = o & 31;
        if (n0 + im < a.B) {
          float best0 = -INFINITY, best1 = -INFINITY;
          uint32_t i0 = 0, i1 = 0, packed = 0;
#pragma unroll
          for (int pos = 0; pos < 9; ++pos) {
            const int prow_i = im * 9 + pos;
            const int rr = (im * 25 + (pos / 3) * 5 + pos % 3);   // the accumulator row that wrote this slot
            const uint32_t w = *reinterpret_cast<const uint32_t*>(pool + prow_i * 128 + ((((cp >> 2) + rr) & 7) << 4) + ((cp & 3) << 2));
            const float2 f = unpack_bf16x2(w);
            if (f.x > best0) { best0 = f.x; i0 = pos; packed = (packed & 0xffff0000u) | (w & 0xffffu); }
            if (f.y > best1) { best1 = f.y; i1 = pos; packed = (packed & 0x0000ffffu) | (w & 0xffff0000u); }
          }
          const long long oidx = static_cast<long long>(n0 + im) * kO + cp * 2;
          *reinterpret_cast<uint32_t*>(a.p2 + oidx) = packed;
          *reinterpret_cast<uint16_t*>(a.idx2 + oidx) = static_cast<uint16_t>(i0 | (i1 << 8));
        }
      }
      asm volatile("bar.sync 1, 128;" ::: "memory");
    }
  }
  tc_fence_before_sync();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after_sync();
    tmem_dealloc<128>(tmem_base);
  }
}

// ------------------------------------------------------------------------------------------------
// weight gradient.  D[(tap parity, c), o] += sum_k A[k, (tap, c)] * G[k, o],  k = (image, window pixel)
//   A = two activation boxes (taps 2p, 2p+1) side by side: MN-major, M = 128
//   G = rows of d(conv out) restricted to the pooling window [B*9, 64]: MN-major, N = 64
// One CTA owns one filter row kh (5 taps = 3 accumulators of 64 TMEM columns) and a contiguous range
// of k-blocks (7 images x 9 pixels = 63 rows; row 63 of every operand slot is zero); its partial is
// added to the packed fp32 gradient [64 o, 1600 = tap*64 + c] with coalesced red.global.add.
// ------------------------------------------------------------------------------------------------
constexpr int kWImgs = 7;
constexpr int kWRows = kWImgs * kWin * kWin;    // 63
constexpr int kWChunk = 64 * 128;               // one [64 k-rows][64 elements] MN-major chunk: 8 KB
constexpr int kWBox = kWRows * 128;             // bytes TMA delivers per chunk
constexpr int kWStage = 6 * kWChunk;            // G + 5 activation boxes = 48 KB
constexpr int kWStages = 4;
constexpr int kWSmem = kWStages * kWStage + kWChunk + 1024;  // + one chunk of slack read by the odd tap's M=128 MMA

struct WgradArgs {
  float* dwp;     // [64, 1600] fp32, packed order k = tap*64 + c; accumulated into
  int num_kb;     // ceil(B / 7)
  int splits;     // CTAs per filter row
};

__global__ void __launch_bounds__(kThreads, 1)
conv2_wgrad_kernel(const __grid_constant__ CUtensorMap tma_x, const __grid_constant__ CUtensorMap tma_g, const WgradArgs a) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
  __shared__ __align__(8) uint64_t full_bar[kWStages], empty_bar[kWStages], done_bar;
  __shared__ uint32_t tmem_base_smem;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int kh = blockIdx.x / a.splits, split = blockIdx.x - kh * a.splits;
  const int per = (a.num_kb + a.splits - 1) / a.splits;
  const int kb0 = split * per, kb1 = min(a.num_kb, kb0 + per);

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tma_x);
    tma_prefetch_desc(&tma_g);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < kWStages; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    mbar_init(&done_bar, 1);
    fence_mbar_init();
  }
  if (warp == 2) tmem_alloc<256>(&tmem_base_smem);
  // k-row 63 of every chunk (TMA writes rows 0..62 only) and the slack chunk: zero, once
  for (int i = threadIdx.x; i < (kWStages * 6 + 1) * 8; i += blockDim.x) {
    const int chunk = i >> 3, j = i & 7;
    *reinterpret_cast<uint4*>(smem + chunk * kWChunk + 63 * 128 + j * 16) = make_uint4(0, 0, 0, 0);
  }
  for (int i = threadIdx.x; i < kWChunk / 16; i += blockDim.x)
    *reinterpret_cast<uint4*>(smem + kWStages * kWStage + i * 16) = make_uint4(0, 0, 0, 0);
  fence_proxy_async_smem();
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  const uint32_t tmem_base = tmem_base_smem;
  pdl_wait();                 // everything above touched only this CTA's shared / tensor memory
  pdl_launch_dependents();

  if (warp == 0) {
    if (elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      for (int kb = kb0; kb < kb1; ++kb) {
        mbar_wait(&empty_bar[stage], phase ^ 1);
        uint8_t* s = smem + stage * kWStage;
        mbar_arrive_expect_tx(&full_bar[stage], 6 * kWBox);
        tma_load_2d(s, &tma_g, &full_bar[stage], 0, kb * kWRows);
#pragma unroll
        for (int kw = 0; kw < kKW; ++kw) tma_load_4d(s + (1 + kw) * kWChunk, &tma_x, &full_bar[stage], 0, kw, kh, kb * kWImgs);
        if (++stage == kWStages) { stage = 0; phase ^= 1; }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    if (elect_one()) {
      constexpr uint32_t idesc = make_idesc_bf16(128, kO, 1, 1);
      int stage = 0;
      uint32_t phase = 0;
      for (int kb = kb0; kb < kb1; ++kb) {
        mbar_wait(&full_bar[stage], phase);
        tc_fence_after_sync();
        const uint32_t g_addr = smem_u32(smem + stage * kWStage);
#pragma unroll
        for (int p = 0; p < 3; ++p) {
          const uint32_t a_addr = g_addr + (1 + 2 * p) * kWChunk;
#pragma unroll
          for (int k = 0; k < 4; ++k)   // 16 k-rows per MMA
            umma_f16(tmem_base + static_cast<uint32_t>(p * kO), make_smem_desc_sw128(a_addr + k * 2048, kWChunk, 1024),
                     make_smem_desc_sw128(g_addr + k * 2048, kWChunk, 1024), idesc, (kb > kb0 || k > 0) ? 1u : 0u);
        }
        umma_commit(&empty_bar[stage]);
        if (++stage == kWStages) { stage = 0; phase ^= 1; }
      }
      umma_commit(&done_bar);
    }
    __syncwarp();
  } else if (warp >= 4 && kb1 > kb0) {
    const int wg = warp & 3;
    const int m = wg * 32 + lane;       // accumulator row: (tap parity, channel)
    const int c = m & 63, odd = m >> 6;
    mbar_wait(&done_bar, 0);
    tc_fence_after_sync();
#pragma unroll 1
    for (int p = 0; p < 3; ++p) {
      const int kw = 2 * p + odd;
      uint32_t r0[32], r1[32];
      const uint32_t taddr = tmem_base + (static_cast<uint32_t>(wg * 32) << 16) + static_cast<uint32_t>(p * kO);
      tmem_ld_32x32b_x32(taddr, r0);
      tmem_ld_32x32b_x32(taddr + 32, r1);
      tmem_ld_wait();
      if (kw < kKW) {
        float* dst = a.dwp + (kh * kKW + kw) * kC + c;
#pragma unroll
        for (int o = 0; o < kO; ++o) atomicAdd(dst + o * (kTaps * kC), __uint_as_float(o < 32 ? r0[o] : r1[o - 32]));
      }
    }
  }
  tc_fence_before_sync();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after_sync();
    tmem_dealloc<256>(tmem_base);
  }
}

// ------------------------------------------------------------------------------------------------
// bf16 operand shadows of the fp32 master weights, rebuilt by ONE launch per step:
//   w1p [64, 80]    conv1 filter, k' = (kh*3 + c)*5 + kw, zero padded       (conv1_fused.cu)
//   w2p [64, 1600]  conv2 filter, k' = tap*64 + c                           (forward B operand)
//   Tt  [576, 4032] conv2 filter Toeplitz-expanded for the transposed convolution (dgrad B operand,
//                   MN-major): Tt[(pos, o), (fy, fx, c)] = W2[o, c, fy - py, fx - px] (0 outside the 5x5
//                   kernel), fy < 7, fx < 9 - the gradient w.r.t. p1 is then one dense GEMM
//                   dp1[B, 7 x 9 x 64] = dA2w[B, 9 x 64] * Tt whose row pitch is that of the full
//                   [B, 9, 9, 64] tensor (columns fx >= 7 come out as zeros, rows fy >= 7 are never
//                   written and stay zero).
// Torch7 layouts of the sources: w1 [64, 75] k = (c*5 + kh)*5 + kw;  w2 [64, 1600] likewise.
// ------------------------------------------------------------------------------------------------
constexpr int kTtRows = kWin * kWin * kO;        // 576
constexpr int kTtFH = 7, kTtFW = 9;
constexpr int kTtCols = kTtFH * kTtFW * kC;      // 4032

// b1 != nullptr: w1p is the [128, 128] operand of the sliding-operand conv1 (bias folded in), else the [64, 80] patch-row one
__global__ void __launch_bounds__(256) cnn_pack_shadows_kernel(const float* __restrict__ w1, const float* __restrict__ b1, const float* __restrict__ w2,
                                                               __nv_bfloat16* __restrict__ w1p, __nv_bfloat16* __restrict__ w2p,
                                                               uint4* __restrict__ Tt) {
  pdl_wait();
  pdl_launch_dependents();
  const int n_w1 = b1 ? 128 * 128 : 64 * 80, n_w2 = kO * kTaps * kC / 8, n_t = kTtRows * kTtCols / 8;
  const int total = n_w1 + n_w2 + n_t;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    if (i < n_w1 && b1) {
      w1p[i] = __float2bfloat16_rn(conv1s_weight_value(w1, b1, i));
    } else if (i < n_w1) {
      const int o = i / 80, k = i % 80;
      float v = 0.f;
      if (k < 75) {
        const int kw = k % 5, c = (k / 5) % 3, kh = k / 15;
        v = w1[o * 75 + (c * 5 + kh) * 5 + kw];
      }
      w1p[i] = __float2bfloat16_rn(v);
    } else if (i < n_w1 + n_w2) {
      const int j = i - n_w1;                 // 8 consecutive channels of one (o, tap)
      const int c8 = j & 7, tap = (j >> 3) % kTaps, o = j / (8 * kTaps);
      const float* src = w2 + o * 1600 + (c8 * 8) * 25 + tap;
      uint4 u;
      u.x = pack_bf16x2(src[0], src[25]);
      u.y = pack_bf16x2(src[50], src[75]);
      u.z = pack_bf16x2(src[100], src[125]);
      u.w = pack_bf16x2(src[150], src[175]);
      reinterpret_cast<uint4*>(w2p)[j] = u;
    } else {
      const int j = i - n_w1 - n_w2;          // 8 consecutive channels of one (pos, o, fy, fx)
      const int c8 = j & 7;
      int t = j >> 3;
      const int fx = t % kTtFW; t /= kTtFW;
      const int fy = t % kTtFH; t /= kTtFH;
      const int o = t & 63, pos = t >> 6;
      const int kh = fy - pos / 3, kw = fx - pos % 3;
      uint4 u = make_uint4(0, 0, 0, 0);
      if (kh >= 0 && kh < kKH && kw >= 0 && kw < kKW) {
        const float* src = w2 + o * 1600 + (c8 * 8) * 25 + kh * 5 + kw;
        u.x = pack_bf16x2(src[0], src[25]);
        u.y = pack_bf16x2(src[50], src[75]);
        u.z = pack_bf16x2(src[100], src[125]);
        u.w = pack_bf16x2(src[150], src[175]);
      }
      Tt[j] = u;
    }
  }
}

// gw2[o, (c*5+kh)*5+kw] += dwp[o, tap*64 + c];  dwp = 0 (ready for the next step's red.adds)
__global__ void __launch_bounds__(256) conv2_wgrad_unpack_kernel(float* __restrict__ dwp, float* __restrict__ gw) {
  pdl_wait();
  pdl_launch_dependents();
  const int total = kO * kTaps * kC;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const int c = i & 63, tap = (i >> 6) % kTaps, o = i / (kTaps * kC);
    gw[o * 1600 + c * 25 + tap] += dwp[i];
    dwp[i] = 0.f;
  }
}

int make_x_map(CUtensorMap* m, const void* p1, int B, int bx, int by, int bn) {
  const unsigned long long gdim[4] = {64ull, 9ull, 9ull, static_cast<unsigned long long>(B)};
  const unsigned long long gstr[3] = {128ull, 9ull * 128ull, 81ull * 128ull};
  const unsigned int box[4] = {64u, static_cast<unsigned>(bx), static_cast<unsigned>(by), static_cast<unsigned>(bn)};
  return encode_tmap_bf16(m, p1, 4, gdim, gstr, box);
}

}  // namespace
}  // namespace tadnn

using namespace tadnn;

extern "C" int tadnn_conv2_pool_forward(const void* p1, const void* w2_packed, const float* bias, void* p2, void* idx2, void* a2,
                                        int B, tadnn_stream_t stream_) {
  TADNN_REQUIRE(p1 && w2_packed && bias && p2 && idx2 && B > 0, "conv2_pool_forward: bad arguments");
  CUtensorMap tx, tw;
  if (int rc = make_x_map(&tx, p1, B, kOW, kOH, kFImgs)) return rc;
  {
    const unsigned long long gdim[2] = {static_cast<unsigned long long>(kTaps * kC), static_cast<unsigned long long>(kO)};
    const unsigned long long gstr[1] = {static_cast<unsigned long long>(kTaps * kC * 2)};
    const unsigned int box[2] = {64u, static_cast<unsigned>(kO)};
    if (int rc = encode_tmap_bf16(&tw, w2_packed, 2, gdim, gstr, box)) return rc;
  }
  static std::atomic<unsigned long long> attr{0};
  if (int rc = ensure_dyn_smem(conv2_fwd_kernel, kFSmem, attr)) return rc;
  FwdArgs a;
  a.bias = bias;
  a.p2 = static_cast<__nv_bfloat16*>(p2);
  a.idx2 = static_cast<unsigned char*>(idx2);
  a.a2 = static_cast<__nv_bfloat16*>(a2);
  a.B = B;
  a.num_tiles = (B + kFImgs - 1) / kFImgs;
  const int grid = a.num_tiles < sm_count() ? a.num_tiles : sm_count();
  TADNN_CHECK_CUDA(launch_pdl(conv2_fwd_kernel, dim3(grid), dim3(kThreads), kFSmem, static_cast<cudaStream_t>(stream_), tx, tw, a));
  TADNN_LAUNCH_CHECK("conv2_pool_fwd_tcgen05");
  return 0;
}

extern "C" int tadnn_conv2_wgrad(const void* p1, const void* da2w, float* dw_packed, int B, tadnn_stream_t stream_) {
  TADNN_REQUIRE(p1 && da2w && dw_packed && B > 0, "conv2_wgrad: bad arguments");
  CUtensorMap tx, tg;
  if (int rc = make_x_map(&tx, p1, B, kWin, kWin, kWImgs)) return rc;
  {
    const unsigned long long gdim[2] = {64ull, static_cast<unsigned long long>(B) * 9ull};
    const unsigned long long gstr[1] = {128ull};
    const unsigned int box[2] = {64u, static_cast<unsigned>(kWRows)};
    if (int rc = encode_tmap_bf16(&tg, da2w, 2, gdim, gstr, box)) return rc;
  }
  static std::atomic<unsigned long long> attr{0};
  if (int rc = ensure_dyn_smem(conv2_wgrad_kernel, kWSmem, attr)) return rc;
  WgradArgs a;
  a.dwp = dw_packed;
  a.num_kb = (B + kWImgs - 1) / kWImgs;
  int splits = sm_count() / kKH;
  if (splits > a.num_kb) splits = a.num_kb;
  if (splits < 1) splits = 1;
  a.splits = splits;
  TADNN_CHECK_CUDA(launch_pdl(conv2_wgrad_kernel, dim3(kKH * splits), dim3(kThreads), kWSmem, static_cast<cudaStream_t>(stream_), tx, tg, a));
  TADNN_LAUNCH_CHECK("conv2_wgrad_tcgen05");
  return 0;
}

extern "C" int tadnn_conv2_wgrad_unpack(float* dw_packed, float* grad_weight, tadnn_stream_t stream_) {
  TADNN_REQUIRE(dw_packed && grad_weight, "conv2_wgrad_unpack: null argument");
  TADNN_CHECK_CUDA(launch_pdl(conv2_wgrad_unpack_kernel, dim3((kO * kTaps * kC + 255) / 256), dim3(256), 0, static_cast<cudaStream_t>(stream_), dw_packed, grad_weight));
  TADNN_LAUNCH_CHECK("conv2_wgrad_unpack");
  return 0;
}

extern "C" int tadnn_cnn_pack_shadows(const float* w1, const float* w2, void* w1_packed, void* w2_packed, void* toeplitz_t,
                                      tadnn_stream_t stream_) {
  TADNN_REQUIRE(w1 && w2 && w1_packed && w2_packed && toeplitz_t, "cnn_pack_shadows: null argument");
  TADNN_REQUIRE(((reinterpret_cast<uintptr_t>(w2_packed) | reinterpret_cast<uintptr_t>(toeplitz_t)) & 15) == 0, "cnn_pack_shadows: outputs must be 16-byte aligned");
  const int total = 64 * 80 + kO * kTaps * kC / 8 + kTtRows * kTtCols / 8;
  TADNN_CHECK_CUDA(launch_pdl(cnn_pack_shadows_kernel, dim3((total + 255) / 256), dim3(256), 0, static_cast<cudaStream_t>(stream_), w1,
                              static_cast<const float*>(nullptr), w2, static_cast<__nv_bfloat16*>(w1_packed), static_cast<__nv_bfloat16*>(w2_packed),
                              static_cast<uint4*>(toeplitz_t)));
  TADNN_LAUNCH_CHECK("cnn_pack_shadows");
  return 0;
}

extern "C" int tadnn_cnn_pack_shadows_sliding(const float* w1, const float* b1, const float* w2, void* w1x_packed, void* w2_packed,
                                              void* toeplitz_t, tadnn_stream_t stream_) {
  TADNN_REQUIRE(w1 && b1 && w2 && w1x_packed && w2_packed && toeplitz_t, "cnn_pack_shadows_sliding: null argument");
  TADNN_REQUIRE(((reinterpret_cast<uintptr_t>(w2_packed) | reinterpret_cast<uintptr_t>(toeplitz_t) | reinterpret_cast<uintptr_t>(w1x_packed)) & 15) == 0,
                "cnn_pack_shadows_sliding: outputs must be 16-byte aligned");
  const int total = 128 * 128 + kO * kTaps * kC / 8 + kTtRows * kTtCols / 8;
  TADNN_CHECK_CUDA(launch_pdl(cnn_pack_shadows_kernel, dim3((total + 255) / 256), dim3(256), 0, static_cast<cudaStream_t>(stream_), w1, b1, w2,
                              static_cast<__nv_bfloat16*>(w1x_packed), static_cast<__nv_bfloat16*>(w2_packed), static_cast<uint4*>(toeplitz_t)));
  TADNN_LAUNCH_CHECK("cnn_pack_shadows");
  return 0;
}
