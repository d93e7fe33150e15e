// Fused first convolution stage of the example CNN, forward and backward, as implicit GEMMs on
// tcgen05 with operands generated on the fly in shared memory (no im2col buffer, no un-pooled
// activation or gradient ever touches HBM).
//
//   forward : p1, idx1 = maxpool3x3/3( relu( conv5x5(x; W1) + b1 ) )
//   backward: dW1 += sum_pixels patch(x)^T * d(conv out),  with d(conv out) regenerated from
//             (dp1, p1, idx1): the pooled gradient routed to each window's arg-max where p1 > 0
//
// Replaces, for the 3x32x32 -> 64x28x28 -> 64x9x9 stage of
// cifar_example/sgd-torchad_nn-cifar.lua:91-93, the cunn calls SpatialConvolutionMM (im2col +
// SGEMM), ReLU and SpatialMaxPooling and their backward counterparts.
//
// Pixel order: pooling-window major.  GEMM row r of tile t is pixel (dy,dx) = (r%9 / 3, r%9 % 3) of
// window 14*t + r/9 (windows numbered b*81 + py*9 + px), so the 9 conv outputs a pooled value
// needs sit in 9 consecutive accumulator rows and pooling happens in the epilogue.  Only the
// 27x27 pixels that some window reads are computed (a fused conv+pool has no use for the rest).
// Patch order: k = (kh*3 + c)*5 + kw, 75 values padded to 80 (5 MMAs of K=16).
#include "common.cuh"
#include "sm100.cuh"

namespace tadnn {
namespace {

using namespace sm100;

constexpr int kC = 3, kH = 32, kW = 32, kKH = 5, kKW = 5, kO = 64;
constexpr int kK = kC * kKH * kKW;   // 75
constexpr int kKpad = 80;
constexpr int kPool = 3, kPH = 9, kPW = 9;
constexpr int kWinPerImg = kPH * kPW;        // 81
constexpr int kWinPerTile = 14;              // 14 windows x 9 pixels = 126 rows of a 128-row tile
constexpr int kRows = 128;
constexpr int kBlockBytes = kRows * 128;     // one [128 rows x 64 bf16] SWIZZLE_128B block: 16 KB
constexpr int kThreads = 288;                // warp 0: MMA/TMEM, warps 1-4: producers, warps 5-8: second group

struct Conv1Args {
  const __nv_bfloat16* x;     // [B, 3, 32, 32]
  const __nv_bfloat16* w;     // packed [64, 80], k = (kh*3+c)*5+kw
  const float* bias;          // [64]
  __nv_bfloat16* p1;          // [B, 9, 9, 64]
  unsigned char* idx1;        // [B, 9, 9, 64]
  const __nv_bfloat16* dp1;   // backward: [B, 9, 9, 64]
  float* gw;                  // backward: [64, 75] Torch7 layout (c,kh,kw), accumulated with red.add
  int B;
};

// byte offset of 16-byte chunk j of row r inside a SWIZZLE_128B [rows x 128 B] block
__device__ __forceinline__ uint32_t sw128(int r, int j) { return static_cast<uint32_t>(r * 128 + ((j ^ (r & 7)) << 4)); }

// Build this thread's im2col row (80 bf16 = 10 chunks) for tile row `r` of tile `tile` and store it
// into the two swizzled blocks: block0 holds k 0..63, block1 holds k 64..79 in its chunks 0..1.
__device__ __forceinline__ void build_patch_row(const __nv_bfloat16* __restrict__ x, long long total_win, int tile, int r,
                                                uint8_t* blk0, uint8_t* blk1) {
  const long long win = static_cast<long long>(tile) * kWinPerTile + r / 9;
  const bool valid = (r < kWinPerTile * 9) && (win < total_win);
  __align__(16) __nv_bfloat16 v[kKpad];
  if (valid) {
    const int pos = r % 9;
    const int b = static_cast<int>(win / kWinPerImg);
    const int wi = static_cast<int>(win % kWinPerImg);
    const int oy = (wi / kPW) * kPool + pos / 3;
    const int ox = (wi % kPW) * kPool + pos % 3;
    const __nv_bfloat16* img = x + static_cast<long long>(b) * (kC * kH * kW) + oy * kW + ox;
#pragma unroll
    for (int kh = 0; kh < kKH; ++kh)
#pragma unroll
      for (int c = 0; c < kC; ++c)
#pragma unroll
        for (int kw = 0; kw < kKW; ++kw)
          v[(kh * kC + c) * kKW + kw] = __ldg(img + (c * kH + kh) * kW + kw);
#pragma unroll
    for (int k = kK; k < kKpad; ++k) v[k] = __float2bfloat16_rn(0.f);
  } else {
#pragma unroll
    for (int k = 0; k < kKpad; ++k) v[k] = __float2bfloat16_rn(0.f);
  }
  const uint4* q = reinterpret_cast<const uint4*>(v);
#pragma unroll
  for (int j = 0; j < 8; ++j) *reinterpret_cast<uint4*>(blk0 + sw128(r, j)) = q[j];
#pragma unroll
  for (int j = 0; j < 2; ++j) *reinterpret_cast<uint4*>(blk1 + sw128(r, j)) = q[8 + j];
}

// ======================================= forward =======================================
// smem: W block0 | W block1 (64 rows each, 8 KB) | A[2] x (block0 16 KB + block1 16 KB) | E 16 KB
constexpr int kFwdSmem = 2 * 8192 + 2 * 2 * kBlockBytes + kRows * 128 + 1024;

__global__ void __launch_bounds__(kThreads, 1) conv1_fwd_kernel(const Conv1Args a) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
  uint8_t* w0 = smem;
  uint8_t* w1 = smem + 8192;
  uint8_t* abuf = smem + 16384;                       // stage s: abuf + s*32768 (+0: block0, +16384: block1)
  uint8_t* ebuf = abuf + 2 * 2 * kBlockBytes;         // [128 rows][64 ch] bf16, plain row-major
  __shared__ __align__(8) uint64_t a_full[2], a_empty[2], t_full[2], t_empty[2];
  __shared__ uint32_t tmem_base_smem;
  __shared__ float s_bias[kO];

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long long total_win = static_cast<long long>(a.B) * kWinPerImg;
  const int num_tiles = static_cast<int>((total_win + kWinPerTile - 1) / kWinPerTile);

  if (threadIdx.x == 0) {
    for (int s = 0; s < 2; ++s) {
      mbar_init(&a_full[s], 128);
      mbar_init(&a_empty[s], 1);
      mbar_init(&t_full[s], 1);
      mbar_init(&t_empty[s], 128);
    }
    fence_mbar_init();
  }
  if (warp == 0) tmem_alloc<128>(&tmem_base_smem);
  // weights -> shared memory (K-major, swizzled), bias
  for (int i = threadIdx.x; i < kO * 10; i += blockDim.x) {
    const int row = i / 10, j = i % 10;
    const uint4 v = *reinterpret_cast<const uint4*>(a.w + row * kKpad + j * 8);
    if (j < 8) *reinterpret_cast<uint4*>(w0 + sw128(row, j)) = v;
    else *reinterpret_cast<uint4*>(w1 + sw128(row, j - 8)) = v;
  }
  if (threadIdx.x < kO) s_bias[threadIdx.x] = a.bias[threadIdx.x];
  fence_proxy_async_smem();
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  const uint32_t tmem_base = tmem_base_smem;

  if (warp == 0) {
    // ---------------- MMA issuer ----------------
    if (elect_one()) {
      constexpr uint32_t idesc = make_idesc_bf16(128, kO, 0, 0);
      const uint32_t w0a = smem_u32(w0), w1a = smem_u32(w1);
      int i = 0;
      for (int t = blockIdx.x; t < num_tiles; t += gridDim.x, ++i) {
        const int s = i & 1;
        const uint32_t ph = (i >> 1) & 1;
        mbar_wait(&t_empty[s], ph ^ 1);
        mbar_wait(&a_full[s], ph);
        tc_fence_after_sync();
        const uint32_t a0 = smem_u32(abuf + s * 2 * kBlockBytes), a1 = a0 + kBlockBytes;
        const uint32_t d = tmem_base + static_cast<uint32_t>(s * kO);
#pragma unroll
        for (int k = 0; k < 4; ++k)
          umma_f16(d, make_smem_desc_sw128(a0 + k * 32, 16, 1024), make_smem_desc_sw128(w0a + k * 32, 16, 1024), idesc, k > 0);
        umma_f16(d, make_smem_desc_sw128(a1, 16, 1024), make_smem_desc_sw128(w1a, 16, 1024), idesc, 1u);
        umma_commit(&a_empty[s]);
        umma_commit(&t_full[s]);
      }
    }
    __syncwarp();
  } else if (warp <= 4) {
    // ---------------- producers: one im2col row per thread ----------------
    const int r = threadIdx.x - 32;
    int i = 0;
    for (int t = blockIdx.x; t < num_tiles; t += gridDim.x, ++i) {
      const int s = i & 1;
      const uint32_t ph = (i >> 1) & 1;
      mbar_wait(&a_empty[s], ph ^ 1);
      uint8_t* b0 = abuf + s * 2 * kBlockBytes;
      build_patch_row(a.x, total_win, t, r, b0, b0 + kBlockBytes);
      fence_proxy_async_smem();
      mbar_arrive(&a_full[s]);
    }
  } else {
    // ---------------- epilogue: bias + ReLU + 3x3 max-pool with arg-max ----------------
    const int q = warp & 3;                 // TMEM lane quarter of this warp
    const int row = q * 32 + lane;          // accumulator row = tile pixel
    const int et = threadIdx.x - 160;       // 0..127 inside the epilogue group
    int i = 0;
    for (int t = blockIdx.x; t < num_tiles; t += gridDim.x, ++i) {
      const int s = i & 1;
      const uint32_t ph = (i >> 1) & 1;
      mbar_wait(&t_full[s], ph);
      tc_fence_after_sync();
      uint32_t r0[32], r1[32];
      const uint32_t taddr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + static_cast<uint32_t>(s * kO);
      tmem_ld_32x32b_x32(taddr, r0);
      tmem_ld_32x32b_x32(taddr + 32, r1);
      tmem_ld_wait();
      tc_fence_before_sync();
      mbar_arrive(&t_empty[s]);
      uint4* erow = reinterpret_cast<uint4*>(ebuf + row * 128);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        float v[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          const int c = j * 8 + e;
          v[e] = fmaxf(__uint_as_float(c < 32 ? r0[c] : r1[c - 32]) + s_bias[c], 0.f);
        }
        // rotate the chunk position by the row so that the 32 rows of a warp spread over all banks
        erow[(j + row) & 7] = make_uint4(pack_bf16x2(v[0], v[1]), pack_bf16x2(v[2], v[3]), pack_bf16x2(v[4], v[5]), pack_bf16x2(v[6], v[7]));
      }
      asm volatile("bar.sync 1, 128;" ::: "memory");
      // 14 windows x 32 channel pairs
      for (int o = et; o < kWinPerTile * 32; o += 128) {
        const int wl = o >> 5, cp = o & 31;
        const long long win = static_cast<long long>(t) * kWinPerTile + wl;
        if (win < total_win) {
          float b0 = -INFINITY, b1 = -INFINITY;
          int i0 = 0, i1 = 0;
#pragma unroll
          for (int pos = 0; pos < 9; ++pos) {
            const int rr = wl * 9 + pos;
            const int chunk = ((cp >> 2) + rr) & 7;
            const uint32_t u = *reinterpret_cast<const uint32_t*>(ebuf + rr * 128 + chunk * 16 + (cp & 3) * 4);
            const float2 f = unpack_bf16x2(u);
            if (f.x > b0) { b0 = f.x; i0 = pos; }
            if (f.y > b1) { b1 = f.y; i1 = pos; }
          }
          *reinterpret_cast<uint32_t*>(a.p1 + win * kO + cp * 2) = pack_bf16x2(b0, b1);
          *reinterpret_cast<unsigned short*>(a.idx1 + win * kO + cp * 2) = static_cast<unsigned short>(i0 | (i1 << 8));
        }
      }
      asm volatile("bar.sync 1, 128;" ::: "memory");
    }
  }
  tc_fence_before_sync();
  __syncthreads();
  if (warp == 0) {
    tc_fence_after_sync();
    tmem_dealloc<128>(tmem_base);
  }
}

// ======================================= backward =======================================
// D[patch k (128 lanes, 80 used), channel (64 cols)] += sum over pixels  col[pix, k] * g[pix, ch]
//   A = col tile, MN-major: rows = pixels, two 64-wide chunks of k (block0, block1)
//   B = g tile,   MN-major: rows = pixels, 64 channels
// smem: A[2] x (block0 + block1) | G[2] x 16 KB
constexpr int kBwdSmem = 2 * 2 * kBlockBytes + 2 * kBlockBytes + 1024;

__global__ void __launch_bounds__(kThreads, 1) conv1_bwd_kernel(const Conv1Args a) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
  uint8_t* abuf = smem;
  uint8_t* gbuf = smem + 2 * 2 * kBlockBytes;
  __shared__ __align__(8) uint64_t ab_full[2], ab_empty[2], done_bar;
  __shared__ uint32_t tmem_base_smem;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long long total_win = static_cast<long long>(a.B) * kWinPerImg;
  const int num_tiles = static_cast<int>((total_win + kWinPerTile - 1) / kWinPerTile);

  if (threadIdx.x == 0) {
    for (int s = 0; s < 2; ++s) {
      mbar_init(&ab_full[s], 256);
      mbar_init(&ab_empty[s], 1);
    }
    mbar_init(&done_bar, 1);
    fence_mbar_init();
  }
  if (warp == 0) tmem_alloc<64>(&tmem_base_smem);
  // block1 of both stages: chunks 2..7 (k = 80..127) are read by the M=128 MMA and never written
  // by the producers -> zero them once
  for (int i = threadIdx.x; i < 2 * kRows * 6; i += blockDim.x) {
    const int s = i / (kRows * 6), rem = i % (kRows * 6);
    const int r = rem / 6, j = 2 + rem % 6;
    *reinterpret_cast<uint4*>(abuf + s * 2 * kBlockBytes + kBlockBytes + sw128(r, j)) = make_uint4(0, 0, 0, 0);
  }
  fence_proxy_async_smem();
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  const uint32_t tmem_base = tmem_base_smem;
  const bool has_work = static_cast<int>(blockIdx.x) < num_tiles;

  if (warp == 0) {
    if (elect_one()) {
      constexpr uint32_t idesc = make_idesc_bf16(128, kO, 1, 1);
      int i = 0;
      for (int t = blockIdx.x; t < num_tiles; t += gridDim.x, ++i) {
        const int s = i & 1;
        const uint32_t ph = (i >> 1) & 1;
        mbar_wait(&ab_full[s], ph);
        tc_fence_after_sync();
        const uint32_t a0 = smem_u32(abuf + s * 2 * kBlockBytes);
        const uint32_t g0 = smem_u32(gbuf + s * kBlockBytes);
#pragma unroll
        for (int k = 0; k < kRows / 16; ++k)  // 16 pixels per MMA
          umma_f16(tmem_base, make_smem_desc_sw128(a0 + k * 2048, kBlockBytes, 1024),
                   make_smem_desc_sw128(g0 + k * 2048, kBlockBytes, 1024), idesc, (i > 0 || k > 0) ? 1u : 0u);
        umma_commit(&ab_empty[s]);
      }
      umma_commit(&done_bar);
    }
    __syncwarp();
  } else if (warp <= 4) {
    const int r = threadIdx.x - 32;
    int i = 0;
    for (int t = blockIdx.x; t < num_tiles; t += gridDim.x, ++i) {
      const int s = i & 1;
      const uint32_t ph = (i >> 1) & 1;
      mbar_wait(&ab_empty[s], ph ^ 1);
      uint8_t* b0 = abuf + s * 2 * kBlockBytes;
      build_patch_row(a.x, total_win, t, r, b0, b0 + kBlockBytes);
      fence_proxy_async_smem();
      mbar_arrive(&ab_full[s]);
    }
  } else {
    // gradient rows: the pooled gradient routed to the arg-max pixel of each window (ReLU folded in)
    const int r = threadIdx.x - 160;
    int i = 0;
    for (int t = blockIdx.x; t < num_tiles; t += gridDim.x, ++i) {
      const int s = i & 1;
      const uint32_t ph = (i >> 1) & 1;
      mbar_wait(&ab_empty[s], ph ^ 1);
      uint8_t* g = gbuf + s * kBlockBytes;
      const long long win = static_cast<long long>(t) * kWinPerTile + r / 9;
      const bool valid = (r < kWinPerTile * 9) && (win < total_win);
      const unsigned pos = static_cast<unsigned>(r % 9);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        uint4 o = make_uint4(0, 0, 0, 0);
        if (valid) {
          const uint4 dv = *reinterpret_cast<const uint4*>(a.dp1 + win * kO + j * 8);
          const uint4 pv = *reinterpret_cast<const uint4*>(a.p1 + win * kO + j * 8);
          const uint2 iv = *reinterpret_cast<const uint2*>(a.idx1 + win * kO + j * 8);
          const uint32_t dw[4] = {dv.x, dv.y, dv.z, dv.w};
          const uint32_t pw[4] = {pv.x, pv.y, pv.z, pv.w};
          uint32_t ow[4];
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            const uint32_t ib = (e < 2 ? iv.x : iv.y) >> (16 * (e & 1));
            const float2 d2 = unpack_bf16x2(dw[e]);
            const float2 p2 = unpack_bf16x2(pw[e]);
            ow[e] = pack_bf16x2(((ib & 0xffu) == pos && p2.x > 0.f) ? d2.x : 0.f,
                                (((ib >> 8) & 0xffu) == pos && p2.y > 0.f) ? d2.y : 0.f);
          }
          o = make_uint4(ow[0], ow[1], ow[2], ow[3]);
        }
        *reinterpret_cast<uint4*>(g + sw128(r, j)) = o;
      }
      fence_proxy_async_smem();
      mbar_arrive(&ab_full[s]);
    }
    // final epilogue: accumulator rows = patch index k, columns = output channel
    if (has_work) {
      mbar_wait(&done_bar, 0);
      tc_fence_after_sync();
      const int q = warp & 3;
      const int k = q * 32 + lane;
      uint32_t r0[32], r1[32];
      const uint32_t taddr = tmem_base + (static_cast<uint32_t>(q * 32) << 16);
      tmem_ld_32x32b_x32(taddr, r0);
      tmem_ld_32x32b_x32(taddr + 32, r1);
      tmem_ld_wait();
      if (k < kK) {
        const int kw = k % kKW, c = (k / kKW) % kC, kh = k / (kKW * kC);
        float* dst = a.gw + (c * kKH + kh) * kKW + kw;   // Torch7 order inside a row of [64, 75]
#pragma unroll
        for (int ch = 0; ch < kO; ++ch)
          atomicAdd(dst + ch * kK, __uint_as_float(ch < 32 ? r0[ch] : r1[ch - 32]));
      }
    }
  }
  tc_fence_before_sync();
  __syncthreads();
  if (warp == 0) {
    tc_fence_after_sync();
    tmem_dealloc<64>(tmem_base);
  }
}

// Torch7 weight [64, 75] fp32 (k = (c*5+kh)*5+kw)  ->  bf16 [64, 80], k' = (kh*3+c)*5+kw, zero padded
__global__ void conv1_weight_pack_kernel(const float* __restrict__ w, __nv_bfloat16* __restrict__ wp) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= kO * kKpad) return;
  const int o = i / kKpad, k = i % kKpad;
  float v = 0.f;
  if (k < kK) {
    const int kw = k % kKW, c = (k / kKW) % kC, kh = k / (kKW * kC);
    v = w[o * kK + (c * kKH + kh) * kKW + kw];
  }
  wp[i] = __float2bfloat16_rn(v);
}

}  // namespace
}  // namespace tadnn

using namespace tadnn;

extern "C" int tadnn_conv1_weight_pack(const float* w, void* wp, tadnn_stream_t stream) {
  conv1_weight_pack_kernel<<<(kO * kKpad + 255) / 256, 256, 0, static_cast<cudaStream_t>(stream)>>>(w, static_cast<__nv_bfloat16*>(wp));
  TADNN_LAUNCH_CHECK("conv1_weight_pack");
  return 0;
}

extern "C" int tadnn_conv1_pool_forward(const void* x, const void* w_packed, const float* bias, void* p1, void* idx1, int B,
                                        tadnn_stream_t stream) {
  TADNN_REQUIRE(B > 0, "conv1_pool_forward: empty batch");
  TADNN_REQUIRE(((reinterpret_cast<uintptr_t>(w_packed) | reinterpret_cast<uintptr_t>(p1)) & 15) == 0, "conv1_pool_forward: operands must be 16-byte aligned");
  static bool attr = false;
  if (!attr) { TADNN_CHECK_CUDA(cudaFuncSetAttribute(conv1_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kFwdSmem)); attr = true; }
  Conv1Args a;
  memset(&a, 0, sizeof(a));
  a.x = static_cast<const __nv_bfloat16*>(x); a.w = static_cast<const __nv_bfloat16*>(w_packed); a.bias = bias;
  a.p1 = static_cast<__nv_bfloat16*>(p1); a.idx1 = static_cast<unsigned char*>(idx1); a.B = B;
  const long long tiles = (static_cast<long long>(B) * kWinPerImg + kWinPerTile - 1) / kWinPerTile;
  const int grid = tiles < sm_count() ? static_cast<int>(tiles) : sm_count();
  conv1_fwd_kernel<<<grid, kThreads, kFwdSmem, static_cast<cudaStream_t>(stream)>>>(a);
  TADNN_LAUNCH_CHECK("conv1_pool_fwd_tcgen05");
  return 0;
}

extern "C" int tadnn_conv1_pool_backward(const void* x, const void* dp1, const void* p1, const void* idx1, float* grad_weight,
                                         int B, tadnn_stream_t stream) {
  TADNN_REQUIRE(B > 0, "conv1_pool_backward: empty batch");
  TADNN_REQUIRE(((reinterpret_cast<uintptr_t>(dp1) | reinterpret_cast<uintptr_t>(p1) | reinterpret_cast<uintptr_t>(idx1)) & 15) == 0,
                "conv1_pool_backward: operands must be 16-byte aligned");
  static bool attr = false;
  if (!attr) { TADNN_CHECK_CUDA(cudaFuncSetAttribute(conv1_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kBwdSmem)); attr = true; }
  Conv1Args a;
  memset(&a, 0, sizeof(a));
  a.x = static_cast<const __nv_bfloat16*>(x); a.dp1 = static_cast<const __nv_bfloat16*>(dp1);
  a.p1 = const_cast<__nv_bfloat16*>(static_cast<const __nv_bfloat16*>(p1));
  a.idx1 = const_cast<unsigned char*>(static_cast<const unsigned char*>(idx1));
  a.gw = grad_weight; a.B = B;
  const long long tiles = (static_cast<long long>(B) * kWinPerImg + kWinPerTile - 1) / kWinPerTile;
  const int grid = tiles < sm_count() ? static_cast<int>(tiles) : sm_count();
  conv1_bwd_kernel<<<grid, kThreads, kBwdSmem, static_cast<cudaStream_t>(stream)>>>(a);
  TADNN_LAUNCH_CHECK("conv1_pool_bwd_tcgen05");
  return 0;
}
