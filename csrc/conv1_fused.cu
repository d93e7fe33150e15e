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
constexpr int kThreads = 416;                // backward: warp 0 MMA/TMEM, warps 1-8 two im2col producer groups (one per smem stage), warps 9-12 gradient rows + final epilogue
constexpr int kFwdThreads = 544;             // forward: same, plus warps 9-12 / 13-16 = two epilogue groups (one per TMEM stage)

struct Conv1Args {
  const __nv_bfloat16* x;     // [B, 3, 32, 32]
  const __nv_bfloat16* w;     // packed [64, 80], k = (kh*3+c)*5+kw
  const float* bias;          // [64]
  __nv_bfloat16* p1;          // [B, 9, 9, 64]
  unsigned char* idx1;        // [B, 9, 9, 64]
  const __nv_bfloat16* dp1;   // backward: [B, 9, 9, 64]
  float* gw;                  // backward: [64, 75] Torch7 layout (c,kh,kw), accumulated with red.add
  float* gb;                  // backward, optional: [64] bias gradient (+=) = sum over windows of dp1 * (p1 > 0)
  int B;
  int dph, dpw;               // backward: dp1 is stored compactly as [B, dph, dpw, 64] (zero outside), dph,dpw <= 9
};

// byte offset of 16-byte chunk j of row r inside a SWIZZLE_128B [rows x 128 B] block
__device__ __forceinline__ uint32_t sw128(int r, int j) { return static_cast<uint32_t>(r * 128 + ((j ^ (r & 7)) << 4)); }

// Stage the (at most two) images a tile touches in shared memory ([2][3][32][32] bf16 = 12 KB) with
// 1-D bulk async copies (cp.async.bulk, completion on an mbarrier): issued by one thread for the
// NEXT tile while the group builds the current one.
constexpr int kImgBytes = kC * kH * kW * 2;  // 6144
__device__ __forceinline__ void stage_images_async(const __nv_bfloat16* __restrict__ x, int B, int b0, uint8_t* img, uint64_t* bar) {
  const int n = (b0 + 1 < B) ? 2 : 1;
  mbar_arrive_expect_tx(bar, static_cast<uint32_t>(n * kImgBytes));
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(img)),
               "l"(x + static_cast<long long>(b0) * (kC * kH * kW)), "r"(n * kImgBytes), "r"(smem_u32(bar))
               : "memory");
}

// Build this thread's im2col row (80 bf16 = 10 chunks) for tile row `r` of tile `tile` from the
// staged images and store it into the two swizzled blocks: block0 holds k 0..63, block1 holds
// k 64..79 in its chunks 0..1.  Each (kh, c) run of 5 contiguous input pixels is fetched as three
// aligned 32-bit words and funnel-shifted by the pixel's column parity; the 2-byte shuffles into
// patch order are compile-time byte permutes.
// ONES: the first padding slot (k = 75) of a valid row holds 1.0, so that a weight row whose slot 75 is
// the bias makes the MMA add it (forward only).
template <bool ONES>
__device__ __forceinline__ void build_patch_row(const uint32_t* __restrict__ img_words, int total_win, int tile, int b0,
                                                int r, uint8_t* blk0, uint8_t* blk1) {
  const int win = tile * kWinPerTile + r / 9;
  const bool valid = (r < kWinPerTile * 9) && (win < total_win);
  uint32_t out[kKpad / 2];
  if (valid) {
    const int pos = r % 9;
    const int b = win / kWinPerImg;
    const int wi = win - b * kWinPerImg;
    const int oy = (wi / kPW) * kPool + pos / 3;
    const int ox = (wi % kPW) * kPool + pos % 3;
    const uint32_t* base = img_words + (b - b0) * (kC * kH * kW / 2) + oy * (kW / 2) + (ox >> 1);
    const uint32_t sh = static_cast<uint32_t>(ox & 1) * 16u;
    uint32_t A[kKH * kC], Bv[kKH * kC], Cv[kKH * kC];
#pragma unroll
    for (int kh = 0; kh < kKH; ++kh)
#pragma unroll
      for (int c = 0; c < kC; ++c) {
        const uint32_t* p = base + (c * kH + kh) * (kW / 2);
        const uint32_t w0 = p[0], w1 = p[1], w2 = p[2];
        const int g = kh * kC + c;
        A[g] = __funnelshift_r(w0, w1, sh);    // elements 0,1 of the run
        Bv[g] = __funnelshift_r(w1, w2, sh);   // elements 2,3
        Cv[g] = (w2 >> sh) & 0xffffu;          // element 4
      }
#pragma unroll
    for (int m = 0; m < kKpad / 2; ++m) {
      uint32_t half[2];
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        const int k = 2 * m + h;
        if (k < kK) {
          const int g = k / kKW, e = k % kKW;
          const uint32_t wsrc = e < 2 ? A[g] : (e < 4 ? Bv[g] : Cv[g]);
          half[h] = (e == 1 || e == 3) ? (wsrc >> 16) : (wsrc & 0xffffu);
        } else {
          half[h] = (ONES && k == kK) ? 0x3f80u : 0u;
        }
      }
      out[m] = half[0] | (half[1] << 16);
    }
  } else {
#pragma unroll
    for (int m = 0; m < kKpad / 2; ++m) out[m] = 0u;
  }
#pragma unroll
  for (int j = 0; j < 8; ++j)
    *reinterpret_cast<uint4*>(blk0 + sw128(r, j)) = make_uint4(out[4 * j], out[4 * j + 1], out[4 * j + 2], out[4 * j + 3]);
#pragma unroll
  for (int j = 0; j < 2; ++j)
    *reinterpret_cast<uint4*>(blk1 + sw128(r, j)) = make_uint4(out[32 + 4 * j], out[33 + 4 * j], out[34 + 4 * j], out[35 + 4 * j]);
}

// ======================================= forward =======================================
// smem: W block0 | W block1 (64 rows each, 8 KB) | A[2] x (block0 16 KB + block1 16 KB) | E 16 KB
constexpr int kImgStageBytes = 2 * 2 * kC * kH * kW * 2;  // per producer group: double-buffered pair of staged images
constexpr int kEPitch = 144;                  // staging row pitch: 128 B + 16 B pad = conflict-free row writes, constant read offsets
constexpr int kEBytes = kRows * kEPitch;      // per epilogue group
constexpr int kFwdSmem = 2 * 8192 + 2 * 2 * kBlockBytes + 2 * kEBytes + 2 * kImgStageBytes + 1024;

__global__ void __launch_bounds__(kFwdThreads, 1) conv1_fwd_kernel(const Conv1Args a) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
  uint8_t* w0 = smem;
  uint8_t* w1 = smem + 8192;
  uint8_t* abuf = smem + 16384;                       // stage s: abuf + s*32768 (+0: block0, +16384: block1)
  uint8_t* ebuf = abuf + 2 * 2 * kBlockBytes;         // 2 x [128 rows][64 ch] bf16, row pitch kEPitch
  uint8_t* imgbuf = ebuf + 2 * kEBytes;                // staged input images (2 * kEBytes = 36 KB keeps them 1 KB aligned)
  __shared__ __align__(8) uint64_t a_full[2], a_empty[2], t_full[2], t_empty[2], img_full[2][2];
  __shared__ uint32_t tmem_base_smem;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long long total_win = static_cast<long long>(a.B) * kWinPerImg;
  const int num_tiles = static_cast<int>((total_win + kWinPerTile - 1) / kWinPerTile);

  if (threadIdx.x == 0) {
    for (int s = 0; s < 2; ++s) {
      mbar_init(&a_full[s], 128);
      mbar_init(&a_empty[s], 1);
      mbar_init(&t_full[s], 1);
      mbar_init(&t_empty[s], 128);
      mbar_init(&img_full[s][0], 1);
      mbar_init(&img_full[s][1], 1);
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
  __syncthreads();
  // the bias rides in the GEMM: weight slot k = 75 (the first padding column; the producers put 1.0 there)
  if (threadIdx.x < kO)
    *reinterpret_cast<__nv_bfloat16*>(w1 + sw128(threadIdx.x, 1) + 2 * (kK - 72)) = __float2bfloat16_rn(a.bias[threadIdx.x]);
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
  } else if (warp <= 8) {
    // ---------------- producers: one im2col row per thread; group g owns smem stage g ----------------
    const int grp = (warp - 1) >> 2;
    const int r = (threadIdx.x - 32) & 127;
    uint8_t* b0p = abuf + grp * 2 * kBlockBytes;
    uint8_t* img = imgbuf + grp * kImgStageBytes;   // two buffers of 2 images
    const long long t_first = blockIdx.x + static_cast<long long>(grp) * gridDim.x, t_step = 2ll * gridDim.x;
    if (r == 0 && t_first < num_tiles)
      stage_images_async(a.x, a.B, static_cast<int>(t_first * kWinPerTile / kWinPerImg), img, &img_full[grp][0]);
    int n = 0;
    for (long long t = t_first; t < num_tiles; t += t_step, ++n) {
      const int buf = n & 1;
      const int b0 = static_cast<int>(t * kWinPerTile / kWinPerImg);
      if (r == 0 && t + t_step < num_tiles)  // prefetch the next tile's images into the other buffer
        stage_images_async(a.x, a.B, static_cast<int>((t + t_step) * kWinPerTile / kWinPerImg), img + (buf ^ 1) * 2 * kImgBytes,
                           &img_full[grp][buf ^ 1]);
      mbar_wait(&img_full[grp][buf], (n >> 1) & 1);
      mbar_wait(&a_empty[grp], (n & 1) ^ 1);
      build_patch_row<true>(reinterpret_cast<const uint32_t*>(img + buf * 2 * kImgBytes), static_cast<int>(total_win), static_cast<int>(t), b0, r, b0p,
                            b0p + kBlockBytes);
      fence_proxy_async_smem();
      mbar_arrive(&a_full[grp]);
      asm volatile("bar.sync %0, 128;" ::"r"(2 + grp) : "memory");  // everyone is done with img[buf] before it is refilled
    }
  } else {
    // ---------------- epilogue: bias + ReLU + 3x3 max-pool with arg-max ----------------
    // two groups of 4 warps; group g drains TMEM stage g (tiles i with i % 2 == g) through its own
    // staging buffer, so one group pools while the other reads the next accumulator.
    const int grp = (warp - 9) >> 2;
    const int q = warp & 3;                 // TMEM lane quarter of this warp
    const int row = q * 32 + lane;          // accumulator row = tile pixel
    const int et = (threadIdx.x - 288) & 127;
    uint8_t* eb = ebuf + grp * kEBytes;
    int n = 0;
    for (long long t = blockIdx.x + static_cast<long long>(grp) * gridDim.x; t < num_tiles; t += 2ll * gridDim.x, ++n) {
      mbar_wait(&t_full[grp], n & 1);
      tc_fence_after_sync();
      uint32_t r0[32], r1[32];
      const uint32_t taddr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + static_cast<uint32_t>(grp * kO);
      tmem_ld_32x32b_x32(taddr, r0);
      tmem_ld_32x32b_x32(taddr + 32, r1);
      tmem_ld_wait();
      tc_fence_before_sync();
      mbar_arrive(&t_empty[grp]);
      // the accumulator already contains the bias; ReLU commutes with rounding and with max, so it is applied to the
      // pooled value: per pixel only the bf16 pack remains
      uint4* erow = reinterpret_cast<uint4*>(eb + row * kEPitch);
#pragma unroll
      for (int j = 0; j < 4; ++j)
        erow[j] = make_uint4(pack_bf16x2(__uint_as_float(r0[8 * j]), __uint_as_float(r0[8 * j + 1])),
                             pack_bf16x2(__uint_as_float(r0[8 * j + 2]), __uint_as_float(r0[8 * j + 3])),
                             pack_bf16x2(__uint_as_float(r0[8 * j + 4]), __uint_as_float(r0[8 * j + 5])),
                             pack_bf16x2(__uint_as_float(r0[8 * j + 6]), __uint_as_float(r0[8 * j + 7])));
#pragma unroll
      for (int j = 0; j < 4; ++j)
        erow[4 + j] = make_uint4(pack_bf16x2(__uint_as_float(r1[8 * j]), __uint_as_float(r1[8 * j + 1])),
                                 pack_bf16x2(__uint_as_float(r1[8 * j + 2]), __uint_as_float(r1[8 * j + 3])),
                                 pack_bf16x2(__uint_as_float(r1[8 * j + 4]), __uint_as_float(r1[8 * j + 5])),
                                 pack_bf16x2(__uint_as_float(r1[8 * j + 6]), __uint_as_float(r1[8 * j + 7])));
      asm volatile("bar.sync %0, 128;" ::"r"(4 + grp) : "memory");
      // 14 windows x 16 channel quads, packed bf16x2 max + first-arg-max
      for (int o = et; o < kWinPerTile * 16; o += 128) {
        const int wl = o >> 4, q4 = o & 15;
        const long long win = t * kWinPerTile + wl;
        if (win < total_win) {
          const uint8_t* base = eb + wl * 9 * kEPitch + q4 * 8;
          uint2 val[9];
#pragma unroll
          for (int pos = 0; pos < 9; ++pos) val[pos] = *reinterpret_cast<const uint2*>(base + pos * kEPitch);
          uint32_t mx[2], ix[2];
#pragma unroll
          for (int h = 0; h < 2; ++h) {
            __nv_bfloat162 m = *reinterpret_cast<const __nv_bfloat162*>(h ? &val[0].y : &val[0].x);
#pragma unroll
            for (int pos = 1; pos < 9; ++pos)
              m = __hmax2(m, *reinterpret_cast<const __nv_bfloat162*>(h ? &val[pos].y : &val[pos].x));
            uint32_t idx = 0;
#pragma unroll
            for (int pos = 8; pos >= 0; --pos) {  // descending: the smallest matching position wins
              const uint32_t eq = __heq2_mask(*reinterpret_cast<const __nv_bfloat162*>(h ? &val[pos].y : &val[pos].x), m);
              idx = (idx & ~eq) | ((static_cast<uint32_t>(pos) * 0x00010001u) & eq);
            }
            m = __hmax2(m, __float2bfloat162_rn(0.f));   // ReLU
            mx[h] = *reinterpret_cast<const uint32_t*>(&m);
            ix[h] = idx;  // two 16-bit positions
          }
          *reinterpret_cast<uint2*>(a.p1 + win * kO + q4 * 4) = make_uint2(mx[0], mx[1]);
          const uint32_t packed = (ix[0] & 0xffu) | ((ix[0] >> 8) & 0xff00u) | ((ix[1] & 0xffu) << 16) | ((ix[1] & 0xff0000u) << 8);
          *reinterpret_cast<uint32_t*>(a.idx1 + win * kO + q4 * 4) = packed;
        }
      }
      asm volatile("bar.sync %0, 128;" ::"r"(4 + grp) : "memory");
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
// Gradient-side inputs of a tile (14 consecutive windows of dp1, p1 and idx1: three contiguous runs) are
// fetched kGDist tiles ahead with bulk async copies into a ring of kGStages slots: with the one-tile register
// prefetch this replaced, a tile (1.2 us) was shorter than the HBM latency of its own inputs and the group sat
// on the scoreboard for half of the kernel (ncu: 48 % of its samples).
constexpr int kGStages = 8, kGDist = 6;
constexpr int kGSlot = 4608;   // 14*128 (dp1) + 14*128 (p1) + 14*64 (idx1) = 4480, rounded up
constexpr int kBwdSmem = 2 * 2 * kBlockBytes + 2 * kBlockBytes + 2 * kImgStageBytes + kGStages * kGSlot + 1024;

template <bool COMPACT>
__global__ void __launch_bounds__(kThreads, 1) conv1_bwd_kernel(const Conv1Args a) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
  uint8_t* abuf = smem;
  uint8_t* gbuf = smem + 2 * 2 * kBlockBytes;
  uint8_t* imgbuf = gbuf + 2 * kBlockBytes;
  uint8_t* gstage = imgbuf + 2 * kImgStageBytes;
  __shared__ __align__(8) uint64_t ab_full[2], ab_empty[2], done_bar, img_full[2][2], g_full[kGStages];
  __shared__ uint32_t tmem_base_smem;
  __shared__ float s_gb[16][kO];   // bias-gradient partials of the 16 window slots

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long long total_win = static_cast<long long>(a.B) * kWinPerImg;
  const int num_tiles = static_cast<int>((total_win + kWinPerTile - 1) / kWinPerTile);

  if (threadIdx.x == 0) {
    for (int s = 0; s < 2; ++s) {
      mbar_init(&ab_full[s], 256);
      mbar_init(&ab_empty[s], 1);
      mbar_init(&img_full[s][0], 1);
      mbar_init(&img_full[s][1], 1);
    }
    mbar_init(&done_bar, 1);
    for (int g = 0; g < kGStages; ++g) mbar_init(&g_full[g], 1);
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
  } else if (warp <= 8) {
    const int grp = (warp - 1) >> 2;
    const int r = (threadIdx.x - 32) & 127;
    uint8_t* b0p = abuf + grp * 2 * kBlockBytes;
    uint8_t* img = imgbuf + grp * kImgStageBytes;   // two buffers of 2 images
    const long long t_first = blockIdx.x + static_cast<long long>(grp) * gridDim.x, t_step = 2ll * gridDim.x;
    if (r == 0 && t_first < num_tiles)
      stage_images_async(a.x, a.B, static_cast<int>(t_first * kWinPerTile / kWinPerImg), img, &img_full[grp][0]);
    int n = 0;
    for (long long t = t_first; t < num_tiles; t += t_step, ++n) {
      const int buf = n & 1;
      const int b0 = static_cast<int>(t * kWinPerTile / kWinPerImg);
      if (r == 0 && t + t_step < num_tiles)  // prefetch the next tile's images into the other buffer
        stage_images_async(a.x, a.B, static_cast<int>((t + t_step) * kWinPerTile / kWinPerImg), img + (buf ^ 1) * 2 * kImgBytes,
                           &img_full[grp][buf ^ 1]);
      mbar_wait(&img_full[grp][buf], (n >> 1) & 1);
      mbar_wait(&ab_empty[grp], (n & 1) ^ 1);
      build_patch_row<false>(reinterpret_cast<const uint32_t*>(img + buf * 2 * kImgBytes), static_cast<int>(total_win), static_cast<int>(t), b0, r, b0p,
                             b0p + kBlockBytes);
      fence_proxy_async_smem();
      mbar_arrive(&ab_full[grp]);
      asm volatile("bar.sync %0, 128;" ::"r"(2 + grp) : "memory");  // everyone is done with img[buf] before it is refilled
    }
  } else {
    // gradient rows: the pooled gradient routed to the arg-max pixel of each window (ReLU folded in).
    // Thread (window wl, chunk j) loads that window's dp1/p1/idx1 chunk once and emits the 9 pixel
    // rows of the window; the next tile's vectors are fetched before the current ones are used.
    const int gt = threadIdx.x - 288;
    const int wl = gt >> 3, j = gt & 7;          // wl 0..15 (14 real windows), j = 16-byte chunk
    // one thread streams the tiles' inputs into the ring; `issue(i)` fetches the i-th tile of this CTA
    auto issue = [&](int i) {
      const long long t = blockIdx.x + static_cast<long long>(i) * gridDim.x;
      if (t >= num_tiles) return;
      const long long w0 = t * kWinPerTile;
      const int nw = static_cast<int>(min(static_cast<long long>(kWinPerTile), total_win - w0));
      uint8_t* slot = gstage + (i % kGStages) * kGSlot;
      uint64_t* bar = &g_full[i % kGStages];
      mbar_arrive_expect_tx(bar, static_cast<uint32_t>(nw * (COMPACT ? 192 : 320)));
      if (!COMPACT)
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(slot)),
                     "l"(a.dp1 + w0 * kO), "r"(nw * 128), "r"(smem_u32(bar)) : "memory");
      asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(slot + 1792)),
                   "l"(a.p1 + w0 * kO), "r"(nw * 128), "r"(smem_u32(bar)) : "memory");
      asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(slot + 3584)),
                   "l"(a.idx1 + w0 * kO), "r"(nw * 64), "r"(smem_u32(bar)) : "memory");
    };
    if (gt == 0)
      for (int i = 0; i < kGDist; ++i) issue(i);
    uint4 dv = make_uint4(0, 0, 0, 0), pv = dv;
    uint2 iv = make_uint2(0, 0);
    bool ok = false;
    float gbias[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    int i = 0;
    for (long long t = blockIdx.x; t < num_tiles; t += gridDim.x, ++i) {
      const int s = i & 1;
      const uint32_t ph = (i >> 1) & 1;
      mbar_wait(&ab_empty[s], ph ^ 1);   // the MMA of tile i-2 has retired => every thread is done with ring slot (i-2) % kGStages
      if (gt == 0) issue(i + kGDist);      // (i + kGDist) % kGStages == (i - 2) % kGStages
      mbar_wait(&g_full[i % kGStages], (i / kGStages) & 1);
      {
        const uint8_t* slot = gstage + (i % kGStages) * kGSlot;
        const long long win = t * kWinPerTile + wl;
        ok = (wl < kWinPerTile) && (win < total_win);
        if (ok) {
          if (!COMPACT) {
            dv = *reinterpret_cast<const uint4*>(slot + wl * 128 + j * 16);
          } else {
            // compact corner [B, dph, dpw, 64]: zero outside
            const int w32 = static_cast<int>(win);  // B*81 < 2^31 is checked on the host; 32-bit division is cheap
            const int b = w32 / kWinPerImg, wi = w32 - b * kWinPerImg;
            const int py = wi / kPW, px = wi - py * kPW;
            dv = make_uint4(0, 0, 0, 0);
            if (py < a.dph && px < a.dpw)
              dv = __ldg(reinterpret_cast<const uint4*>(a.dp1 + ((static_cast<long long>(b) * a.dph + py) * a.dpw + px) * kO + j * 8));
          }
          pv = *reinterpret_cast<const uint4*>(slot + 1792 + wl * 128 + j * 16);
          iv = *reinterpret_cast<const uint2*>(slot + 3584 + wl * 64 + j * 8);
        }
      }
      uint8_t* g = gbuf + s * kBlockBytes;
      if (wl < kWinPerTile) {
        // integer domain: keep dp1 where the byte arg-max equals this pixel position and p1 != 0
        // (p1 is a ReLU output, so "> 0" is "bits != 0")
        uint32_t dw[4] = {dv.x, dv.y, dv.z, dv.w};
        const uint32_t pw[4] = {pv.x, pv.y, pv.z, pv.w};
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const uint32_t nz = ((pw[e] & 0xffffu) ? 0xffffu : 0u) | ((pw[e] & 0xffff0000u) ? 0xffff0000u : 0u);
          dw[e] = ok ? (dw[e] & nz) : 0u;
          const float2 gf = unpack_bf16x2(dw[e]);   // conv bias gradient: the pooled gradient where ReLU was active
          gbias[2 * e] += gf.x;
          gbias[2 * e + 1] += gf.y;
        }
#pragma unroll
        for (int pos = 0; pos < 9; ++pos) {
          const uint32_t m0 = __vcmpeq4(iv.x, static_cast<uint32_t>(pos) * 0x01010101u);  // channels 0..3
          const uint32_t m1 = __vcmpeq4(iv.y, static_cast<uint32_t>(pos) * 0x01010101u);  // channels 4..7
          *reinterpret_cast<uint4*>(g + sw128(wl * 9 + pos, j)) =
              make_uint4(dw[0] & __byte_perm(m0, 0, 0x1100), dw[1] & __byte_perm(m0, 0, 0x3322),
                         dw[2] & __byte_perm(m1, 0, 0x1100), dw[3] & __byte_perm(m1, 0, 0x3322));
        }
      } else if (wl == kWinPerTile) {
        // padding rows 126 and 127
        *reinterpret_cast<uint4*>(g + sw128(126, j)) = make_uint4(0, 0, 0, 0);
        *reinterpret_cast<uint4*>(g + sw128(127, j)) = make_uint4(0, 0, 0, 0);
      }
      fence_proxy_async_smem();
      mbar_arrive(&ab_full[s]);
    }
    if (a.gb != nullptr) {
#pragma unroll
      for (int e = 0; e < 8; ++e) s_gb[wl][j * 8 + e] = gbias[e];
      asm volatile("bar.sync 6, 128;" ::: "memory");
      if (gt < kO) {
        float v = 0.f;
#pragma unroll
        for (int w = 0; w < kWinPerTile; ++w) v += s_gb[w][gt];
        atomicAdd(a.gb + gt, v);
      }
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
  TADNN_REQUIRE(B > 0 && B <= 20000000, "conv1_pool_forward: batch %d out of range", B);
  TADNN_REQUIRE(((reinterpret_cast<uintptr_t>(w_packed) | reinterpret_cast<uintptr_t>(p1)) & 15) == 0, "conv1_pool_forward: operands must be 16-byte aligned");
  static std::atomic<unsigned long long> attr{0};
  if (int rc = ensure_dyn_smem(conv1_fwd_kernel, kFwdSmem, attr)) return rc;
  Conv1Args a;
  memset(&a, 0, sizeof(a));
  a.x = static_cast<const __nv_bfloat16*>(x); a.w = static_cast<const __nv_bfloat16*>(w_packed); a.bias = bias;
  a.p1 = static_cast<__nv_bfloat16*>(p1); a.idx1 = static_cast<unsigned char*>(idx1); a.B = B;
  const long long tiles = (static_cast<long long>(B) * kWinPerImg + kWinPerTile - 1) / kWinPerTile;
  const int grid = tiles < sm_count() ? static_cast<int>(tiles) : sm_count();
  conv1_fwd_kernel<<<grid, kFwdThreads, kFwdSmem, static_cast<cudaStream_t>(stream)>>>(a);
  TADNN_LAUNCH_CHECK("conv1_pool_fwd_tcgen05");
  return 0;
}

extern "C" int tadnn_conv1_pool_backward(const void* x, const void* dp1, int dp1_h, int dp1_w, const void* p1, const void* idx1,
                                         float* grad_weight, int B, tadnn_stream_t stream) {
  return tadnn_conv1_pool_backward_bias(x, dp1, dp1_h, dp1_w, p1, idx1, grad_weight, nullptr, B, stream);
}

extern "C" int tadnn_conv1_pool_backward_bias(const void* x, const void* dp1, int dp1_h, int dp1_w, const void* p1, const void* idx1,
                                              float* grad_weight, float* grad_bias, int B, tadnn_stream_t stream) {
  TADNN_REQUIRE(B > 0 && B <= 20000000, "conv1_pool_backward: batch %d out of range", B);
  TADNN_REQUIRE(dp1_h >= 1 && dp1_h <= kPH && dp1_w >= 1 && dp1_w <= kPW, "conv1_pool_backward: dp1 extent %dx%d out of range", dp1_h, dp1_w);
  TADNN_REQUIRE(((reinterpret_cast<uintptr_t>(dp1) | reinterpret_cast<uintptr_t>(p1) | reinterpret_cast<uintptr_t>(idx1)) & 15) == 0,
                "conv1_pool_backward: operands must be 16-byte aligned");
  static std::atomic<unsigned long long> attr_full{0}, attr_compact{0};
  if (int rc = ensure_dyn_smem(conv1_bwd_kernel<false>, kBwdSmem, attr_full)) return rc;
  if (int rc = ensure_dyn_smem(conv1_bwd_kernel<true>, kBwdSmem, attr_compact)) return rc;
  Conv1Args a;
  memset(&a, 0, sizeof(a));
  a.x = static_cast<const __nv_bfloat16*>(x); a.dp1 = static_cast<const __nv_bfloat16*>(dp1);
  a.p1 = const_cast<__nv_bfloat16*>(static_cast<const __nv_bfloat16*>(p1));
  a.idx1 = const_cast<unsigned char*>(static_cast<const unsigned char*>(idx1));
  a.gw = grad_weight; a.gb = grad_bias; a.B = B; a.dph = dp1_h; a.dpw = dp1_w;
  const long long tiles = (static_cast<long long>(B) * kWinPerImg + kWinPerTile - 1) / kWinPerTile;
  const int grid = tiles < sm_count() ? static_cast<int>(tiles) : sm_count();
  if (dp1_h == kPH && dp1_w == kPW) conv1_bwd_kernel<false><<<grid, kThreads, kBwdSmem, static_cast<cudaStream_t>(stream)>>>(a);
  else conv1_bwd_kernel<true><<<grid, kThreads, kBwdSmem, static_cast<cudaStream_t>(stream)>>>(a);
  TADNN_LAUNCH_CHECK("conv1_pool_bwd_tcgen05");
  return 0;
}
