// First convolution stage of the example CNN as a "sliding descriptor" implicit GEMM: the staged
// image itself is the tcgen05 operand, no im2col row is ever built.
//
//   forward : p1, idx1 = maxpool3x3/3( relu( conv5x5(x; W1) + b1 ) )
//   backward: dW1, db1 += patch(x)^T * d(conv out), d(conv out) regenerated from (dp1, idx1)
//
// Same layer as conv1_fused.cu (cifar_example/sgd-torchad_nn-cifar.lua:91-93: SpatialConvolutionMM + ReLU +
// SpatialMaxPooling 3x3/3 and their backward), different operand scheme.  conv1_fused.cu builds one 80-element
// patch row per output pixel in shared memory (20 KB of st.shared per 128 pixels, shared-memory-wavefront bound);
// here the input is stored NHWC4 (bf16 [B,32,32,4], 4th channel == 1.0 and carries the bias), so that
//
//   * one 16-byte chunk of an image row = the PAIR of pixels (2q, 2q+1) x 4 channels = 8 consecutive K elements;
//   * "position" r = (n, h, q) (pixel pair q of row h of image n) lives at byte 16*r of the input tensor, and the
//     K chunk (kh, j) of the patch of position r is simply the chunk at position r + 16*kh + j.
//
// A no-swizzle K-major UMMA operand is exactly "rows at a 16-byte pitch (8-row groups SBO = 128 B apart), the two
// 8-element K chunks of an MMA LBO bytes apart", so the operand of K chunk (kh, j) for 96 consecutive positions is
// the staged image at offset 16*(16*kh + j): overlapping rows, zero data movement.  An output pixel with even
// column w = 2q uses taps kw = 2j + p of chunks j = 0..2 (p = pixel inside the pair), one with odd column w = 2q+1
// uses kw = 2j + p - 1: the same data against two weight matrices, i.e. ONE GEMM with the 128 rows
// (column parity, channel) on the weight side:
//
//   D[(par, ch), r] = sum_{kh, j, p, c}  W2[(par, ch), (kh, j, p, c)] * x[r + 16 kh + j][p][c]
//
// 15 chunks (+1 zero-weight pad) = 8 MMAs of K = 16 per tile; a tile is N = 96 positions (6 conv rows = 2 pooled
// rows) or N = 48 (conv rows 24..26), M = 128.  The epilogue thread of lane (par, ch) holds every position of its
// channel in registers: pooling is thread-local along the accumulator columns, the two column parities are merged
// through a shared exchange.  "accumulator | (8 - window position)" keys make max + first-arg-max one integer max.
#include "common.cuh"
#include "sm100.cuh"
#include "conv1s_pack.cuh"
#include <cuda_fp16.h>
#include <stdlib.h>

#include <utility>

namespace tadnn {
namespace {

using namespace sm100;

constexpr int kO = 64, kK = 75;
constexpr int kImgBytes = 32 * 32 * 4 * 2;   // 8192: one NHWC4 image
constexpr int kPairsPerRow = 16;
constexpr int kWinPerImg = 81;

// no-swizzle shared-memory matrix descriptor (layout type 0, sm_100 "version 1")
__device__ __forceinline__ uint64_t make_smem_desc_nosw(uint32_t smem_addr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  uint64_t d = 0;
  d |= static_cast<uint64_t>((smem_addr >> 4) & 0x3FFFu);
  d |= static_cast<uint64_t>((lbo_bytes >> 4) & 0x3FFFu) << 16;
  d |= static_cast<uint64_t>((sbo_bytes >> 4) & 0x3FFFu) << 32;
  d |= static_cast<uint64_t>(1) << 46;
  return d;
}


__device__ __forceinline__ void tmem_ld_32x32b_x16(uint32_t taddr, uint32_t* r) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]),
        "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr)
      : "memory");
}

// spin on the barrier phase without the suspend-time hint of sm100::mbar_wait: a tile lasts a few hundred cycles here,
// far less than a suspended warp takes to come back
template <bool SPIN>
__device__ __forceinline__ void mbar_wait_t(uint64_t* bar, uint32_t parity) {
  if (!SPIN) {
    mbar_wait(bar, parity);
    return;
  }
  uint32_t ok;
  do {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
        "selp.u32 %0, 1, 0, p;\n"
        "}\n"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
  } while (!ok);
}
__device__ __forceinline__ void sts32(uint32_t addr, int v) { asm volatile("st.shared.b32 [%0], %1;" ::"r"(addr), "r"(v) : "memory"); }
__device__ __forceinline__ int lds32(uint32_t addr) {
  int v;
  asm volatile("ld.shared.b32 %0, [%1];" : "=r"(v) : "r"(addr) : "memory");
  return v;
}

// registers -> TMEM: this warp's 32 lanes x 32 consecutive 32-bit columns
__device__ __forceinline__ void tmem_st_32x32b_x32(uint32_t taddr, const uint32_t* r) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x32.b32 [%32], "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31};" ::"r"(r[0]),
      "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]), "r"(r[10]), "r"(r[11]),
      "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]), "r"(r[16]), "r"(r[17]), "r"(r[18]), "r"(r[19]), "r"(r[20]), "r"(r[21]),
      "r"(r[22]), "r"(r[23]), "r"(r[24]), "r"(r[25]), "r"(r[26]), "r"(r[27]), "r"(r[28]), "r"(r[29]), "r"(r[30]), "r"(r[31]),
      "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

// D[tmem] (+)= A[tmem] * B[smem desc]: the A operand (128 lanes x 8 columns of packed bf16 pairs per K = 16) is read
// from tensor memory, so only B crosses the shared-memory port
__device__ __forceinline__ void umma_f16_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t desc_b, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "setp.ne.b32 p, %4, 0;\n"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n"
      "}\n" ::"r"(tmem_d),
      "r"(tmem_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
      : "memory");
}

// K chunk order of the 8 MMAs: step s < 5 = chunks (kh = s, j = 0), (kh = s, j = 1); step 5 = (0, 2), (1, 2);
// step 6 = (2, 2), (3, 2); step 7 = (4, 2), zero-weight pad (reads chunk (4, 3): finite image data).
__host__ __device__ constexpr int step_first_chunk(int s) { return s < 5 ? s * kPairsPerRow : (s == 5 ? 2 : (s == 6 ? 2 * kPairsPerRow + 2 : 4 * kPairsPerRow + 2)); }
__host__ __device__ constexpr int step_lbo(int s) { return (s == 5 || s == 6) ? kPairsPerRow * 16 : 16; }

struct Conv1sArgs {
  const __nv_bfloat16* x;   // NHWC4 [B, 32, 32, 4], channel 3 == 1.0
  const __nv_bfloat16* w;   // packed [128, 128]: row (par, ch), column (step, half, p, c)
  __nv_bfloat16* p1;        // [B, 9, 9, 64]
  unsigned char* idx1;      // [B, 9, 9, 64]: 8 - (window position 0..8 of the arg-max), 15 where the pooled ReLU output is 0
  const __nv_bfloat16* dp1; // backward: [B, 9, 9, 64]
  float* gw;                // backward: [64, 75] Torch7 layout, +=
  float* gb;                // backward: [64], +=
  int B;
  int rows;                 // backward: pooled rows 0..rows-1 of dp1 can be non-zero
};

// ======================================= forward =======================================
// A tile = 6 conv rows (2 pooled rows, N = 96 positions) or conv rows 24..26 (N = 48): 5 tiles per image, 8 MMAs each.
// Tensor memory: four accumulator stages of 96 columns, one per epilogue group (4 warps = the 128 accumulator lanes; with
// 16 epilogue warps every scheduler has four of them to hide tcgen05.ld and shared-memory latency behind), then 64
// columns that hold the weight operand for the whole kernel (the MMAs take A from tensor memory: with A in shared
// memory an N = 96 MMA is bound by the operand port at 56 cycles instead of 48, tools/experiments/umma_rate.cu).
constexpr int kSlots = 8;                     // staged images
constexpr int kGroups = 4;
constexpr int kTilesPerImg = 5;
constexpr int kIssuers = 2;
constexpr int kFwdThreads = 32 * (kIssuers + 1 + 4 * kGroups);   // warps 0-1: MMA issuers (even / odd tiles), warp 2: loader, then the epilogue groups
constexpr int kXchBytes = 18 * kO * 4;        // per epilogue group
constexpr int kFwdSmem = kSlots * kImgBytes + 2 * kGroups * kXchBytes + 1024;
constexpr int kStageCols = 96;                // TMEM columns per accumulator stage
constexpr int kWCol = kGroups * kStageCols;   // the weights: 64 columns (8 K steps x 8 columns of bf16 pairs) behind the accumulators

// The 9 window keys of one pooled row (3 conv rows x 16 pairs of accumulator columns) for column parity PAR.
// key = fp32 accumulator with its lowest mantissa byte replaced by (8 - window position): for positive values
// (the only ones that survive the ReLU) a signed integer max is the maximum with ties going to the first position.
template <int CODE>
__device__ __forceinline__ int make_key(uint32_t v, uint32_t mask) {
  int k;
  asm("lop3.b32 %0, %1, %2, %3, 0xEA;" : "=r"(k) : "r"(v), "r"(mask), "n"(CODE));   // (v & mask) | CODE in one LOP3
  return k;
}

template <int PAR, int IQ>
__device__ __forceinline__ void key_step(const uint32_t (&v)[48], int (&key)[9], uint32_t mask) {
  constexpr int i = IQ / 14, q = IQ % 14, w = 2 * q + PAR;   // conv row i of the pooled row, pixel pair q, column w
  if constexpr (w <= 26) key[w / 3] = max(key[w / 3], make_key<8 - (i * 3 + w % 3)>(v[i * 16 + q], mask));
}
template <int PAR, int... IQ>
__device__ __forceinline__ void key_steps(const uint32_t (&v)[48], int (&key)[9], uint32_t mask, std::integer_sequence<int, IQ...>) {
  (key_step<PAR, IQ>(v, key, mask), ...);
}
template <int PAR>
__device__ __forceinline__ void pool_row_keys(const uint32_t (&v)[48], int (&key)[9]) {
  uint32_t mask;
  asm volatile("mov.u32 %0, 0xFFFFFF00;" : "=r"(mask));   // kept in a register (volatile: not folded back into an immediate)
#pragma unroll
  for (int w = 0; w < 9; ++w) key[w] = static_cast<int>(0x80000000u);
  key_steps<PAR>(v, key, mask, std::make_integer_sequence<int, 42>{});
}

template <bool SPIN>
__global__ void __launch_bounds__(kFwdThreads, 1) conv1s_fwd_kernel(const Conv1sArgs a) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
  uint8_t* imgbuf = smem;                             // kSlots images
  const uint32_t xch = smem_u32(imgbuf + kSlots * kImgBytes);
  __shared__ __align__(8) uint64_t img_full[kSlots], img_empty[kSlots], t_full[kGroups], t_empty[kGroups];
  __shared__ uint32_t tmem_base_smem;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    for (int s = 0; s < kSlots; ++s) {
      mbar_init(&img_full[s], 1);
      mbar_init(&img_empty[s], kIssuers);   // each issuer warp has at least two of the 5 tiles of an image
    }
    for (int s = 0; s < kGroups; ++s) {
      mbar_init(&t_full[s], 1);
      mbar_init(&t_empty[s], 4);      // one arrival per epilogue warp
    }
    fence_mbar_init();
  }
  if (warp == 0) tmem_alloc<512>(&tmem_base_smem);
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  const uint32_t tmem_base = tmem_base_smem;
  pdl_wait();                 // everything above touched only this CTA's shared / tensor memory
  pdl_launch_dependents();
  // The weight operand lives in tensor memory for the whole kernel: row m = (parity, channel) in lane m, its 128 K
  // values as 64 columns of bf16 pairs.  Written by the first epilogue group (one warp per lane quarter).
  if (warp >= kIssuers + 1 && warp < kIssuers + 5) {
    const int q = warp & 3;
    const uint4* src = reinterpret_cast<const uint4*>(a.w + (q * 32 + lane) * 128);
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      uint32_t r[32];
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const uint4 v = __ldg(src + h * 8 + j);
        r[4 * j] = v.x; r[4 * j + 1] = v.y; r[4 * j + 2] = v.z; r[4 * j + 3] = v.w;
      }
      tmem_st_32x32b_x32(tmem_base + (static_cast<uint32_t>(q * 32) << 16) + static_cast<uint32_t>(kWCol + h * 32), r);
    }
    tmem_st_wait();
  }
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();

  if (warp < kIssuers) {
    // ---------------- MMA issuers ----------------
    // Warp w issues the tiles tc = w, w + 2, ... (tile t of this CTA's i-th image is tc = 5 i + t, accumulator stage
    // tc & 3): the 8 MMAs of a tile are 400 cycles of math behind some 70 instructions of waits and descriptor
    // set-up, more than one warp issues in that time.  The whole warp runs the loop (warp-uniform control flow),
    // one elected lane issues.  Descriptor words that do not depend on the tile are built once.
    constexpr uint32_t idesc96 = make_idesc_bf16(128, 96, 0, 0), idesc48 = make_idesc_bf16(128, 48, 0, 0);
    const uint32_t img0 = smem_u32(imgbuf);
    const uint32_t wt = tmem_base + static_cast<uint32_t>(kWCol);
    uint64_t bdesc[8];
#pragma unroll
    for (int st = 0; st < 8; ++st)
      bdesc[st] = make_smem_desc_nosw(img0 + step_first_chunk(st) * 16, step_lbo(st), 128);   // + (slot * 512 + t * 96) in the address field
    const int my_imgs = (a.B - static_cast<int>(blockIdx.x) + static_cast<int>(gridDim.x) - 1) / static_cast<int>(gridDim.x);
    const int my_tiles = my_imgs * kTilesPerImg;
    int last_i = -1;
#pragma unroll 1
    for (int tc = warp; tc < my_tiles; tc += kIssuers) {
      const int s = tc & (kGroups - 1);
      const uint32_t d = tmem_base + static_cast<uint32_t>(s * kStageCols);
      const int i = tc / kTilesPerImg, t = tc - i * kTilesPerImg;
      const int slot = i % kSlots;
      if (i != last_i) {
        if (last_i >= 0 && elect_one()) umma_commit(&img_empty[last_i % kSlots]);   // this warp's MMAs on that image are all issued
        __syncwarp();
        mbar_wait_t<SPIN>(&img_full[slot], (i / kSlots) & 1);
        last_i = i;
      }
      mbar_wait_t<SPIN>(&t_empty[s], ((tc >> 2) & 1) ^ 1);
      tc_fence_after_sync();
      const uint64_t boff = static_cast<uint64_t>(slot * (kImgBytes / 16) + t * 96);
      if (elect_one()) {
        if (t < 4) {
#pragma unroll
          for (int st = 0; st < 8; ++st) umma_f16_ts(d, wt + st * 8, bdesc[st] + boff, idesc96, st > 0);
        } else {
#pragma unroll
          for (int st = 0; st < 8; ++st) umma_f16_ts(d, wt + st * 8, bdesc[st] + boff, idesc48, st > 0);
        }
        umma_commit(&t_full[s]);
      }
      __syncwarp();
    }
    if (last_i >= 0 && elect_one()) umma_commit(&img_empty[last_i % kSlots]);
    __syncwarp();
  } else if (warp == kIssuers) {
    // ---------------- loader: one 8 KB bulk copy per image ----------------
    if (elect_one()) {
      int i = 0;
      for (int n = blockIdx.x; n < a.B; n += gridDim.x, ++i) {
        const int slot = i % kSlots;
        mbar_wait_t<true>(&img_empty[slot], ((i / kSlots) & 1) ^ 1);
        mbar_arrive_expect_tx(&img_full[slot], kImgBytes);
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                         smem_u32(imgbuf + slot * kImgBytes)),
                     "l"(a.x + static_cast<long long>(n) * (kImgBytes / 2)), "r"(kImgBytes), "r"(smem_u32(&img_full[slot]))
                     : "memory");
      }
    }
    __syncwarp();
  } else {
    // ---------------- epilogue: pooling in registers, parity merge through shared memory ----------------
    const int grp = (warp - kIssuers - 1) >> 2;
    const int q = warp & 3;                  // TMEM lane quarter
    const int par = q >> 1;                  // column parity of this thread's accumulator row
    const int ch = (q & 1) * 32 + lane;
    const uint32_t xg = xch + static_cast<uint32_t>(grp * kXchBytes + ch * 4);
    const uint32_t taddr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + static_cast<uint32_t>(grp * kStageCols);
    unsigned short* p1u = reinterpret_cast<unsigned short*>(a.p1);
    const int my_imgs = (a.B - static_cast<int>(blockIdx.x) + static_cast<int>(gridDim.x) - 1) / static_cast<int>(gridDim.x);
    const int my_tiles = my_imgs * kTilesPerImg;
    // this group's tiles: tc = grp, grp + 4, ... (tile t of this CTA's i-th image is tc = 5 i + t)
#pragma unroll 1
    for (int tc = grp; tc < my_tiles; tc += kGroups) {
      const int i = tc / kTilesPerImg, t = tc - i * kTilesPerImg;
      const long long n = static_cast<long long>(blockIdx.x) + static_cast<long long>(i) * gridDim.x;
      mbar_wait_t<SPIN>(&t_full[grp], (tc >> 2) & 1);
      tc_fence_after_sync();
      const bool two = t < 4;        // two pooled rows (18 windows), or the last one of the image (9 windows)
      int key[18];
      // the accumulator stage goes back to the MMA issuers as soon as its last column is in registers, before the keys
      // of that row are computed
      auto release = [&]() {
        tc_fence_before_sync();
        __syncwarp();
        if (lane == 0) mbar_arrive(&t_empty[grp]);   // 128 arrivals on one shared-memory word would serialise
      };
      {
        uint32_t v[48];
        tmem_ld_32x32b_x32(taddr, *reinterpret_cast<uint32_t(*)[32]>(&v[0]));
        tmem_ld_32x32b_x16(taddr + 32, &v[32]);
        tmem_ld_wait();
        if (!two) release();
        if (par == 0) pool_row_keys<0>(v, *reinterpret_cast<int(*)[9]>(&key[0]));
        else pool_row_keys<1>(v, *reinterpret_cast<int(*)[9]>(&key[0]));
      }
      if (two) {
        uint32_t v[48];
        tmem_ld_32x32b_x32(taddr + 48, *reinterpret_cast<uint32_t(*)[32]>(&v[0]));
        tmem_ld_32x32b_x16(taddr + 80, &v[32]);
        tmem_ld_wait();
        release();
        if (par == 0) pool_row_keys<0>(v, *reinterpret_cast<int(*)[9]>(&key[9]));
        else pool_row_keys<1>(v, *reinterpret_cast<int(*)[9]>(&key[9]));
      }
      // Window wi is finished by the thread whose parity equals wi & 1; the other one hands its key over through the
      // exchange buffer of this tile's parity (double-buffered: one barrier per tile is enough).
      const uint32_t xb = xg + static_cast<uint32_t>(((tc >> 2) & 1) * (kGroups * kXchBytes));
#pragma unroll
      for (int wi = 0; wi < 9; ++wi)
        if ((wi & 1) != par) sts32(xb + wi * (kO * 4), key[wi]);
      if (two) {
#pragma unroll
        for (int wi = 9; wi < 18; ++wi)
          if ((wi & 1) != par) sts32(xb + wi * (kO * 4), key[wi]);
      }
      asm volatile("bar.sync %0, 128;" ::"r"(2 + grp) : "memory");
      unsigned short* po = p1u + (n * kWinPerImg + t * 18) * kO + ch;
      unsigned char* io = a.idx1 + (n * kWinPerImg + t * 18) * kO + ch;
      // max of the two parities and 0 (ReLU) in one 3-input max; a dead window has no value bits left: key <= 255
      auto finish = [&](int wi) {
        const int k = max(max(key[wi], lds32(xb + wi * (kO * 4))), 0);
        const __nv_bfloat16 hv = __float2bfloat16_rn(__int_as_float(k));   // the code byte is far below bf16 resolution
        po[wi * kO] = *reinterpret_cast<const unsigned short*>(&hv);
        io[wi * kO] = static_cast<unsigned char>(k > 255 ? k : 15);
      };
#pragma unroll
      for (int wi = 0; wi < 9; ++wi)
        if ((wi & 1) == par) finish(wi);
      if (two) {
#pragma unroll
        for (int wi = 9; wi < 18; ++wi)
          if ((wi & 1) == par) finish(wi);
      }
    }
  }
  tc_fence_before_sync();
  __syncthreads();
  if (warp == 0) {
    tc_fence_after_sync();
    tmem_dealloc<512>(tmem_base);
  }
}

// ======================================= backward =======================================
//   D_j[(par, ch), (kh', p, c)] += sum over positions r = (h, q) of  dY[(par, ch), r] * x[r + 16 kh' + j][p][c]      j = 0, 1, 2
// and dW[ch][c][kh][kw] = sum over (par, j, p) with 2 j + p - par == kw of D_j[(par, ch), (kh, p, c)];  channel c = 3 is the
// constant 1.0, so its tap (kh, kw) = (0, 0) is the bias gradient.
//   A = dY^T, in TENSOR MEMORY: lane (par, ch), K = the 16 pair positions of one conv row = 8 columns of bf16 pairs,
//       written with tcgen05.st by the thread of that lane from (dp1, idx1): the pooled gradient goes to the window's
//       arg-max position when the window is alive and that position has this lane's column parity.
//   B = the staged image, MN-major no-swizzle: N = (kh' = 0..5, 8 elements of a pair chunk) with the kh' groups one
//       image row (SBO = 256 B) apart, K = 16 positions at a 16-byte pitch (8-position groups LBO = 128 B apart);
//       kh' = 5 is read (N must be a multiple of 16) and ignored.
// One K step = one conv row, 3 MMAs (j = 0, 1, 2; M = 128, N = 48) into three accumulators that live for the whole
// kernel; a pooled row = 3 conv rows = 24 operand columns; the operand ring in tensor memory has two slots of 7 pooled
// rows (the 7 rows the example CNN back-propagates through = one sample; 9 rows = a slot of 7 and a slot of 2), so an
// issuer waits once and issues 21 MMAs back to back.
constexpr int kXS = 8;                         // staged images (B operand)
constexpr int kGS = 4;                         // staged (dp1, idx1) images
constexpr int kGIdxOff = 81 * kO * 2;          // 10368: idx1 part of a gradient slot
constexpr int kGSlotBytes = 15616;             // 10368 + 5184 = 15552, rounded up to a multiple of 128
constexpr int kDStride = 64;                   // accumulator j: columns 64 j .. 64 j + 47
constexpr int kASlots = 2, kRowCols = 24, kSlotRows = 7, kASlotCols = kSlotRows * kRowCols;
constexpr int kACol0 = 176;                    // operand ring: columns 176 .. 511
static_assert(kACol0 + kASlots * kASlotCols <= 512, "operand ring exceeds tensor memory");
constexpr int kBGroups = 4;
constexpr int kBwdThreads = 32 * (4 + 4 * kBGroups);   // warps 0-2: MMA issuers (one per accumulator j), warp 3: loader, then the builder groups
constexpr int kBwdSmem = kXS * kImgBytes + 128 + kGS * kGSlotBytes + 2 * 64 * 77 * 4 + 1024;

__device__ __forceinline__ void tmem_st_32x32b_x8(uint32_t taddr, const uint32_t* r) {
  asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%8], {%0, %1, %2, %3, %4, %5, %6, %7};" ::"r"(r[0]), "r"(r[1]), "r"(r[2]),
               "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(taddr)
               : "memory");
}
__device__ __forceinline__ uint32_t lds_u16(uint32_t addr) {
  uint32_t v;
  asm volatile("ld.shared.u16 %0, [%1];" : "=r"(v) : "r"(addr) : "memory");
  return v;
}
__device__ __forceinline__ uint32_t lds_u8(uint32_t addr) {
  uint32_t v;
  asm volatile("ld.shared.u8 %0, [%1];" : "=r"(v) : "r"(addr) : "memory");
  return v;
}
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst), "l"(src), "r"(bytes),
               "r"(smem_u32(bar))
               : "memory");
}

// dY^T of one pooled row for column parity PAR: word K2 of conv row I holds positions q = 2 K2 (low half) and 2 K2 + 1
// (high half), i.e. columns w = 4 K2 + PAR and 4 K2 + 2 + PAR.  A position gets its window's pooled gradient when the
// window's arg-max code equals the position's code (8 - (3 I + w % 3)); codes are compared two at a time as the fp16
// numbers their bit patterns denote (0..15: zero and subnormals, all distinct; fp16 compares do not flush them), the
// pair of gradients and the pair of codes are shared by the three rows.
template <int PAR, int K2>
__device__ __forceinline__ void dy_words(const uint32_t (&g)[9], const uint32_t (&cdh)[9], uint32_t& r0, uint32_t& r1, uint32_t& r2) {
  constexpr int wa = 4 * K2 + PAR, wb = 4 * K2 + 2 + PAR;
  constexpr bool vb = wb <= 26;
  constexpr int pa = wa / 3, pb = vb ? wb / 3 : pa;
  const uint32_t P = __byte_perm(g[pa], g[pb], 0x5410);
  const uint32_t C = __byte_perm(cdh[pa], cdh[pb], 0x5410);
  const __half2 c2 = *reinterpret_cast<const __half2*>(&C);
#define TADNN_DY_ROW(I, OUT)                                                                                          \
  {                                                                                                                   \
    constexpr uint32_t k = static_cast<uint32_t>(8 - (3 * I + wa % 3)) |                                              \
                           ((vb ? static_cast<uint32_t>(8 - (3 * I + wb % 3)) : 0x7C00u) << 16);                      \
    const uint32_t kk = k;                                                                                            \
    OUT = P & __heq2_mask(c2, *reinterpret_cast<const __half2*>(&kk));                                               \
  }
  TADNN_DY_ROW(0, r0)
  TADNN_DY_ROW(1, r1)
  TADNN_DY_ROW(2, r2)
#undef TADNN_DY_ROW
}
template <int PAR>
__device__ __forceinline__ void dy_rows(const uint32_t (&g)[9], const uint32_t (&cdh)[9], uint32_t (&r0)[8], uint32_t (&r1)[8], uint32_t (&r2)[8]) {
  dy_words<PAR, 0>(g, cdh, r0[0], r1[0], r2[0]);
  dy_words<PAR, 1>(g, cdh, r0[1], r1[1], r2[1]);
  dy_words<PAR, 2>(g, cdh, r0[2], r1[2], r2[2]);
  dy_words<PAR, 3>(g, cdh, r0[3], r1[3], r2[3]);
  dy_words<PAR, 4>(g, cdh, r0[4], r1[4], r2[4]);
  dy_words<PAR, 5>(g, cdh, r0[5], r1[5], r2[5]);
  dy_words<PAR, 6>(g, cdh, r0[6], r1[6], r2[6]);
  r0[7] = r1[7] = r2[7] = 0u;
}

// launch bound of 896 threads = at most 72 registers: with the 640 threads actually launched that leaves room on the SM for a
// 512-thread communication kernel of the data-parallel step to run next to this one
__global__ void __launch_bounds__(896, 1) conv1s_bwd_kernel(const Conv1sArgs a) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
  uint8_t* ximg = smem;                                   // kXS images (+128 B of zeros: kh' = 5 of the last rows reads past the last one)
  uint8_t* gbuf = smem + kXS * kImgBytes + 128;           // kGS x (dp1 rows | idx1 rows)
  __shared__ __align__(8) uint64_t x_full[kXS], x_empty[kXS], g_full[kGS], g_empty[kGS], a_full[kASlots], a_empty[kASlots], done_bar;
  __shared__ uint32_t tmem_base_smem;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int rows = a.rows;
  if (threadIdx.x == 0) {
    for (int s = 0; s < kXS; ++s) { mbar_init(&x_full[s], 1); mbar_init(&x_empty[s], 3); }
    // consumer barriers count one arrival per warp (lane 0 after __syncwarp): 128 arrivals on one shared-memory word serialise
    for (int s = 0; s < kGS; ++s) { mbar_init(&g_full[s], 1); mbar_init(&g_empty[s], static_cast<uint32_t>(rows * 4)); }
    // slot 0 takes pooled rows 0..6 of a sample, slot 1 the next sample's (rows <= 7) or rows 7, 8 of the same one
    mbar_init(&a_full[0], static_cast<uint32_t>((rows < kSlotRows ? rows : kSlotRows) * 4));
    mbar_init(&a_full[1], static_cast<uint32_t>((rows > kSlotRows ? rows - kSlotRows : rows) * 4));
    for (int s = 0; s < kASlots; ++s) mbar_init(&a_empty[s], 3);
    mbar_init(&done_bar, 3);
    fence_mbar_init();
  }
  if (warp == 0) tmem_alloc<512>(&tmem_base_smem);
  if (threadIdx.x < 32) reinterpret_cast<uint32_t*>(ximg + kXS * kImgBytes)[threadIdx.x] = 0u;
  fence_proxy_async_smem();
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  const uint32_t tmem_base = tmem_base_smem;
  pdl_wait();                 // everything above touched only this CTA's shared / tensor memory
  pdl_launch_dependents();
  const int my_imgs = (a.B - static_cast<int>(blockIdx.x) + static_cast<int>(gridDim.x) - 1) / static_cast<int>(gridDim.x);

  if (warp < 3) {
    // ---------------- MMA issuers: warp j owns accumulator j (whole warp in the loop, one elected lane issues) ----------------
    // One warp issues an instruction every ~8 cycles here (waits, uniform-register set-up, the MMAs themselves): with
    // all 63 MMAs of a sample in one instruction stream the tensor pipe was busy 40 % of the time.
    constexpr uint32_t idesc = make_idesc_bf16(128, 48, 0, 1);
    const int j = warp;
    const uint32_t x0 = smem_u32(ximg);
    const uint64_t bdesc0 = make_smem_desc_nosw(x0, 128, 256) + static_cast<uint64_t>(j);   // MN-major no-swizzle: LBO = next 8 K rows, SBO = next 8 N elements
    const uint32_t acol = tmem_base + static_cast<uint32_t>(kACol0);
    const uint32_t d = tmem_base + static_cast<uint32_t>(j * kDStride);
    const int nsc = rows > kSlotRows ? 2 : 1;   // operand slots per sample
#pragma unroll 1
    for (int i = 0; i < my_imgs; ++i) {
      const int xs = i % kXS;
      mbar_wait_t<true>(&x_full[xs], (i / kXS) & 1);
      const uint64_t bimg = bdesc0 + static_cast<uint64_t>(xs * (kImgBytes / 16));
#pragma unroll
      for (int c = 0; c < 2; ++c) {
        if (c < nsc) {
          const int sc = i * nsc + c, slot = sc & 1;
          mbar_wait_t<true>(&a_full[slot], (sc >> 1) & 1);
          tc_fence_after_sync();
          const uint32_t at = acol + static_cast<uint32_t>(slot * kASlotCols);
          if (elect_one()) {
#pragma unroll
            for (int prl = 0; prl < kSlotRows; ++prl)
              if (c * kSlotRows + prl < rows) {
#pragma unroll
                for (int r = 0; r < 3; ++r)
                  umma_f16_ts(d, at + prl * kRowCols + r * 8, bimg + static_cast<uint64_t>((3 * (c * kSlotRows + prl) + r) * kPairsPerRow), idesc,
                              (i > 0 || c > 0 || prl > 0 || r > 0) ? 1u : 0u);
              }
            umma_commit(&a_empty[slot]);
          }
          __syncwarp();
        }
      }
      if (elect_one()) umma_commit(&x_empty[xs]);
      __syncwarp();
    }
    if (elect_one()) umma_commit(&done_bar);
    __syncwarp();
  } else if (warp == 3) {
    // ---------------- loader: image, dp1 rows and idx1 rows of one sample per step ----------------
    if (elect_one()) {
      const uint32_t gbytes = static_cast<uint32_t>(rows * 9 * kO * 2), ibytes = static_cast<uint32_t>(rows * 9 * kO);
      for (int i = 0; i < my_imgs; ++i) {
        const long long n = static_cast<long long>(blockIdx.x) + static_cast<long long>(i) * gridDim.x;
        const int xs = i % kXS, gs = i % kGS;
        mbar_wait_t<true>(&g_empty[gs], ((i / kGS) & 1) ^ 1);
        mbar_arrive_expect_tx(&g_full[gs], gbytes + ibytes);
        bulk_g2s(smem_u32(gbuf + gs * kGSlotBytes), a.dp1 + n * (kWinPerImg * kO), gbytes, &g_full[gs]);
        bulk_g2s(smem_u32(gbuf + gs * kGSlotBytes + kGIdxOff), a.idx1 + n * (kWinPerImg * kO), ibytes, &g_full[gs]);
        mbar_wait_t<true>(&x_empty[xs], ((i / kXS) & 1) ^ 1);
        mbar_arrive_expect_tx(&x_full[xs], kImgBytes);
        bulk_g2s(smem_u32(ximg + xs * kImgBytes), a.x + n * (kImgBytes / 2), kImgBytes, &x_full[xs]);
      }
    }
    __syncwarp();
  } else {
    // ---------------- builders: dY^T of one pooled row (3 conv rows) per step, straight into tensor memory ----------------
    const int grp = (warp - 4) >> 2;
    const int q = warp & 3;
    const int par = q >> 1;
    const int ch = (q & 1) * 32 + lane;
    const uint32_t g0 = smem_u32(gbuf) + static_cast<uint32_t>(ch * 2), i0 = smem_u32(gbuf) + static_cast<uint32_t>(kGIdxOff + ch);
    const uint32_t arow = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + static_cast<uint32_t>(kACol0);
    // pooled row cg = i * rows + pr of this CTA's work; this group takes cg = grp, grp + 4, ...
    const int nsc = rows > kSlotRows ? 2 : 1;
    int i = 0, pr = grp, waited = -1;
    while (pr >= rows) { pr -= rows; ++i; }
#pragma unroll 1
    while (i < my_imgs) {
      const int gs = i & (kGS - 1);
      if (i != waited) {   // this thread still owes an arrival on g_empty for every row it reads: the slot cannot be refilled under it
        mbar_wait_t<true>(&g_full[gs], (i / kGS) & 1);
        waited = i;
      }
      uint32_t g[9], cdh[9];
      const uint32_t goff = static_cast<uint32_t>(gs * kGSlotBytes + pr * 9 * kO * 2), ioff = static_cast<uint32_t>(gs * kGSlotBytes + pr * 9 * kO);
#pragma unroll
      for (int pw = 0; pw < 9; ++pw) {
        g[pw] = lds_u16(g0 + goff + pw * (kO * 2));
        cdh[pw] = lds_u8(i0 + ioff + pw * kO);
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(&g_empty[gs]);
      uint32_t r0[8], r1[8], r2[8];
      if (par == 0) dy_rows<0>(g, cdh, r0, r1, r2);
      else dy_rows<1>(g, cdh, r0, r1, r2);
      const int c = pr >= kSlotRows ? 1 : 0;
      const int sc = i * nsc + c, slot = sc & 1;
      mbar_wait_t<true>(&a_empty[slot], ((sc >> 1) & 1) ^ 1);
      tc_fence_after_sync();
      const uint32_t at = arow + static_cast<uint32_t>(slot * kASlotCols + (pr - c * kSlotRows) * kRowCols);
      tmem_st_32x32b_x8(at, r0);
      tmem_st_32x32b_x8(at + 8, r1);
      tmem_st_32x32b_x8(at + 16, r2);
      tmem_st_wait();
      tc_fence_before_sync();
      __syncwarp();
      if (lane == 0) mbar_arrive(&a_full[slot]);
      pr += kBGroups;
      while (pr >= rows) { pr -= rows; ++i; }
    }
    // ---------------- final reduction: accumulators -> gw, gb ----------------
    // The first builder group moves its accumulator rows to shared memory in Torch7 order, one plane per column
    // parity; then every builder thread adds the two planes into the global gradients.  CTA b starts its sweep at a
    // different element: the CTAs finish together, and 148 atomics queued on the same line at the same time serialise.
    const uint32_t red = smem_u32(gbuf + kGS * kGSlotBytes);   // [2][64][77] floats
    if (my_imgs > 0) {
      if (grp == 0) {
        mbar_wait_t<true>(&done_bar, 0);
        tc_fence_after_sync();
        const uint32_t drow = tmem_base + (static_cast<uint32_t>(q * 32) << 16);
        const uint32_t mine = red + static_cast<uint32_t>((par * kO + ch) * 77 * 4);
#pragma unroll
        for (int j = 0; j < 3; ++j) {
          uint32_t v[48];
          tmem_ld_32x32b_x32(drow + j * kDStride, *reinterpret_cast<uint32_t(*)[32]>(&v[0]));
          tmem_ld_32x32b_x16(drow + j * kDStride + 32, &v[32]);
          tmem_ld_wait();
#pragma unroll
          for (int kh = 0; kh < 5; ++kh)
#pragma unroll
            for (int e = 0; e < 8; ++e) {
              const int p = e >> 2, c = e & 3;
              // k = Torch7 index (c*5+kh)*5+kw for a weight tap, 75 for the bias (channel 3 is the constant 1.0)
              if (par == 0) {
                const int kw = 2 * j + p;
                if (kw <= 4 && (c < 3 || (kh == 0 && kw == 0))) sts32(mine + (c < 3 ? (c * 5 + kh) * 5 + kw : 75) * 4, static_cast<int>(v[kh * 8 + e]));
              } else {
                const int kw = 2 * j + p - 1;
                if (kw >= 0 && kw <= 4 && (c < 3 || (kh == 0 && kw == 0))) sts32(mine + (c < 3 ? (c * 5 + kh) * 5 + kw : 75) * 4, static_cast<int>(v[kh * 8 + e]));
              }
            }
        }
      }
      asm volatile("bar.sync 2, %0;" ::"n"(128 * kBGroups) : "memory");
      constexpr int kFlushThreads = 128 * kBGroups, kElems = kO * 76;
      constexpr int kSweeps = (kElems + kFlushThreads - 1) / kFlushThreads;
      const int t = threadIdx.x - 128;
      for (int it = 0; it < kSweeps; ++it) {
        const int e = ((it + static_cast<int>(blockIdx.x)) % kSweeps) * kFlushThreads + t;
        if (e < kElems) {
          const int c2 = e / 76, k = e - c2 * 76;
          const float val = __int_as_float(lds32(red + (c2 * 77 + k) * 4)) + __int_as_float(lds32(red + ((kO + c2) * 77 + k) * 4));
          if (k < kK) atomicAdd(a.gw + c2 * kK + k, val);
          else if (a.gb != nullptr) atomicAdd(a.gb + c2, val);
        }
      }
    }
  }
  tc_fence_before_sync();
  __syncthreads();
  if (warp == 0) {
    tc_fence_after_sync();
    tmem_dealloc<512>(tmem_base);
  }
}

// Torch7 weight [64, 75] fp32 (k = (c*5+kh)*5+kw) + bias [64]  ->  bf16 [128, 128]
__global__ void conv1s_weight_pack_kernel(const float* __restrict__ w, const float* __restrict__ bias, __nv_bfloat16* __restrict__ wp) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < 128 * 128) wp[i] = __float2bfloat16_rn(conv1s_weight_value(w, bias, i));
}

// NCHW {uint8, bf16, fp32} [B, 3, 32, 32] -> NHWC4 bf16 [B, 32, 32, 4]: y = x * scale[c] + shift[c], channel 3 = 1.0
template <typename T>
__device__ __forceinline__ float ld_as_float(const T* p) { return static_cast<float>(*p); }
template <>
__device__ __forceinline__ float ld_as_float<__nv_bfloat16>(const __nv_bfloat16* p) { return __bfloat162float(*p); }

struct Affine3 { float scale[3], shift[3]; };

template <typename T>
__global__ void nhwc4_pack_kernel(const T* __restrict__ x, __nv_bfloat16* __restrict__ y, long long npix, const Affine3 af) {
  const long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= npix) return;
  const long long n = i >> 10;
  const int hw = static_cast<int>(i & 1023);
  const T* src = x + n * 3072 + hw;
  const float c0 = ld_as_float(src) * af.scale[0] + af.shift[0];
  const float c1 = ld_as_float(src + 1024) * af.scale[1] + af.shift[1];
  const float c2 = ld_as_float(src + 2048) * af.scale[2] + af.shift[2];
  *reinterpret_cast<uint2*>(y + i * 4) = make_uint2(pack_bf16x2(c0, c1), pack_bf16x2(c2, 1.0f));
}

}  // namespace
}  // namespace tadnn

using namespace tadnn;

extern "C" int tadnn_conv1s_weight_pack(const float* w, const float* bias, void* wp, tadnn_stream_t stream) {
  conv1s_weight_pack_kernel<<<64, 256, 0, static_cast<cudaStream_t>(stream)>>>(w, bias, static_cast<__nv_bfloat16*>(wp));
  TADNN_LAUNCH_CHECK("conv1s_weight_pack");
  return 0;
}

extern "C" int tadnn_nhwc4_pack(const void* x, int src_dtype, void* y, long long B, const float* host_scale, const float* host_shift,
                                tadnn_stream_t stream) {
  TADNN_REQUIRE(B > 0 && B <= 20000000, "nhwc4_pack: batch %lld out of range", B);
  TADNN_REQUIRE((reinterpret_cast<uintptr_t>(y) & 7) == 0, "nhwc4_pack: output must be 8-byte aligned");
  Affine3 af;
  for (int c = 0; c < 3; ++c) {
    af.scale[c] = host_scale ? host_scale[c] : 1.f;
    af.shift[c] = host_shift ? host_shift[c] : 0.f;
  }
  const long long npix = B * 1024;
  const unsigned grid = static_cast<unsigned>((npix + 255) / 256);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (src_dtype == 0) nhwc4_pack_kernel<unsigned char><<<grid, 256, 0, s>>>(static_cast<const unsigned char*>(x), static_cast<__nv_bfloat16*>(y), npix, af);
  else if (src_dtype == 1) nhwc4_pack_kernel<__nv_bfloat16><<<grid, 256, 0, s>>>(static_cast<const __nv_bfloat16*>(x), static_cast<__nv_bfloat16*>(y), npix, af);
  else if (src_dtype == 2) nhwc4_pack_kernel<float><<<grid, 256, 0, s>>>(static_cast<const float*>(x), static_cast<__nv_bfloat16*>(y), npix, af);
  else TADNN_REQUIRE(false, "nhwc4_pack: source dtype %d (0 = uint8, 1 = bf16, 2 = fp32)", src_dtype);
  TADNN_LAUNCH_CHECK("nhwc4_pack");
  return 0;
}

extern "C" int tadnn_conv1s_pool_forward(const void* x_nhwc4, const void* w_packed, void* p1, void* idx1, int B, tadnn_stream_t stream) {
  TADNN_REQUIRE(B > 0 && B <= 20000000, "conv1s_pool_forward: batch %d out of range", B);
  TADNN_REQUIRE(((reinterpret_cast<uintptr_t>(x_nhwc4) | reinterpret_cast<uintptr_t>(w_packed)) & 15) == 0,
                "conv1s_pool_forward: operands must be 16-byte aligned");
  static std::atomic<unsigned long long> attr{0};
  if (int rc = ensure_dyn_smem(conv1s_fwd_kernel<true>, kFwdSmem, attr)) return rc;
  Conv1sArgs a;
  memset(&a, 0, sizeof(a));
  a.x = static_cast<const __nv_bfloat16*>(x_nhwc4); a.w = static_cast<const __nv_bfloat16*>(w_packed);
  a.p1 = static_cast<__nv_bfloat16*>(p1); a.idx1 = static_cast<unsigned char*>(idx1); a.B = B;
  const int grid = B < sm_count() ? B : sm_count();
  TADNN_CHECK_CUDA(launch_pdl(conv1s_fwd_kernel<true>, dim3(grid), dim3(kFwdThreads), kFwdSmem, static_cast<cudaStream_t>(stream), a));
  TADNN_LAUNCH_CHECK("conv1s_pool_fwd_tcgen05");
  return 0;
}

extern "C" int tadnn_conv1s_pool_backward(const void* x_nhwc4, const void* dp1, const void* idx1, float* grad_weight, float* grad_bias,
                                          int rows, int B, tadnn_stream_t stream) {
  TADNN_REQUIRE(B > 0 && B <= 20000000, "conv1s_pool_backward: batch %d out of range", B);
  TADNN_REQUIRE(rows >= 1 && rows <= 9, "conv1s_pool_backward: %d pooled rows (1..9)", rows);
  TADNN_REQUIRE(((reinterpret_cast<uintptr_t>(x_nhwc4) | reinterpret_cast<uintptr_t>(dp1) | reinterpret_cast<uintptr_t>(idx1)) & 15) == 0,
                "conv1s_pool_backward: operands must be 16-byte aligned");
  static std::atomic<unsigned long long> attr{0};
  if (int rc = ensure_dyn_smem(conv1s_bwd_kernel, kBwdSmem, attr)) return rc;
  Conv1sArgs a;
  memset(&a, 0, sizeof(a));
  a.x = static_cast<const __nv_bfloat16*>(x_nhwc4); a.dp1 = static_cast<const __nv_bfloat16*>(dp1);
  a.idx1 = const_cast<unsigned char*>(static_cast<const unsigned char*>(idx1));
  a.gw = grad_weight; a.gb = grad_bias; a.B = B; a.rows = rows;
  const int grid = B < sm_count() ? B : sm_count();
  TADNN_CHECK_CUDA(launch_pdl(conv1s_bwd_kernel, dim3(grid), dim3(kBwdThreads), kBwdSmem, static_cast<cudaStream_t>(stream), a));
  TADNN_LAUNCH_CHECK("conv1s_pool_bwd_tcgen05");
  return 0;
}
