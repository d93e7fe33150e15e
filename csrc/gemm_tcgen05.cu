// Persistent, warp-specialised bf16 GEMM for sm_100a:
//   TMA (cp.async.bulk.tensor, SWIZZLE_128B)  ->  shared memory ring (mbarrier full/empty)
//   -> tcgen05.mma (one elected thread, fp32 accumulators in TMEM, double-buffered)
//   -> tcgen05.ld epilogue (bias + tanh/relu + cast, optional replicated "all-gather"
//      stores to peer-mapped buffers, optional split-K red.global.add).
//
// Replaces the cuBLAS SGEMM calls under nn.Linear in the reference
// (nodemodule.lua:47,65,72,95,116) and under SpatialConvolutionMM in its example
// (cifar_example/sgd-torchad_nn-cifar.lua:91,96).  The fused epilogue replaces
// THNN.Tanh_updateOutput (nodemodule.lua:134) and, with n_push>0, the collective that
// follows the GEMM in MPInitialLinear/MPBaseLinear (nodemodule.lua:52,103).
//
// Warp roles (256 threads, 1 CTA/SM):
//   warp 0 : TMA producer (one elected lane)
//   warp 1 : MMA issuer   (one elected lane)
//   warp 2 : TMEM allocator / deallocator
//   warp 3 : idle
//   warps 4-7 : epilogue (warp w owns TMEM lanes 32*(w%4) .. +31)
#include "common.cuh"
#include "sm100.cuh"

#include <map>
#include <mutex>
#include <tuple>

namespace tadnn {
namespace {

using namespace sm100;

constexpr int BM = 128;
constexpr int BK = 64;  // 64 bf16 = 128 B = one swizzle row
constexpr int kThreads = 256;
constexpr int kEpiThreads = 128;
constexpr int kABytes = BM * BK * 2;

template <int BN>
struct Cfg {
  static constexpr int kBBytes = BN * BK * 2;
  static constexpr int kStageBytes = kABytes + kBBytes;
  // keep the ring <= ~200 KB
  static constexpr int kStages = (BN == 256) ? 4 : (BN == 128 ? 6 : 8);
  static constexpr int kStagingBytes = 4 * 2 * 4096;  // 4 epilogue warps x 2 buffers x (32 rows x 128 B)
  static constexpr int kSmemBytes = kStages * kStageBytes + kStagingBytes + 1024;  // +1024 alignment slack
  static constexpr uint32_t kTmemCols = (2 * BN <= 128) ? 128 : (2 * BN <= 256 ? 256 : 512);
};

struct DevArgs {
  int M, N, K;
  void* C;
  int ldc;
  int c_fp32;
  const float* bias;
  int act;
  float alpha;
  int atomic;  // accumulate or split-K: red.global.add into fp32 C
  int split_k;
  int n_push;
  void* push[8];
  int push_ldc;
  int push_col_offset;
  int tma_store;  // bf16 output through shared memory + cp.async.bulk.tensor stores
  int rs_width;   // >0: reduce-scatter push: column block q = col/rs_width goes ONLY to push map q
  int rs_slab;    // rs push writes 4 KB slabs (32 rows x 64 cols, swizzled as staged) contiguously: slab-major staging
  const __nv_bfloat16* aux;  // optional [M, ld_aux] bf16; epilogue multiplies by (1 - aux^2) (tanh backward)
  int ld_aux;
  int group_m;   // m-blocks per rasterisation group
  int n_rotate;  // tile column order starts at this n-block (reduce-scatter push: de-synchronises the ranks' targets)
  // fused all-gather -> GEMM (see tadnn_gemm_args)
  uint32_t* sig[8];
  int n_sig, sig_src, flag_stride;
  const uint32_t* wait_flags;
  uint32_t wait_target;
  int wait_cols_per_src;
  // structurally sparse B: bit kb of kmask[n_blk] clear = k-block kb of column tile n_blk is all zeros and is skipped
  int use_kmask;
  uint32_t kmask[32];
};

__device__ __forceinline__ uint32_t ld_acquire_sys_u32(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void red_release_sys_add_u32(uint32_t* p, uint32_t v) {
  asm volatile("red.release.sys.global.add.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
// Consumer side of the fused all-gather: spin until a peer's epilogue has delivered one (32-row block, source)
// pair.  Gives up after 20 s (a dead peer must not wedge the GPU; the result is then garbage and the
// communicator's own failure detection reports the rank).
__device__ __forceinline__ void wait_flag(const uint32_t* f, uint32_t target) {
  if (ld_acquire_sys_u32(f) >= target) return;
  unsigned long long t0;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
  uint32_t spins = 0;
  while (ld_acquire_sys_u32(f) < target) {
    if ((++spins & 1023u) == 0u) {
      unsigned long long t1;
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
      if (t1 - t0 > 20ull * 1000ull * 1000ull * 1000ull) return;
    }
  }
}

struct PushMaps {
  CUtensorMap m[8];
};

__device__ __forceinline__ void red_add_v4(float* p, float a, float b, float c, float d) {
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(a), "f"(b), "f"(c),
               "f"(d)
               : "memory");
}

__device__ __forceinline__ float apply_act(float v, int act) {
  if (act == 1) return tanh_f32(v);
  if (act == 2) return fmaxf(v, 0.0f);
  return v;
}

// Store 32 consecutive output columns of one row (registers v[0..31]) into `base`
// ([rows, ld] of bf16 or fp32), honouring N bounds.
__device__ __forceinline__ void store_row32(void* base, int ld, int c_fp32, int row, int col0,
                                            int N, const float (&v)[32], int atomic) {
  const bool full = (col0 + 32 <= N);
  if (c_fp32) {
    float* p = reinterpret_cast<float*>(base) + static_cast<size_t>(row) * ld + col0;
    const bool vec = full && ((reinterpret_cast<uintptr_t>(p) & 15) == 0);
    if (atomic) {
      if (vec) {
#pragma unroll
        for (int j = 0; j < 32; j += 4) red_add_v4(p + j, v[j], v[j + 1], v[j + 2], v[j + 3]);
      } else {
        for (int j = 0; j < 32; ++j)
          if (col0 + j < N) atomicAdd(p + j, v[j]);
      }
    } else {
      if (vec) {
#pragma unroll
        for (int j = 0; j < 32; j += 4)
          *reinterpret_cast<float4*>(p + j) = make_float4(v[j], v[j + 1], v[j + 2], v[j + 3]);
      } else {
        for (int j = 0; j < 32; ++j)
          if (col0 + j < N) p[j] = v[j];
      }
    }
  } else {
    __nv_bfloat16* p = reinterpret_cast<__nv_bfloat16*>(base) + static_cast<size_t>(row) * ld + col0;
    const bool vec = full && ((reinterpret_cast<uintptr_t>(p) & 15) == 0);
    if (vec) {
#pragma unroll
      for (int j = 0; j < 32; j += 8) {
        uint4 u;
        u.x = pack_bf16x2(v[j], v[j + 1]);
        u.y = pack_bf16x2(v[j + 2], v[j + 3]);
        u.z = pack_bf16x2(v[j + 4], v[j + 5]);
        u.w = pack_bf16x2(v[j + 6], v[j + 7]);
        *reinterpret_cast<uint4*>(p + j) = u;
      }
    } else {
      for (int j = 0; j < 32; ++j)
        if (col0 + j < N) p[j] = __float2bfloat16_rn(v[j]);
    }
  }
}

// Tile rasterisation: m-blocks are walked in groups of `group_m` (m fastest inside a group, then n),
// so that the ~148 tiles in flight form a roughly square patch of the output and re-use each
// other's A and B panels out of L2 (a plain m-fastest sweep streams all of A once per ~2 n-blocks).
__device__ __forceinline__ void decode_tile(int tile, int num_m, int num_n, int group_m, int n_rotate, int& m_blk, int& n_blk) {
  const int per_group = group_m * num_n;
  const int g = tile / per_group;
  const int first_m = g * group_m;
  const int gsz = min(num_m - first_m, group_m);
  const int in_g = tile - g * per_group;
  m_blk = first_m + in_g % gsz;
  n_blk = (in_g / gsz + n_rotate) % num_n;
}

template <bool A_MN, bool B_MN, int BN, bool AUX>
__global__ void __launch_bounds__(kThreads, 1)
gemm_bf16_tcgen05_kernel(const __grid_constant__ CUtensorMap tma_a,
                         const __grid_constant__ CUtensorMap tma_b,
                         const __grid_constant__ CUtensorMap tma_c,
                         const __grid_constant__ PushMaps tma_push, const DevArgs p) {
  using C = Cfg<BN>;
  extern __shared__ uint8_t smem_raw[];
  // SWIZZLE_128B tiles need 1024-byte alignment.
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~static_cast<uintptr_t>(1023));
  __shared__ __align__(8) uint64_t full_bar[C::kStages];
  __shared__ __align__(8) uint64_t empty_bar[C::kStages];
  __shared__ __align__(8) uint64_t tmem_full_bar[2];
  __shared__ __align__(8) uint64_t tmem_empty_bar[2];
  __shared__ uint32_t tmem_base_smem;
  __shared__ int s_ready;   // fused all-gather: work items whose A blocks the watcher warp has seen arrive
  __shared__ __align__(16) float s_bias[4][64];

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  if (threadIdx.x == 0) s_ready = 0;

  const int num_m = (p.M + BM - 1) / BM;
  const int num_n = (p.N + BN - 1) / BN;
  const int num_kb = (p.K + BK - 1) / BK;
  const int splits = p.split_k > 1 ? p.split_k : 1;
  const int kb_per_split = (num_kb + splits - 1) / splits;
  const int num_work = num_m * num_n * splits;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tma_a);
    tma_prefetch_desc(&tma_b);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < C::kStages; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(&tmem_full_bar[s], 1);
      mbar_init(&tmem_empty_bar[s], kEpiThreads);
    }
    fence_mbar_init();
  }
  if (warp == 2) tmem_alloc<C::kTmemCols>(&tmem_base_smem);
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  const uint32_t tmem_base = tmem_base_smem;
  pdl_wait();                 // everything above touched only this CTA's shared / tensor memory
  pdl_launch_dependents();

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      int item = 0;
      for (int w = blockIdx.x; w < num_work; w += gridDim.x, ++item) {
        const int split = w % splits;
        const int tile = w / splits;
        int m_blk, n_blk;
        decode_tile(tile, num_m, num_n, p.group_m, p.n_rotate, m_blk, n_blk);
        const int kb0 = split * kb_per_split;
        const int kb1 = min(num_kb, kb0 + kb_per_split);
        if (p.wait_flags != nullptr) {
          // A is a gathered matrix: the watcher warp has seen every block this work item reads arrive
          while (*reinterpret_cast<volatile int*>(&s_ready) <= item) {
          }
          asm volatile("fence.proxy.async;" ::: "memory");  // the TMA (async proxy) reads below must see those writes
        }
        const uint32_t km = p.use_kmask ? p.kmask[n_blk] : 0xffffffffu;
        for (int kb = kb0; kb < kb1; ++kb) {
          if (!((km >> (kb & 31)) & 1u)) continue;
          mbar_wait(&empty_bar[stage], phase ^ 1);
          uint8_t* sa = smem + stage * C::kStageBytes;
          uint8_t* sb = sa + kABytes;
          mbar_arrive_expect_tx(&full_bar[stage], C::kStageBytes);
          if (!A_MN) {
            tma_load_2d(sa, &tma_a, &full_bar[stage], kb * BK, m_blk * BM);
          } else {
#pragma unroll
            for (int j = 0; j < BM / 64; ++j)
              tma_load_2d(sa + j * (BK * 128), &tma_a, &full_bar[stage], m_blk * BM + j * 64,
                          kb * BK);
          }
          if (!B_MN) {
            tma_load_2d(sb, &tma_b, &full_bar[stage], kb * BK, n_blk * BN);
          } else {
#pragma unroll
            for (int j = 0; j < BN / 64; ++j)
              tma_load_2d(sb + j * (BK * 128), &tma_b, &full_bar[stage], n_blk * BN + j * 64,
                          kb * BK);
          }
          if (++stage == C::kStages) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    if (elect_one()) {
      constexpr uint32_t idesc = make_idesc_bf16(BM, BN, A_MN ? 1u : 0u, B_MN ? 1u : 0u);
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      for (int w = blockIdx.x; w < num_work; w += gridDim.x) {
        const int split = w % splits;
        const int kb0 = split * kb_per_split;
        const int kb1 = min(num_kb, kb0 + kb_per_split);
        uint32_t km = 0xffffffffu;
        if (p.use_kmask) {
          int m_blk, n_blk;
          decode_tile(w / splits, num_m, num_n, p.group_m, p.n_rotate, m_blk, n_blk);
          km = p.kmask[n_blk];
        }
        bool first = true;
        mbar_wait(&tmem_empty_bar[acc], acc_phase ^ 1);
        tc_fence_after_sync();
        const uint32_t tmem_d = tmem_base + static_cast<uint32_t>(acc * BN);
        for (int kb = kb0; kb < kb1; ++kb) {
          if (!((km >> (kb & 31)) & 1u)) continue;
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after_sync();
          const uint32_t a_addr = smem_u32(smem + stage * C::kStageBytes);
          const uint32_t b_addr = a_addr + kABytes;
#pragma unroll
          for (int k = 0; k < BK / 16; ++k) {
            const uint64_t da = A_MN ? make_smem_desc_sw128(a_addr + k * 2048, BK * 128, 1024)
                                     : make_smem_desc_sw128(a_addr + k * 32, 16, 1024);
            const uint64_t db = B_MN ? make_smem_desc_sw128(b_addr + k * 2048, BK * 128, 1024)
                                     : make_smem_desc_sw128(b_addr + k * 32, 16, 1024);
            umma_f16(tmem_d, da, db, idesc, (!first || k > 0) ? 1u : 0u);
          }
          first = false;
          umma_commit(&empty_bar[stage]);  // frees the smem slot once these MMAs retire
          if (++stage == C::kStages) {
            stage = 0;
            phase ^= 1;
          }
        }
        umma_commit(&tmem_full_bar[acc]);  // accumulator ready for the epilogue
        if (++acc == 2) {
          acc = 0;
          acc_phase ^= 1;
        }
      }
    }
    __syncwarp();
  } else if (warp == 3 && p.wait_flags != nullptr) {
    // ===================== flag watcher (fused all-gather -> GEMM) =====================
    // One lane per (32-row block, source rank) pair that the next work item reads: all pairs are polled in
    // parallel (a single thread doing it serially would cost an L2 round trip per pair), then the TMA producer
    // is released through shared memory.
    int item = 0;
    for (int w = blockIdx.x; w < num_work; w += gridDim.x, ++item) {
      const int split = w % splits;
      const int tile = w / splits;
      int m_blk, n_blk;
      decode_tile(tile, num_m, num_n, p.group_m, p.n_rotate, m_blk, n_blk);
      const int kb0 = split * kb_per_split;
      const int kb1 = min(num_kb, kb0 + kb_per_split);
      if (kb1 > kb0) {
        if (!A_MN) {
          // A [M, K] K-major: rows of this m-block, one source per wait_cols_per_src columns of K
          const int src0 = (kb0 * BK) / p.wait_cols_per_src;
          const int src1 = (min(p.K, kb1 * BK) - 1) / p.wait_cols_per_src;
          const int rb_end = (p.M + 31) / 32;
          const int count = (BM / 32) * (src1 - src0 + 1);
          for (int i = lane; i < count; i += 32) {
            const int rb = m_blk * (BM / 32) + (i & 3), src = src0 + (i >> 2);
            if (rb < rb_end) wait_flag(p.wait_flags + static_cast<size_t>(rb) * p.flag_stride + src, p.wait_target);
          }
        } else {
          // A [K, M] MN-major: 32-row blocks of the k-range, sources of this m-block's two 64-column chunks
          const int rb0 = kb0 * (BK / 32);
          const int rb1 = min((p.K + 31) / 32, kb1 * (BK / 32));
          const int count = (rb1 - rb0) * (BM / 64);
          for (int i = lane; i < count; i += 32) {
            const int rb = rb0 + i / (BM / 64), col = m_blk * BM + (i % (BM / 64)) * 64;
            if (col < p.M) wait_flag(p.wait_flags + static_cast<size_t>(rb) * p.flag_stride + col / p.wait_cols_per_src, p.wait_target);
          }
        }
      }
      __syncwarp();
      if (lane == 0) {
        __threadfence_block();
        *reinterpret_cast<volatile int*>(&s_ready) = item + 1;
      }
    }
  } else if (warp >= 4) {
    // ===================== epilogue =====================
    const int wg = warp & 3;  // TMEM lane quarter this warp may touch
    int acc = 0;
    uint32_t acc_phase = 0;
    if (p.tma_store) {
      // bf16 output: TMEM -> registers -> (bias, act, cast) -> SWIZZLE_128B staging in shared
      // memory -> one bulk tensor store per 32-row x 64-column slab (full 128-byte lines, OOB
      // clipped by the TMA unit; the same slab is also stored to every push map = fused all-gather).
      uint8_t* stg = smem + C::kStages * C::kStageBytes + wg * 8192;
      float* sb = s_bias[wg];
      int buf = 0;
      for (int w = blockIdx.x; w < num_work; w += gridDim.x) {
        const int tile = w / splits;
        int m_blk, n_blk;
        decode_tile(tile, num_m, num_n, p.group_m, p.n_rotate, m_blk, n_blk);
        mbar_wait(&tmem_full_bar[acc], acc_phase);
        tc_fence_after_sync();
        const int row0 = m_blk * BM + wg * 32;
        const uint32_t taddr = tmem_base + (static_cast<uint32_t>(wg * 32) << 16) +
                               static_cast<uint32_t>(acc * BN);
        int slabs_done = 0;
        const bool plain = p.alpha == 1.0f && p.bias == nullptr && p.act == 0;
#pragma unroll 1
        for (int c = 0; c < BN / 64; ++c) {
          const int col0 = n_blk * BN + c * 64;
          if (col0 >= p.N) break;  // warp-uniform
          if (p.bias != nullptr) {
            sb[lane] = (col0 + lane < p.N) ? __ldg(p.bias + col0 + lane) : 0.f;
            sb[32 + lane] = (col0 + 32 + lane < p.N) ? __ldg(p.bias + col0 + 32 + lane) : 0.f;
          }
          uint32_t r0[32], r1[32];
          tmem_ld_32x32b_x32(taddr + static_cast<uint32_t>(c * 64), r0);
          tmem_ld_32x32b_x32(taddr + static_cast<uint32_t>(c * 64 + 32), r1);
          if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");  // staging[buf] free
          tmem_ld_wait();
          __syncwarp();
          uint8_t* dst = stg + buf * 4096 + lane * 128;
          const int row = row0 + lane;

          const bool aux_ok = AUX && row < p.M;
          const __nv_bfloat16* aux_row = p.aux + static_cast<size_t>(row < p.M ? row : 0) * p.ld_aux + col0;
          if (!AUX && plain) {
            // nothing to apply: 32 packs + 8 stores per row instead of ~400 instructions (the four epilogue warps, one per
            // scheduler, were the critical path of short-K problems: 6 K cycles per 128x256 tile against 3-5 K of MMA)
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              uint4 u;
              const uint32_t* rr = j < 4 ? &r0[j * 8] : &r1[(j - 4) * 8];
              u.x = pack_bf16x2(__uint_as_float(rr[0]), __uint_as_float(rr[1]));
              u.y = pack_bf16x2(__uint_as_float(rr[2]), __uint_as_float(rr[3]));
              u.z = pack_bf16x2(__uint_as_float(rr[4]), __uint_as_float(rr[5]));
              u.w = pack_bf16x2(__uint_as_float(rr[6]), __uint_as_float(rr[7]));
              *reinterpret_cast<uint4*>(dst + ((j ^ (lane & 7)) << 4)) = u;
            }
          } else
#pragma unroll
          for (int j = 0; j < 8; ++j) {  // 8 x 16-byte chunks = 64 bf16 columns of this row
            float v[8];
            float ay[8];
            if (AUX && aux_ok) {
              if (col0 + j * 8 + 8 <= p.N && ((reinterpret_cast<uintptr_t>(aux_row + j * 8) & 15) == 0)) {
                const uint4 q = *reinterpret_cast<const uint4*>(aux_row + j * 8);
                float2 f;
                f = unpack_bf16x2(q.x); ay[0] = f.x; ay[1] = f.y;
                f = unpack_bf16x2(q.y); ay[2] = f.x; ay[3] = f.y;
                f = unpack_bf16x2(q.z); ay[4] = f.x; ay[5] = f.y;
                f = unpack_bf16x2(q.w); ay[6] = f.x; ay[7] = f.y;
              } else {
#pragma unroll
                for (int e = 0; e < 8; ++e) ay[e] = (col0 + j * 8 + e < p.N) ? __bfloat162float(aux_row[j * 8 + e]) : 0.f;
              }
            }
#pragma unroll
            for (int e = 0; e < 8; ++e) {
              const int col = j * 8 + e;
              float x = __uint_as_float(col < 32 ? r0[col] : r1[col - 32]) * p.alpha;
              if (p.bias != nullptr) x += sb[col];
              x = apply_act(x, p.act);
              if (AUX && aux_ok) x *= (1.0f - ay[e] * ay[e]);
              v[e] = x;
            }
            uint4 u;
            u.x = pack_bf16x2(v[0], v[1]);
            u.y = pack_bf16x2(v[2], v[3]);
            u.z = pack_bf16x2(v[4], v[5]);
            u.w = pack_bf16x2(v[6], v[7]);
            *reinterpret_cast<uint4*>(dst + ((j ^ (lane & 7)) << 4)) = u;
          }
          fence_proxy_async_smem();
          __syncwarp();
          if (lane == 0 && row0 < p.M) {
            if (p.rs_width > 0 && p.rs_slab) {
              // slab-major reduce-scatter push: the staged 32x64 slab (4 KB, still swizzled) goes as ONE
              // contiguous bulk copy into our slot of the owner's staging buffer: 4 KB NVLink writes
              // instead of 32 separate 128-byte rows.  reduce_slots undoes the swizzle.
              const int q = col0 / p.rs_width;
              const size_t slab = static_cast<size_t>(row0 >> 5) * (p.rs_width >> 6) + ((col0 - q * p.rs_width) >> 6);
              uint8_t* dst = static_cast<uint8_t*>(p.push[q]) + slab * 4096;
              asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], 4096;" ::"l"(dst), "r"(smem_u32(stg + buf * 4096)) : "memory");
            } else if (p.rs_width > 0) {
              // reduce-scatter push: this 64-column slab belongs to rank q; store it into OUR slot
              // of q's staging buffer (peer memory over NVLink), nowhere else.
              const int q = col0 / p.rs_width;
              tma_store_2d(&tma_push.m[q], stg + buf * 4096, col0 - q * p.rs_width, row0);
            } else {
              tma_store_2d(&tma_c, stg + buf * 4096, col0, row0);
              for (int q = 0; q < p.n_push; ++q)
                tma_store_2d(&tma_push.m[q], stg + buf * 4096, p.push_col_offset + col0, row0);
            }
            asm volatile("cp.async.bulk.commit_group;" ::: "memory");
          }
          buf ^= 1;
          ++slabs_done;
        }
        tc_fence_before_sync();
        mbar_arrive(&tmem_empty_bar[acc]);
        if (p.n_sig > 0 && row0 < p.M) {
          // fused all-gather, producer side: every slab of this 32-row block has been handed to the TMA unit;
          // once the stores have completed, tell each target how many slabs of ours it now holds
          if (lane == 0) {
            asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
            __threadfence_system();
            for (int q = 0; q < p.n_sig; ++q)
              red_release_sys_add_u32(p.sig[q] + static_cast<size_t>(row0 >> 5) * p.flag_stride + p.sig_src, static_cast<uint32_t>(slabs_done));
          }
          __syncwarp();
        }
        if (++acc == 2) {
          acc = 0;
          acc_phase ^= 1;
        }
      }
      if (lane == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");  // stores landed before exit
      __syncwarp();
    } else {
    for (int w = blockIdx.x; w < num_work; w += gridDim.x) {
      const int tile = w / splits;
      int m_blk, n_blk;
      decode_tile(tile, num_m, num_n, p.group_m, p.n_rotate, m_blk, n_blk);
      mbar_wait(&tmem_full_bar[acc], acc_phase);
      tc_fence_after_sync();
      const int row = m_blk * BM + wg * 32 + lane;
      const uint32_t taddr = tmem_base + (static_cast<uint32_t>(wg * 32) << 16) +
                             static_cast<uint32_t>(acc * BN);
#pragma unroll 1
      for (int c = 0; c < BN / 32; ++c) {
        const int col0 = n_blk * BN + c * 32;
        if (col0 >= p.N) break;  // warp-uniform
        uint32_t r[32];
        tmem_ld_32x32b_x32(taddr + static_cast<uint32_t>(c * 32), r);
        tmem_ld_wait();
        float v[32];
#pragma unroll
        for (int j = 0; j < 32; ++j) {
          float x = __uint_as_float(r[j]) * p.alpha;
          if (p.bias != nullptr && col0 + j < p.N) x += __ldg(p.bias + col0 + j);
          v[j] = apply_act(x, p.act);
        }
        if (row < p.M) {
          store_row32(p.C, p.ldc, p.c_fp32, row, col0, p.N, v, p.atomic);
          for (int q = 0; q < p.n_push; ++q)
            store_row32(p.push[q], p.push_ldc, p.c_fp32, row, p.push_col_offset + col0,
                        p.push_col_offset + p.N, v, 0);
        }
      }
      tc_fence_before_sync();
      mbar_arrive(&tmem_empty_bar[acc]);
      if (++acc == 2) {
        acc = 0;
        acc_phase ^= 1;
      }
    }
    }
  }

  tc_fence_before_sync();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after_sync();
    tmem_dealloc<C::kTmemCols>(tmem_base);
  }
}

// ---- host side: tensor maps -------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                  const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn get_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) ==
            cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  });
  return fn;
}

// 2-D bf16 tensor [rows, cols] (cols contiguous, pitch ld elements); box = 64 cols x box_rows.
int make_tmap(CUtensorMap* out, const void* ptr, long long rows, long long cols, long long ld,
              int box_rows) {
  memset(out, 0, sizeof(*out));
  EncodeTiledFn fn = get_encode_fn();
  TADNN_REQUIRE(fn != nullptr, "cuTensorMapEncodeTiled not available from the driver");
  TADNN_REQUIRE((reinterpret_cast<uintptr_t>(ptr) & 15) == 0, "TMA operand %p is not 16-byte aligned", ptr);
  TADNN_REQUIRE(ld % 8 == 0, "TMA operand pitch %lld is not a multiple of 8 bf16 elements", ld);
  cuuint64_t gdim[2] = {static_cast<cuuint64_t>(cols), static_cast<cuuint64_t>(rows)};
  cuuint64_t gstr[1] = {static_cast<cuuint64_t>(ld) * 2};
  cuuint32_t box[2] = {64u, static_cast<cuuint32_t>(box_rows)};
  cuuint32_t estr[2] = {1u, 1u};
  CUresult r = fn(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(ptr), gdim, gstr, box,
                  estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                  CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  TADNN_REQUIRE(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled failed with CUresult %d (rows=%lld cols=%lld ld=%lld box_rows=%d)",
                static_cast<int>(r), rows, cols, ld, box_rows);
  return 0;
}

template <bool A_MN, bool B_MN, int BN, bool AUX = false>
int launch(const tadnn_gemm_args* a, const DevArgs& d, cudaStream_t stream) {
  using C = Cfg<BN>;
  CUtensorMap ta, tb, tc;
  PushMaps tp;
  memset(&tc, 0, sizeof(tc));
  memset(&tp, 0, sizeof(tp));
  int rc;
  if (d.tma_store) {
    // C / push targets: [M, ld] bf16, stored in 32-row x 64-column slabs
    if (d.rs_width > 0) {
      for (int q = 0; q < d.n_push; ++q) {
        rc = make_tmap(&tp.m[q], d.push[q], a->M, d.rs_width, d.push_ldc, 32);
        if (rc) return rc;
      }
    } else {
      rc = make_tmap(&tc, d.C, a->M, a->N, d.ldc, 32);
      if (rc) return rc;
      for (int q = 0; q < d.n_push; ++q) {
        rc = make_tmap(&tp.m[q], d.push[q], a->M, static_cast<long long>(d.push_col_offset) + a->N, d.push_ldc, 32);
        if (rc) return rc;
      }
    }
  }
  // K-major operand: tensor [MN, K], box 64(K) x tile rows.  MN-major: tensor [K, MN], box 64(MN) x BK.
  rc = A_MN ? make_tmap(&ta, a->A, a->K, a->M, a->lda, BK) : make_tmap(&ta, a->A, a->M, a->K, a->lda, BM);
  if (rc) return rc;
  rc = B_MN ? make_tmap(&tb, a->B, a->K, a->N, a->ldb, BK) : make_tmap(&tb, a->B, a->N, a->K, a->ldb, BN);
  if (rc) return rc;
  auto kern = gemm_bf16_tcgen05_kernel<A_MN, B_MN, BN, AUX>;
  static std::atomic<unsigned long long> attr_set{0};  // per instantiation, one bit per device
  if (int arc = ensure_dyn_smem(kern, C::kSmemBytes, attr_set)) return arc;
  const int num_m = ceil_div(a->M, BM), num_n = ceil_div(a->N, BN);
  const int splits = d.split_k > 1 ? d.split_k : 1;
  const int work = num_m * num_n * splits;
  const int grid = work < sm_count() ? work : sm_count();
  // plain GEMMs may overlap the tail of the kernel before them; the ones that talk to peers keep the strict stream order
  if (d.n_push == 0 && d.n_sig == 0 && d.wait_flags == nullptr) {
    TADNN_CHECK_CUDA(launch_pdl(kern, dim3(grid), dim3(kThreads), C::kSmemBytes, stream, ta, tb, tc, tp, d));
  } else {
    kern<<<grid, kThreads, C::kSmemBytes, stream>>>(ta, tb, tc, tp, d);
  }
  TADNN_LAUNCH_CHECK("gemm_bf16_tcgen05");
  return 0;
}

template <int BN>
int dispatch_major(const tadnn_gemm_args* a, const DevArgs& d, cudaStream_t s) {
  if (!a->a_mn_major && !a->b_mn_major) return launch<false, false, BN>(a, d, s);
  if (!a->a_mn_major && a->b_mn_major) return d.aux != nullptr ? launch<false, true, BN, true>(a, d, s) : launch<false, true, BN>(a, d, s);
  if (a->a_mn_major && a->b_mn_major) return launch<true, true, BN>(a, d, s);
  return launch<true, false, BN>(a, d, s);
}

}  // namespace
}  // namespace tadnn

extern "C" int tadnn_gemm_bf16(const tadnn_gemm_args* a, tadnn_stream_t stream_) {
  using namespace tadnn;
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  TADNN_REQUIRE(a != nullptr, "tadnn_gemm_bf16: null args");
  TADNN_REQUIRE(a->M > 0 && a->N > 0 && a->K > 0, "tadnn_gemm_bf16: bad shape M=%d N=%d K=%d", a->M, a->N, a->K);
  TADNN_REQUIRE(a->n_push >= 0 && a->n_push <= 8, "tadnn_gemm_bf16: n_push=%d out of range", a->n_push);
  // Large plain problems (K-major operands, bf16 output, no collective / split-K epilogue, tile width not
  // forced) go to the 2-CTA kernel: cta_group::2 pairs, 256x256 tiles, 1/3 less L2->shared traffic.
  if (!a->a_mn_major && !a->b_mn_major && !a->c_fp32 && !a->accumulate && a->split_k <= 1 && a->n_push == 0 &&
      a->rs_width == 0 && a->aux_tanh_out == nullptr && a->n_sig == 0 && a->wait_flags == nullptr && a->tile_n == 0 && a->M >= 512 && a->N >= 256 && a->K >= 256 &&
      a->lda % 8 == 0 && a->ldb % 8 == 0 && a->ldc % 8 == 0 &&
      ((reinterpret_cast<uintptr_t>(a->A) | reinterpret_cast<uintptr_t>(a->B) | reinterpret_cast<uintptr_t>(a->C)) & 15) == 0)
    return tadnn_gemm_bf16_2cta(a->A, a->lda, a->B, a->ldb, a->C, a->ldc, a->M, a->N, a->K, a->bias, a->act, a->alpha, stream_);
  const int num_kb = ceil_div(a->K, BK);
  int split = a->split_k > 1 ? a->split_k : 1;
  if (split > num_kb) split = num_kb;
  // make sure no split is empty
  while (split > 1 && (split - 1) * ceil_div(num_kb, split) >= num_kb) --split;
  const int atomic = (a->accumulate || split > 1) ? 1 : 0;
  TADNN_REQUIRE(!atomic || a->c_fp32, "tadnn_gemm_bf16: accumulate/split_k need an fp32 C");
  TADNN_REQUIRE(split == 1 || (a->bias == nullptr && a->act == 0 && a->n_push == 0),
                "tadnn_gemm_bf16: split_k cannot be combined with bias/act/push epilogues");
  if (split > 1 && !a->accumulate) {
    // split-K partials are summed with red.global.add: start from zero.
    if (a->ldc == a->N) {
      TADNN_CHECK_CUDA(cudaMemsetAsync(a->C, 0, sizeof(float) * static_cast<size_t>(a->M) * a->N, stream));
    } else {
      TADNN_CHECK_CUDA(cudaMemset2DAsync(a->C, sizeof(float) * a->ldc, 0, sizeof(float) * a->N, a->M, stream));
    }
  }
  DevArgs d;
  memset(&d, 0, sizeof(d));
  d.M = a->M; d.N = a->N; d.K = a->K;
  d.C = a->C; d.ldc = a->ldc; d.c_fp32 = a->c_fp32;
  d.bias = a->bias; d.act = a->act; d.alpha = a->alpha;
  d.atomic = atomic; d.split_k = split;
  d.n_push = a->n_push;
  for (int i = 0; i < a->n_push; ++i) d.push[i] = a->push[i];
  d.push_ldc = a->push_ldc; d.push_col_offset = a->push_col_offset;
  d.rs_width = a->rs_width;
  d.rs_slab = a->rs_slab_major ? 1 : 0;
  d.n_rotate = a->n_rotate > 0 ? a->n_rotate : 0;
  {
    // group height so that group_m x (148 / group_m) tiles in flight are about square in bytes
    // (an n-block is BN=256 wide, an m-block 128 tall): 16 x 9 for big problems
    const int num_m = ceil_div(a->M, BM);
    d.group_m = a->rs_width > 0 ? num_m : (num_m < 16 ? num_m : 16);
  }
  d.aux = static_cast<const __nv_bfloat16*>(a->aux_tanh_out);
  d.ld_aux = a->ld_aux;
  TADNN_REQUIRE(a->n_sig >= 0 && a->n_sig <= 8, "tadnn_gemm_bf16: n_sig=%d out of range", a->n_sig);
  d.n_sig = a->n_sig;
  for (int i = 0; i < a->n_sig; ++i) d.sig[i] = static_cast<uint32_t*>(a->sig_flags[i]);
  d.sig_src = a->sig_src;
  d.flag_stride = a->flag_stride;
  d.wait_flags = static_cast<const uint32_t*>(a->wait_flags);
  d.wait_target = a->wait_target;
  d.wait_cols_per_src = a->wait_cols_per_src;
  d.use_kmask = 0;
  if (a->kblock_mask != nullptr) {
    const int tn = a->tile_n;
    TADNN_REQUIRE(tn == 64 || tn == 128 || tn == 256, "tadnn_gemm_bf16: kblock_mask needs an explicit tile_n");
    const int nn = (a->N + tn - 1) / tn, nk = (a->K + 63) / 64;
    TADNN_REQUIRE(nn <= 32 && nk <= 32 && a->split_k <= 1 && a->n_kblock_mask == nn,
                  "tadnn_gemm_bf16: kblock_mask supports <= 32 column tiles and <= 32 k-blocks, no split-K, one word per column tile (%d given, %d tiles)",
                  a->n_kblock_mask, nn);
    for (int i = 0; i < nn; ++i) {
      TADNN_REQUIRE((a->kblock_mask[i] & (nk >= 32 ? 0xffffffffu : ((1u << nk) - 1u))) != 0, "tadnn_gemm_bf16: kblock_mask[%d] selects no k-block", i);
      d.kmask[i] = a->kblock_mask[i];
    }
    d.use_kmask = 1;
  }
  TADNN_REQUIRE(d.wait_flags == nullptr || (d.wait_cols_per_src > 0 && d.wait_cols_per_src % 64 == 0 && d.flag_stride > 0),
                "tadnn_gemm_bf16: wait_flags needs wait_cols_per_src (a multiple of 64) and flag_stride");
  if (a->rs_width > 0) {
    TADNN_REQUIRE(!a->c_fp32 && !atomic, "tadnn_gemm_bf16: rs_width needs a bf16, non-accumulating output");
    TADNN_REQUIRE(a->rs_width % 64 == 0 && a->n_push * a->rs_width >= a->N && a->push_ldc % 8 == 0,
                  "tadnn_gemm_bf16: rs_width=%d must be a multiple of 64 with n_push*rs_width >= N and push_ldc %% 8 == 0", a->rs_width);
    for (int i = 0; i < a->n_push; ++i)
      TADNN_REQUIRE((reinterpret_cast<uintptr_t>(a->push[i]) & 15) == 0, "tadnn_gemm_bf16: rs push target %d misaligned", i);
    d.tma_store = 1;
  } else {
    bool ok = !a->c_fp32 && !atomic && (a->ldc % 8 == 0) && ((reinterpret_cast<uintptr_t>(a->C) & 15) == 0);
    for (int i = 0; i < a->n_push && ok; ++i)
      ok = (a->push_ldc % 8 == 0) && ((reinterpret_cast<uintptr_t>(a->push[i]) & 15) == 0) && (a->push_col_offset % 8 == 0);
    d.tma_store = ok ? 1 : 0;
  }
  TADNN_REQUIRE(d.n_sig == 0 || (d.tma_store && a->rs_width == 0 && d.flag_stride > 0), "tadnn_gemm_bf16: sig_flags need the bf16 TMA-store epilogue (all-gather push form)");
  TADNN_REQUIRE(d.aux == nullptr || (d.tma_store && !a->a_mn_major && a->b_mn_major), "tadnn_gemm_bf16: aux_tanh_out needs the bf16 TMA-store epilogue and the dgrad operand layout (a K-major, b MN-major)");

  int bn = a->tile_n;
  if (bn == 0) bn = a->N <= 64 ? 64 : (a->N <= 128 ? 128 : 256);
  switch (bn) {
    case 64: return dispatch_major<64>(a, d, stream);
    case 128: return dispatch_major<128>(a, d, stream);
    case 256: return dispatch_major<256>(a, d, stream);
    default: TADNN_REQUIRE(false, "tadnn_gemm_bf16: tile_n must be 0, 64, 128 or 256 (got %d)", bn);
  }
  return 0;
}
