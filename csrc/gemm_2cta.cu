// 2-CTA ("cta_group::2") variant of the bf16 GEMM for large K-major problems: a thread-block cluster of
// two CTAs on one TPC computes a 256 x 256 output tile with ONE tcgen05.mma stream issued by the leader
// CTA.  Each CTA stages its own 128 rows of A and HALF of the B tile (128 of the 256 N-rows), so a
// k-block costs 32 KB of L2->shared traffic per SM instead of 48 KB and the ring holds 6 stages.
//
//   CTA r of the pair:  A rows  m0 + 128 r .. +128      B rows  n0 + 128 r .. +128
//   full barrier   : lives in the leader; both CTAs' TMA loads complete_tx on it (2 producers)
//   empty / tmem_full barriers : in both CTAs; the leader's tcgen05.commit multicasts to both
//   tmem_empty barrier : in the leader; the epilogue warps of both CTAs arrive on it
//   accumulator    : TMEM of each CTA holds its own 128 rows x 256 columns (x2 stages)
//
// Same role as gemm_tcgen05.cu (replaces the cuBLAS SGEMM under nn.Linear, nodemodule.lua:47,95).
#include "common.cuh"
#include "sm100.cuh"

namespace tadnn {
namespace {

using namespace sm100;

constexpr int BM = 128;          // rows per CTA
constexpr int BN = 256;          // tile width (both CTAs together provide the 256 B rows)
constexpr int BNH = 128;         // B rows staged per CTA
constexpr int BK = 64;
constexpr int kThreads = 256;
constexpr int kStages = 6;
constexpr int kABytes = BM * BK * 2, kBBytes = BNH * BK * 2, kStageBytes = kABytes + kBBytes;  // 32 KB
constexpr int kStagingBytes = 4 * 2 * 4096;
constexpr int kSmemBytes = kStages * kStageBytes + kStagingBytes + 1024;
constexpr uint32_t kTmemCols = 512;

struct Args2 {
  int M, N, K;
  const float* bias;
  int act;
  float alpha;
  int group_m;
};

__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// shared::cluster address of the same shared-memory location in CTA `rank` of the cluster
__device__ __forceinline__ uint32_t mapa(uint32_t local_addr, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(local_addr), "r"(rank));
  return r;
}
__device__ __forceinline__ void mbar_arrive_expect_tx_cluster(uint32_t cluster_addr, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cluster.b64 _, [%0], %1;" ::"r"(cluster_addr), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t cluster_addr) {
  asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
__device__ __forceinline__ void tma_load_2d_2sm(void* smem_dst, const CUtensorMap* tmap, uint32_t bar_cluster_addr, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(
          smem_u32(smem_dst)),
      "l"(tmap), "r"(bar_cluster_addr), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void umma_f16_2cta(uint32_t tmem_d, uint64_t da, uint64_t db, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "setp.ne.b32 p, %4, 0;\n"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n"
      "}\n" ::"r"(tmem_d),
      "l"(da), "l"(db), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void umma_commit_2cta(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(
                   smem_u32(bar)),
               "h"(static_cast<uint16_t>(3))
               : "memory");
}

__device__ __forceinline__ void decode_tile2(int tile, int num_m, int num_n, int group_m, int& m_blk, int& n_blk) {
  const int per_group = group_m * num_n;
  const int g = tile / per_group;
  const int first_m = g * group_m;
  const int gsz = min(num_m - first_m, group_m);
  const int in_g = tile - g * per_group;
  m_blk = first_m + in_g % gsz;
  n_blk = in_g / gsz;
}

__device__ __forceinline__ float act2(float v, int act) {
  if (act == 1) return tanh_f32(v);
  if (act == 2) return fmaxf(v, 0.0f);
  return v;
}

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kThreads, 1)
gemm_bf16_2cta_kernel(const __grid_constant__ CUtensorMap tma_a, const __grid_constant__ CUtensorMap tma_b,
                      const __grid_constant__ CUtensorMap tma_c, const Args2 p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
  __shared__ __align__(8) uint64_t full_bar[kStages];
  __shared__ __align__(8) uint64_t empty_bar[kStages];
  __shared__ __align__(8) uint64_t tmem_full_bar[2];
  __shared__ __align__(8) uint64_t tmem_empty_bar[2];
  __shared__ uint32_t tmem_base_smem;
  __shared__ __align__(16) float s_bias[4][64];

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t cta = cluster_ctarank();
  const bool leader = cta == 0;
  const int cluster_id = blockIdx.x >> 1, num_clusters = gridDim.x >> 1;
  const int num_m = (p.M + 2 * BM - 1) / (2 * BM);   // 256-row tiles
  const int num_n = (p.N + BN - 1) / BN;
  const int num_kb = (p.K + BK - 1) / BK;
  const int num_tiles = num_m * num_n;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tma_a);
    tma_prefetch_desc(&tma_b);
    tma_prefetch_desc(&tma_c);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < kStages; ++s) {
      mbar_init(&full_bar[s], 2);   // one arrive.expect_tx from each CTA's producer
      mbar_init(&empty_bar[s], 1);  // multicast commit
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(&tmem_full_bar[s], 1);   // multicast commit
      mbar_init(&tmem_empty_bar[s], 8);  // 4 epilogue warps x 2 CTAs (leader's copy is the one waited on)
    }
    fence_mbar_init();
  }
  if (warp == 2) {
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_smem)), "r"(kTmemCols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
  }
  tc_fence_before_sync();
  __syncthreads();
  cluster_sync_all();
  tc_fence_after_sync();
  const uint32_t tmem_base = tmem_base_smem;

  if (warp == 0) {
    // ===================== TMA producer (both CTAs) =====================
    if (elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      for (int t = cluster_id; t < num_tiles; t += num_clusters) {
        int m_blk, n_blk;
        decode_tile2(t, num_m, num_n, p.group_m, m_blk, n_blk);
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(&empty_bar[stage], phase ^ 1);
          uint8_t* sa = smem + stage * kStageBytes;
          uint8_t* sb = sa + kABytes;
          const uint32_t lead_full = mapa(smem_u32(&full_bar[stage]), 0);
          mbar_arrive_expect_tx_cluster(lead_full, kStageBytes);
          tma_load_2d_2sm(sa, &tma_a, lead_full, kb * BK, m_blk * 2 * BM + static_cast<int>(cta) * BM);
          tma_load_2d_2sm(sb, &tma_b, lead_full, kb * BK, n_blk * BN + static_cast<int>(cta) * BNH);
          if (++stage == kStages) { stage = 0; phase ^= 1; }
        }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // ===================== MMA issuer (leader CTA only) =====================
    if (leader && elect_one()) {
      constexpr uint32_t idesc = make_idesc_bf16(2 * BM, BN, 0u, 0u);
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      for (int t = cluster_id; t < num_tiles; t += num_clusters) {
        mbar_wait(&tmem_empty_bar[acc], acc_phase ^ 1);
        tc_fence_after_sync();
        const uint32_t tmem_d = tmem_base + static_cast<uint32_t>(acc * BN);
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after_sync();
          const uint32_t a_addr = smem_u32(smem + stage * kStageBytes);
          const uint32_t b_addr = a_addr + kABytes;
#pragma unroll
          for (int k = 0; k < BK / 16; ++k)
            umma_f16_2cta(tmem_d, make_smem_desc_sw128(a_addr + k * 32, 16, 1024), make_smem_desc_sw128(b_addr + k * 32, 16, 1024),
                          idesc, (kb > 0 || k > 0) ? 1u : 0u);
          umma_commit_2cta(&empty_bar[stage]);
          if (++stage == kStages) { stage = 0; phase ^= 1; }
        }
        umma_commit_2cta(&tmem_full_bar[acc]);
        if (++acc == 2) { acc = 0; acc_phase ^= 1; }
      }
    }
    __syncwarp();
  } else if (warp >= 4) {
    // ===================== epilogue (both CTAs, own 128 rows) =====================
    const int wg = warp & 3;
    uint8_t* stg = smem + kStages * kStageBytes + wg * 8192;
    float* sb = s_bias[wg];
    int buf = 0, acc = 0;
    uint32_t acc_phase = 0;
    for (int t = cluster_id; t < num_tiles; t += num_clusters) {
      int m_blk, n_blk;
      decode_tile2(t, num_m, num_n, p.group_m, m_blk, n_blk);
      mbar_wait(&tmem_full_bar[acc], acc_phase);
      tc_fence_after_sync();
      const int row0 = m_blk * 2 * BM + static_cast<int>(cta) * BM + wg * 32;
      const uint32_t taddr = tmem_base + (static_cast<uint32_t>(wg * 32) << 16) + static_cast<uint32_t>(acc * BN);
#pragma unroll 1
      for (int c = 0; c < BN / 64; ++c) {
        const int col0 = n_blk * BN + c * 64;
        if (col0 >= p.N) break;
        if (p.bias != nullptr) {
          sb[lane] = (col0 + lane < p.N) ? __ldg(p.bias + col0 + lane) : 0.f;
          sb[32 + lane] = (col0 + 32 + lane < p.N) ? __ldg(p.bias + col0 + 32 + lane) : 0.f;
        }
        uint32_t r0[32], r1[32];
        tmem_ld_32x32b_x32(taddr + static_cast<uint32_t>(c * 64), r0);
        tmem_ld_32x32b_x32(taddr + static_cast<uint32_t>(c * 64 + 32), r1);
        if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
        tmem_ld_wait();
        __syncwarp();
        uint8_t* dst = stg + buf * 4096 + lane * 128;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          float v[8];
#pragma unroll
          for (int e = 0; e < 8; ++e) {
            const int col = j * 8 + e;
            float x = __uint_as_float(col < 32 ? r0[col] : r1[col - 32]) * p.alpha;
            if (p.bias != nullptr) x += sb[col];
            v[e] = act2(x, p.act);
          }
          uint4 u;
          u.x = pack_bf16x2(v[0], v[1]); u.y = pack_bf16x2(v[2], v[3]);
          u.z = pack_bf16x2(v[4], v[5]); u.w = pack_bf16x2(v[6], v[7]);
          *reinterpret_cast<uint4*>(dst + ((j ^ (lane & 7)) << 4)) = u;
        }
        fence_proxy_async_smem();
        __syncwarp();
        if (lane == 0 && row0 < p.M) {
          tma_store_2d(&tma_c, stg + buf * 4096, col0, row0);
          asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        }
        buf ^= 1;
      }
      tc_fence_before_sync();
      __syncwarp();
      if (lane == 0) mbar_arrive_cluster(mapa(smem_u32(&tmem_empty_bar[acc]), 0));  // release this accumulator to the leader's MMA
      if (++acc == 2) { acc = 0; acc_phase ^= 1; }
    }
    if (lane == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
    __syncwarp();
  }

  tc_fence_before_sync();
  __syncthreads();
  cluster_sync_all();  // the peer's shared memory / TMEM must stay alive until both CTAs are done
  if (warp == 2) {
    tc_fence_after_sync();
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(kTmemCols) : "memory");
  }
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

int make_tmap2(CUtensorMap* out, const void* ptr, long long rows, long long cols, long long ld, int box_rows) {
  static EncodeTiledFn fn = nullptr;
  if (fn == nullptr) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  }
  TADNN_REQUIRE(fn != nullptr, "cuTensorMapEncodeTiled not available from the driver");
  memset(out, 0, sizeof(*out));
  cuuint64_t gdim[2] = {static_cast<cuuint64_t>(cols), static_cast<cuuint64_t>(rows)};
  cuuint64_t gstr[1] = {static_cast<cuuint64_t>(ld) * 2};
  cuuint32_t box[2] = {64u, static_cast<cuuint32_t>(box_rows)};
  cuuint32_t estr[2] = {1u, 1u};
  CUresult r = fn(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(ptr), gdim, gstr, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                  CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  TADNN_REQUIRE(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled failed with CUresult %d", static_cast<int>(r));
  return 0;
}

}  // namespace
}  // namespace tadnn

using namespace tadnn;

// C[M,N] (bf16) = act(alpha * A[M,K] * B[N,K]^T + bias); A, B K-major bf16.  2-CTA clusters.
extern "C" int tadnn_gemm_bf16_2cta(const void* A, int lda, const void* B, int ldb, void* C, int ldc, int M, int N, int K,
                                    const float* bias, int act, float alpha, tadnn_stream_t stream_) {
  TADNN_REQUIRE(M > 0 && N > 0 && K > 0, "gemm_bf16_2cta: bad shape");
  TADNN_REQUIRE(lda % 8 == 0 && ldb % 8 == 0 && ldc % 8 == 0, "gemm_bf16_2cta: pitches must be multiples of 8 elements");
  TADNN_REQUIRE(((reinterpret_cast<uintptr_t>(A) | reinterpret_cast<uintptr_t>(B) | reinterpret_cast<uintptr_t>(C)) & 15) == 0,
                "gemm_bf16_2cta: operands must be 16-byte aligned");
  CUtensorMap ta, tb, tc;
  if (int rc = make_tmap2(&ta, A, M, K, lda, BM)) return rc;
  if (int rc = make_tmap2(&tb, B, N, K, ldb, BNH)) return rc;
  if (int rc = make_tmap2(&tc, C, M, N, ldc, 32)) return rc;
  static std::atomic<unsigned long long> attr{0};
  if (int rc = ensure_dyn_smem(gemm_bf16_2cta_kernel, kSmemBytes, attr)) return rc;
  Args2 p;
  p.M = M; p.N = N; p.K = K; p.bias = bias; p.act = act; p.alpha = alpha;
  const int num_m = ceil_div(M, 2 * BM), num_n = ceil_div(N, BN);
  p.group_m = num_m < 8 ? num_m : 8;
  const int tiles = num_m * num_n;
  const int pairs = sm_count() / 2;
  const int clusters = tiles < pairs ? tiles : pairs;
  gemm_bf16_2cta_kernel<<<2 * clusters, kThreads, kSmemBytes, static_cast<cudaStream_t>(stream_)>>>(ta, tb, tc, p);
  TADNN_LAUNCH_CHECK("gemm_bf16_2cta");
  return 0;
}
