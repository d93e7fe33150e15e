// tadnn native core — linear layers with a tiny output width ("tiny-N case (o=10) needs a
// split-K/GEMV-style kernel, not a 128x256 tile").
//
// Reference: nodemodule.lua:86-119 (MPBaseLinear as the classifier layer, o = 10 in the README MLP).
// A 128xN tcgen05 tile with N = 10 wastes >90 % of the MMA and, worse, gives only M/128 CTAs to pull
// the activations out of HBM.  These three kernels are bandwidth-shaped instead: x is read exactly
// once with 16-byte loads spread over the whole GPU, W (N x K, a few KB) stays in L1, products are
// accumulated in fp32 registers.  Measured (profiles/smalln_vs_gemm.json): they win while the layer is
// latency-bound (M*K up to a few million elements: 5.1 vs 6.7 us forward, 6.2 vs 16.7 us dgrad at
// 4096x256); beyond that the 10 FMAs per loaded element cost more than the MMA tile's wasted columns,
// so the Python dispatch (ops/smalln.py prefer_*) hands large forward/wgrad shapes back to tcgen05.
//
//   smalln_fwd   : out[m, n] (fp32) = sum_k x[m, k] W[n, k] (+ bias[n])        a warp (narrow K) .. a block (wide K) per R rows
//   smalln_dgrad : dx[m, k] (bf16) = sum_n g[m, n] W[n, k]  (* (1 - a[m,k]^2))  one thread per R rows x 8 k
//   smalln_wgrad : dW[n, k] (fp32) += sum_m g[m, n] x[m, k];  db[n] += sum_m g[m, n]   (atomics per row slab)
#include "common.cuh"

namespace tadnn {
namespace {

constexpr int kRows = 4;  // rows that share one load of a W chunk

__device__ __forceinline__ void unpack8(const uint4& v, float* f) {
  float2 a = unpack_bf16x2(v.x), b = unpack_bf16x2(v.y), c = unpack_bf16x2(v.z), d = unpack_bf16x2(v.w);
  f[0] = a.x; f[1] = a.y; f[2] = b.x; f[3] = b.y; f[4] = c.x; f[5] = c.y; f[6] = d.x; f[7] = d.y;
}

// Warp-wide sum of NT values per lane with NT-1 + log2(32/NT) shuffles instead of 5*NT: every step
// halves the number of live values (the lane keeps one half and receives the partner's copy of it).
// Afterwards every lane holds the total of value index lane / (32 / NT) in v[0].
template <int NT>
__device__ __forceinline__ float warp_sum_transpose(float (&v)[NT], int lane) {
  static_assert(NT == 4 || NT == 8 || NT == 16, "power of two <= 16");
  int off = 16;
#pragma unroll
  for (int half = NT / 2; half >= 1; half >>= 1) {
    const bool up = (lane & off) != 0;
#pragma unroll
    for (int i = 0; i < half; ++i) {
      const float send = up ? v[i] : v[i + half];
      const float keep = up ? v[i + half] : v[i];
      v[i] = keep + __shfl_xor_sync(0xffffffffu, send, off);
    }
    off >>= 1;
  }
  float r = v[0];
  for (; off > 0; off >>= 1) r += __shfl_xor_sync(0xffffffffu, r, off);
  return r;
}

// 256 threads = (256 / tk) row groups x tk threads along K (tk a power of two, 32..256); a row group
// owns kRows rows.  Wide K therefore gets a whole block per row group (all of a row's 16-byte loads
// in flight at once), narrow K a warp.
template <int NT>
__global__ void __launch_bounds__(256) smalln_fwd_kernel(const __nv_bfloat16* __restrict__ x, int ldx, const __nv_bfloat16* __restrict__ W,
                                                         int ldw, const float* __restrict__ bias, float* __restrict__ out, int ldo,
                                                         int M, int N, int K, int tk) {
  __shared__ float red[8][kRows][NT];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int groups = 256 / tk;
  const int kt = threadIdx.x & (tk - 1), rg = threadIdx.x / tk;
  const long long m0 = (static_cast<long long>(blockIdx.x) * groups + rg) * kRows;
  float acc[kRows][NT];
#pragma unroll
  for (int r = 0; r < kRows; ++r)
#pragma unroll
    for (int n = 0; n < NT; ++n) acc[r][n] = 0.f;
  if (m0 < M) {
    for (int k = kt * 8; k < K; k += tk * 8) {
      float xf[kRows][8];
#pragma unroll
      for (int r = 0; r < kRows; ++r) {
        uint4 v = make_uint4(0, 0, 0, 0);
        if (m0 + r < M) v = *reinterpret_cast<const uint4*>(x + (m0 + r) * ldx + k);
        unpack8(v, xf[r]);
      }
#pragma unroll
      for (int n = 0; n < NT; ++n) {
        if (n < N) {
          float wf[8];
          unpack8(__ldg(reinterpret_cast<const uint4*>(W + static_cast<long long>(n) * ldw + k)), wf);
#pragma unroll
          for (int r = 0; r < kRows; ++r)
#pragma unroll
            for (int j = 0; j < 8; ++j) acc[r][n] = fmaf(xf[r][j], wf[j], acc[r][n]);
        }
      }
    }
  }
  constexpr int kShare = 32 / NT;  // lanes that end up with the same value index
#pragma unroll
  for (int r = 0; r < kRows; ++r) {
    const float s = warp_sum_transpose<NT>(acc[r], lane);
    if (lane % kShare == 0) red[warp][r][lane / kShare] = s;
  }
  __syncthreads();
  const int wpg = tk / 32;  // warps per row group
  for (int t = threadIdx.x; t < groups * kRows * NT; t += 256) {
    const int n = t % NT, r = (t / NT) % kRows, g = t / (NT * kRows);
    const long long m = (static_cast<long long>(blockIdx.x) * groups + g) * kRows + r;
    if (n < N && m < M) {
      float s = bias ? bias[n] : 0.f;
      for (int w = 0; w < wpg; ++w) s += red[g * wpg + w][r][n];
      out[m * ldo + n] = s;
    }
  }
}

template <int NT>
__global__ void __launch_bounds__(256) smalln_dgrad_kernel(const __nv_bfloat16* __restrict__ g, int ldg, const __nv_bfloat16* __restrict__ W,
                                                           int ldw, const __nv_bfloat16* __restrict__ aux, int ld_aux,
                                                           __nv_bfloat16* __restrict__ dx, int lddx, int M, int N, int K) {
  const int kc = K >> 3;  // 8-wide chunks per row
  const long long id = blockIdx.x * 256ll + threadIdx.x;
  const long long grp = id / kc;
  const int k = static_cast<int>(id - grp * kc) * 8;
  const long long m0 = grp * kRows;
  if (m0 >= M) return;
  float gf[kRows][NT];
#pragma unroll
  for (int r = 0; r < kRows; ++r)
#pragma unroll
    for (int n = 0; n < NT; ++n) gf[r][n] = (n < N && m0 + r < M) ? bf16_to_f32(g[(m0 + r) * ldg + n]) : 0.f;
  float acc[kRows][8];
#pragma unroll
  for (int r = 0; r < kRows; ++r)
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[r][j] = 0.f;
#pragma unroll
  for (int n = 0; n < NT; ++n) {
    if (n < N) {
      float wf[8];
      unpack8(__ldg(reinterpret_cast<const uint4*>(W + static_cast<long long>(n) * ldw + k)), wf);
#pragma unroll
      for (int r = 0; r < kRows; ++r)
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[r][j] = fmaf(gf[r][n], wf[j], acc[r][j]);
    }
  }
#pragma unroll
  for (int r = 0; r < kRows; ++r) {
    if (m0 + r >= M) break;
    if (aux) {
      float af[8];
      unpack8(*reinterpret_cast<const uint4*>(aux + (m0 + r) * ld_aux + k), af);
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[r][j] *= 1.f - af[j] * af[j];
    }
    uint4 o;
    o.x = pack_bf16x2(acc[r][0], acc[r][1]);
    o.y = pack_bf16x2(acc[r][2], acc[r][3]);
    o.z = pack_bf16x2(acc[r][4], acc[r][5]);
    o.w = pack_bf16x2(acc[r][6], acc[r][7]);
    *reinterpret_cast<uint4*>(dx + (m0 + r) * lddx + k) = o;
  }
}

// grid.x covers 32 k-chunks (256 columns) per block, grid.y the row slabs.  Inside a block the 8 warps
// take interleaved rows of the slab (8 independent 512-byte row segments in flight per block step),
// merge their register tiles through shared-memory atomics, and the block issues ONE coalesced set of
// global reductions.
template <int NT>
__global__ void __launch_bounds__(256) smalln_wgrad_kernel(const __nv_bfloat16* __restrict__ g, int ldg, const __nv_bfloat16* __restrict__ x,
                                                           int ldx, float* __restrict__ dW, int lddw, float* __restrict__ db, int M, int N,
                                                           int K, int rows_per_slab) {
  __shared__ float tile[NT * 8 * 33];
  __shared__ float bs[NT];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int k = (blockIdx.x * 32 + lane) * 8;
  const bool k_ok = k < K;
  const long long m_begin = static_cast<long long>(blockIdx.y) * rows_per_slab;
  long long m_end = m_begin + rows_per_slab;
  if (m_end > M) m_end = M;
  for (int i = threadIdx.x; i < NT * 8 * 33; i += 256) tile[i] = 0.f;
  if (threadIdx.x < NT) bs[threadIdx.x] = 0.f;
  float acc[NT][8];
  float bsum[NT];
#pragma unroll
  for (int n = 0; n < NT; ++n) {
    bsum[n] = 0.f;
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[n][j] = 0.f;
  }
#pragma unroll 2
  for (long long m = m_begin + warp; m < m_end; m += 8) {
    float xf[8];
    uint4 v = make_uint4(0, 0, 0, 0);
    if (k_ok) v = *reinterpret_cast<const uint4*>(x + m * ldx + k);
    unpack8(v, xf);
#pragma unroll
    for (int n = 0; n < NT; ++n) {
      if (n < N) {
        const float gv = bf16_to_f32(g[m * ldg + n]);  // same address across the warp: one broadcast load
        bsum[n] += gv;
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[n][j] = fmaf(gv, xf[j], acc[n][j]);
      }
    }
  }
  __syncthreads();
#pragma unroll
  for (int n = 0; n < NT; ++n) {
    if (n < N) {
#pragma unroll
      for (int j = 0; j < 8; ++j) atomicAdd(&tile[(n * 8 + j) * 33 + lane], acc[n][j]);
      if (lane == 0) atomicAdd(&bs[n], bsum[n]);
    }
  }
  __syncthreads();
  for (int i = threadIdx.x; i < N * 256; i += 256) {
    const int n = i >> 8, c = i & 255;  // c: column inside the block's 256-wide strip
    const int col = blockIdx.x * 256 + c;
    if (col < K) atomicAdd(dW + static_cast<long long>(n) * lddw + col, tile[(n * 8 + (c & 7)) * 33 + (c >> 3)]);
  }
  if (db && blockIdx.x == 0 && threadIdx.x < N) atomicAdd(db + threadIdx.x, bs[threadIdx.x]);
}

}  // namespace
}  // namespace tadnn

using namespace tadnn;

#define SMALLN_DISPATCH(N, CALL)          \
  do {                                    \
    if ((N) <= 4) { constexpr int NT = 4; CALL; }        \
    else if ((N) <= 8) { constexpr int NT = 8; CALL; }   \
    else { constexpr int NT = 16; CALL; }                \
  } while (0)

static int smalln_check(const char* what, const void* a, int lda, const void* b, int ldb, int M, int N, int K) {
  TADNN_REQUIRE(M > 0 && N >= 1 && N <= 16, "%s: output width %d is outside 1..16 (use tadnn_gemm_bf16)", what, N);
  TADNN_REQUIRE(K >= 8 && K % 8 == 0, "%s: K = %d must be a positive multiple of 8", what, K);
  TADNN_REQUIRE(lda % 8 == 0 && ldb % 8 == 0 && (reinterpret_cast<uintptr_t>(a) & 15) == 0 && (reinterpret_cast<uintptr_t>(b) & 15) == 0,
                "%s: the K-wide operands need 16-byte aligned rows (pitches %d, %d)", what, lda, ldb);
  return 0;
}

extern "C" int tadnn_linear_smalln_fwd(const void* x, int ldx, const void* W, int ldw, const float* bias, float* out, int ldo, int M, int N,
                                       int K, tadnn_stream_t stream_) {
  if (int rc = smalln_check("linear_smalln_fwd", x, ldx, W, ldw, M, N, K)) return rc;
  TADNN_REQUIRE(out && ldo >= N, "linear_smalln_fwd: bad output");
  int tk = 32;  // threads along K: enough that one pass covers the row, capped at the block
  while (tk < 256 && tk * 8 < K) tk <<= 1;
  const long long groups = (static_cast<long long>(M) + kRows - 1) / kRows;
  const int blocks = ceil_div(groups, 256 / tk);
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  SMALLN_DISPATCH(N, (smalln_fwd_kernel<NT><<<blocks, 256, 0, stream>>>(static_cast<const __nv_bfloat16*>(x), ldx,
                                                                        static_cast<const __nv_bfloat16*>(W), ldw, bias, out, ldo, M, N, K, tk)));
  TADNN_LAUNCH_CHECK("linear_smalln_fwd");
  return 0;
}

extern "C" int tadnn_linear_smalln_dgrad(const void* g, int ldg, const void* W, int ldw, const void* aux_tanh_out, int ld_aux, void* dx,
                                         int lddx, int M, int N, int K, tadnn_stream_t stream_) {
  if (int rc = smalln_check("linear_smalln_dgrad", dx, lddx, W, ldw, M, N, K)) return rc;
  TADNN_REQUIRE(g && ldg >= N, "linear_smalln_dgrad: bad gradient pitch");
  TADNN_REQUIRE(!aux_tanh_out || (ld_aux % 8 == 0 && (reinterpret_cast<uintptr_t>(aux_tanh_out) & 15) == 0),
                "linear_smalln_dgrad: the tanh output needs 16-byte aligned rows");
  const long long threads = ((static_cast<long long>(M) + kRows - 1) / kRows) * (K / 8);
  const int blocks = ceil_div(threads, 256);
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  SMALLN_DISPATCH(N, (smalln_dgrad_kernel<NT><<<blocks, 256, 0, stream>>>(
                         static_cast<const __nv_bfloat16*>(g), ldg, static_cast<const __nv_bfloat16*>(W), ldw,
                         static_cast<const __nv_bfloat16*>(aux_tanh_out), ld_aux, static_cast<__nv_bfloat16*>(dx), lddx, M, N, K)));
  TADNN_LAUNCH_CHECK("linear_smalln_dgrad");
  return 0;
}

extern "C" int tadnn_linear_smalln_wgrad(const void* g, int ldg, const void* x, int ldx, float* dW, int lddw, float* db, int M, int N, int K,
                                         int accumulate, tadnn_stream_t stream_) {
  if (int rc = smalln_check("linear_smalln_wgrad", x, ldx, x, ldx, M, N, K)) return rc;
  TADNN_REQUIRE(g && ldg >= N && dW && lddw >= K, "linear_smalln_wgrad: bad arguments");
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (!accumulate) {
    TADNN_CHECK_CUDA(cudaMemset2DAsync(dW, static_cast<size_t>(lddw) * sizeof(float), 0, static_cast<size_t>(K) * sizeof(float), N, stream));
    if (db) TADNN_CHECK_CUDA(cudaMemsetAsync(db, 0, N * sizeof(float), stream));
  }
  const int bx = ceil_div(K, 256);
  // one block per SM holds the register tile; aim at one wave, but keep >= 64 rows per slab so the
  // global reductions at the end stay a small fraction of the work
  long long slabs = sm_count() / bx;
  if (slabs < 1) slabs = 1;
  long long rows = (M + slabs - 1) / slabs;
  if (rows < 64) rows = 64;
  slabs = (M + rows - 1) / rows;
  SMALLN_DISPATCH(N, (smalln_wgrad_kernel<NT><<<dim3(bx, static_cast<unsigned>(slabs)), 256, 0, stream>>>(
                         static_cast<const __nv_bfloat16*>(g), ldg, static_cast<const __nv_bfloat16*>(x), ldx, dW, lddw, db, M, N, K,
                         static_cast<int>(rows))));
  TADNN_LAUNCH_CHECK("linear_smalln_wgrad");
  return 0;
}
