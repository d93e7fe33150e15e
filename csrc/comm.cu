// Symmetric-memory communication context and the collective kernels that run on it
// (K1 one-shot / K2 two-shot / K3 NVLS all-reduce-average, K4 broadcast, K5 device barrier,
//  K7 reduce-scatter, K8 all-gather).  All data movement is done by sm_100a kernels that
// load/store peer-mapped memory over NVLink 5 / NVSwitch; there is no NCCL on this path.
//
// What it replaces in the reference: the TorchMPI library calls made from
// synchronizeModel (datamodule.lua:211-224: per-tensor allreduceTensor + div), from
// parallelize (datamodule.lua:33 synchronizeParameters, :50 barrier) and from the MP layers
// (nodemodule.lua:52,66,103,117 allreduceTensor; :166,266 allgatherTensor).
#include "comm_internal.cuh"

namespace tadnn {
namespace {

__device__ __forceinline__ int my_rank(const CommDev& c) {
  return c.rank >= 0 ? c.rank : static_cast<int>(blockIdx.y);
}

__device__ __forceinline__ unsigned long long globaltimer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
__device__ __forceinline__ uint32_t cas_release_sys(uint32_t* p, uint32_t cmp, uint32_t val) {
  uint32_t old;
  asm volatile("atom.global.release.sys.cas.b32 %0, [%1], %2, %3;" : "=r"(old) : "l"(p), "r"(cmp), "r"(val) : "memory");
  return old;
}
__device__ __forceinline__ uint32_t cas_acquire_sys(uint32_t* p, uint32_t cmp, uint32_t val) {
  uint32_t old;
  asm volatile("atom.global.acquire.sys.cas.b32 %0, [%1], %2, %3;" : "=r"(old) : "l"(p), "r"(cmp), "r"(val) : "memory");
  return old;
}

// K5: block b of every rank meets block b of every other rank.  Flag protocol: signal =
// CAS(peer slot 0->1) with release, wait = CAS(own slot 1->0) with acquire; self-resetting,
// so launches can be replayed (CUDA graphs) without host-side epochs.
// A peer that never shows up (dead or wedged process) does not hang or trap the survivors: after
// c.timeout_ns the waiting thread records (rank, awaited peer, channel, phase) in the host-visible status
// word, poisons the context and returns; see CommDev.
__device__ __noinline__ void barrier_report(const CommDev& c, int rank, int peer, int channel, int phase) {
  const uint32_t code = 1u + ((static_cast<uint32_t>(rank) << 20) | (static_cast<uint32_t>(peer) << 12) |
                              (static_cast<uint32_t>(channel) << 4) | static_cast<uint32_t>(phase));
  if (c.poison != nullptr && atomicCAS(c.poison, 0u, code) == 0u && c.status != nullptr) {
    *reinterpret_cast<volatile uint32_t*>(c.status) = code;
    __threadfence_system();
  }
}
__device__ __forceinline__ bool barrier_spin_check(const CommDev& c, uint32_t& spins, unsigned long long t0, int rank, int peer,
                                                   int channel, int phase) {
  if ((++spins & 255u) != 0u) return false;
  if (c.poison != nullptr && *reinterpret_cast<volatile uint32_t*>(c.poison) != 0u) return true;
  if (globaltimer_ns() - t0 > c.timeout_ns) {
    barrier_report(c, rank, peer, channel, phase);
    return true;
  }
  return false;
}
__device__ __forceinline__ void block_barrier(const CommDev& c, int channel) {
  __syncthreads();
  if (threadIdx.x < c.world) {
    const int rank = my_rank(c);
    const int peer = threadIdx.x;
    const size_t slot = (static_cast<size_t>(channel) * kMaxBlocks + blockIdx.x) * kMaxRanks;
    uint32_t* remote = c.pad[peer] + slot + rank;
    uint32_t* local = c.pad[rank] + slot + peer;
    __threadfence_system();
    const unsigned long long t0 = globaltimer_ns();
    uint32_t spins = 0;
    bool failed = false;
    while (cas_release_sys(remote, 0u, 1u) != 0u) {
      if (barrier_spin_check(c, spins, t0, rank, peer, channel, 0)) { failed = true; break; }
    }
    while (!failed && cas_acquire_sys(local, 1u, 0u) != 1u) {
      if (barrier_spin_check(c, spins, t0, rank, peer, channel, 1)) break;
    }
  }
  __syncthreads();
}

__device__ __forceinline__ uint4 ld_peer16(const void* p) {
  uint4 v;
  asm volatile("ld.volatile.global.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_peer16(void* p, const uint4& v) {
  asm volatile("st.global.v4.u32 [%0], {%1, %2, %3, %4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}

template <int DT>
struct Vec;
template <>
struct Vec<0> {  // fp32 x4
  static constexpr int kN = 4;
  __device__ static void add(float (&a)[8], const uint4& v) {
    a[0] += __uint_as_float(v.x); a[1] += __uint_as_float(v.y);
    a[2] += __uint_as_float(v.z); a[3] += __uint_as_float(v.w);
  }
  __device__ static uint4 pack(const float (&a)[8], float s) {
    return make_uint4(__float_as_uint(a[0] * s), __float_as_uint(a[1] * s), __float_as_uint(a[2] * s), __float_as_uint(a[3] * s));
  }
};
template <>
struct Vec<1> {  // bf16 x8, fp32 accumulation
  static constexpr int kN = 8;
  __device__ static void add(float (&a)[8], const uint4& v) {
    float2 f;
    f = unpack_bf16x2(v.x); a[0] += f.x; a[1] += f.y;
    f = unpack_bf16x2(v.y); a[2] += f.x; a[3] += f.y;
    f = unpack_bf16x2(v.z); a[4] += f.x; a[5] += f.y;
    f = unpack_bf16x2(v.w); a[6] += f.x; a[7] += f.y;
  }
  __device__ static uint4 pack(const float (&a)[8], float s) {
    return make_uint4(pack_bf16x2(a[0] * s, a[1] * s), pack_bf16x2(a[2] * s, a[3] * s),
                      pack_bf16x2(a[4] * s, a[5] * s), pack_bf16x2(a[6] * s, a[7] * s));
  }
};

// ---- K1: one-shot ----------------------------------------------------------
// Every rank reads all P copies of every element, sums in rank order (fp32), scales,
// and -- after a second barrier, so nobody is still reading -- overwrites its own copy.
template <int DT>
__global__ void __launch_bounds__(512) allreduce_oneshot_kernel(CommDev c, size_t off, size_t nvec, float scale, int channel) {
  constexpr int U = 2;
  const size_t stride = static_cast<size_t>(gridDim.x) * blockDim.x;
  const size_t tid = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  const int rank = my_rank(c);
  block_barrier(c, channel);
  for (size_t base = 0; base < nvec; base += stride * U) {
    uint4 out[U];
    uint4 v[U][kMaxRanks];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const size_t i = base + tid + u * stride;
      if (i < nvec) {
#pragma unroll
        for (int r = 0; r < kMaxRanks; ++r)
          if (r < c.world) v[u][r] = ld_peer16(c.heap[r] + off + i * 16);
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      float a[8] = {0, 0, 0, 0, 0, 0, 0, 0};
#pragma unroll
      for (int r = 0; r < kMaxRanks; ++r)
        if (r < c.world) Vec<DT>::add(a, v[u][r]);
      out[u] = Vec<DT>::pack(a, scale);
    }
    block_barrier(c, channel);
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const size_t i = base + tid + u * stride;
      if (i < nvec) *reinterpret_cast<uint4*>(c.heap[rank] + off + i * 16) = out[u];
    }
  }
}

// ---- K2: two-shot ----------------------------------------------------------
// Rank r owns chunk r: it pulls that chunk from all P heaps, reduces, scales and pushes the
// result into chunk r of all P heaps.  In place and race-free: only rank r ever touches chunk r.
template <int DT>
__global__ void __launch_bounds__(512) allreduce_twoshot_kernel(CommDev c, size_t off, size_t nvec, float scale, int channel) {
  constexpr int U = 2;
  const int rank = my_rank(c);
  const size_t chunk = (nvec + c.world - 1) / c.world;
  const size_t lo = min(nvec, chunk * rank), hi = min(nvec, lo + chunk);
  const size_t stride = static_cast<size_t>(gridDim.x) * blockDim.x;
  const size_t tid = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  block_barrier(c, channel);
  for (size_t base = lo; base < hi; base += stride * U) {
    uint4 v[U][kMaxRanks];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const size_t i = base + tid + u * stride;
      if (i < hi) {
#pragma unroll
        for (int r = 0; r < kMaxRanks; ++r)
          if (r < c.world) v[u][r] = ld_peer16(c.heap[r] + off + i * 16);
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const size_t i = base + tid + u * stride;
      if (i < hi) {
        float a[8] = {0, 0, 0, 0, 0, 0, 0, 0};
#pragma unroll
        for (int r = 0; r < kMaxRanks; ++r)
          if (r < c.world) Vec<DT>::add(a, v[u][r]);
        const uint4 o = Vec<DT>::pack(a, scale);
#pragma unroll
        for (int r = 0; r < kMaxRanks; ++r)
          if (r < c.world) st_peer16(c.heap[r] + off + i * 16, o);
      }
    }
  }
  block_barrier(c, channel);
}

// ---- K3: NVLS (multimem) ---------------------------------------------------
// The switch reduces: one multimem.ld_reduce of the multicast address returns the sum over
// all ranks, one multimem.st broadcasts the result.  Rank r handles chunk r.
__device__ __forceinline__ uint4 mm_ld_reduce_f32(const uint8_t* p) {
  uint4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ uint4 mm_ld_reduce_bf16(const uint8_t* p) {
  uint4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void mm_st(uint8_t* p, const uint4& v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
template <int DT>
__device__ __forceinline__ uint4 scale_vec(const uint4& v, float s) {
  if (DT == 0)
    return make_uint4(__float_as_uint(__uint_as_float(v.x) * s), __float_as_uint(__uint_as_float(v.y) * s),
                      __float_as_uint(__uint_as_float(v.z) * s), __float_as_uint(__uint_as_float(v.w) * s));
  float2 a = unpack_bf16x2(v.x), b = unpack_bf16x2(v.y), c = unpack_bf16x2(v.z), d = unpack_bf16x2(v.w);
  return make_uint4(pack_bf16x2(a.x * s, a.y * s), pack_bf16x2(b.x * s, b.y * s), pack_bf16x2(c.x * s, c.y * s), pack_bf16x2(d.x * s, d.y * s));
}
// Four 16-byte switch reductions in flight per thread before the first result is used (the loop was
// latency-bound with one: each multimem.ld_reduce is a round trip to the switch).
template <int DT>
__global__ void __launch_bounds__(512) allreduce_nvls_kernel(CommDev c, size_t off, size_t nvec, float scale, int channel) {
  constexpr int U = 4;
  const int rank = my_rank(c);
  const size_t chunk = (nvec + c.world - 1) / c.world;
  const size_t lo = min(nvec, chunk * rank), hi = min(nvec, lo + chunk);
  const size_t stride = static_cast<size_t>(gridDim.x) * blockDim.x;
  const size_t tid = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  block_barrier(c, channel);
  for (size_t base = lo; base < hi; base += stride * U) {
    uint4 v[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const size_t i = base + tid + u * stride;
      if (i < hi) v[u] = DT == 0 ? mm_ld_reduce_f32(c.mc_heap + off + i * 16) : mm_ld_reduce_bf16(c.mc_heap + off + i * 16);
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const size_t i = base + tid + u * stride;
      if (i < hi) mm_st(c.mc_heap + off + i * 16, scale_vec<DT>(v[u], scale));
    }
  }
  block_barrier(c, channel);
}

// ---- all-reduce-average of the gradients FUSED with the SGD step (fp32) -------------------------------
// The rank that holds a reduced gradient vector in registers applies  p -= lr * (g + wd * p)  right there
// and clears the gradient, so the synchronised training step needs no second pass over the flat buffer
// and no parameter half in the message (the reference sends both, datamodule.lua:214-223, and then runs
// the optimiser separately).  Valid when the replicas hold identical parameters, which the caller
// guarantees (sync every step from a broadcast start); every rank then computes the same update in the
// same rank order, so the parameters stay bit-identical without ever being communicated.
//   one-shot: every rank reads all P gradient copies and updates its own parameters; second barrier
//             before the gradients are cleared (peers may still be reading them).
//   nvls    : rank r owns chunk r: switch-reduced load, update of its own parameter chunk, and the new
//             parameters and the cleared gradients are multicast-stored to every rank.
__global__ void __launch_bounds__(512) allreduce_sgd_oneshot_kernel(CommDev c, size_t g_off, size_t p_off, size_t nvec, float scale,
                                                                    float lr, float wd, int channel) {
  const size_t stride = static_cast<size_t>(gridDim.x) * blockDim.x;
  const size_t tid = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  const int rank = my_rank(c);
  block_barrier(c, channel);
  for (size_t i = tid; i < nvec; i += stride) {
    uint4 v[kMaxRanks];
#pragma unroll
    for (int r = 0; r < kMaxRanks; ++r)
      if (r < c.world) v[r] = ld_peer16(c.heap[r] + g_off + i * 16);
    float a[8] = {0, 0, 0, 0, 0, 0, 0, 0};
#pragma unroll
    for (int r = 0; r < kMaxRanks; ++r)
      if (r < c.world) Vec<0>::add(a, v[r]);
    float4* pp = reinterpret_cast<float4*>(c.heap[rank] + p_off + i * 16);
    float4 w = *pp;
    w.x -= lr * (a[0] * scale + wd * w.x); w.y -= lr * (a[1] * scale + wd * w.y);
    w.z -= lr * (a[2] * scale + wd * w.z); w.w -= lr * (a[3] * scale + wd * w.w);
    *pp = w;
  }
  block_barrier(c, channel);
  for (size_t i = tid; i < nvec; i += stride) *reinterpret_cast<uint4*>(c.heap[rank] + g_off + i * 16) = make_uint4(0, 0, 0, 0);
}

__global__ void __launch_bounds__(512) allreduce_sgd_nvls_kernel(CommDev c, size_t g_off, size_t p_off, size_t nvec, float scale, float lr,
                                                                 float wd, int channel) {
  constexpr int U = 4;
  const int rank = my_rank(c);
  const size_t chunk = (nvec + c.world - 1) / c.world;
  const size_t lo = min(nvec, chunk * rank), hi = min(nvec, lo + chunk);
  const size_t stride = static_cast<size_t>(gridDim.x) * blockDim.x;
  const size_t tid = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  block_barrier(c, channel);
  for (size_t base = lo; base < hi; base += stride * U) {
    uint4 v[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const size_t i = base + tid + u * stride;
      if (i < hi) v[u] = mm_ld_reduce_f32(c.mc_heap + g_off + i * 16);
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const size_t i = base + tid + u * stride;
      if (i < hi) {
        float4 w = *reinterpret_cast<const float4*>(c.heap[rank] + p_off + i * 16);
        w.x -= lr * (__uint_as_float(v[u].x) * scale + wd * w.x); w.y -= lr * (__uint_as_float(v[u].y) * scale + wd * w.y);
        w.z -= lr * (__uint_as_float(v[u].z) * scale + wd * w.z); w.w -= lr * (__uint_as_float(v[u].w) * scale + wd * w.w);
        mm_st(c.mc_heap + p_off + i * 16, make_uint4(__float_as_uint(w.x), __float_as_uint(w.y), __float_as_uint(w.z), __float_as_uint(w.w)));
        mm_st(c.mc_heap + g_off + i * 16, make_uint4(0, 0, 0, 0));
      }
    }
  }
  block_barrier(c, channel);
}

// two-shot form of the same for contexts without a multicast mapping
__global__ void __launch_bounds__(512) allreduce_sgd_twoshot_kernel(CommDev c, size_t g_off, size_t p_off, size_t nvec, float scale, float lr,
                                                                    float wd, int channel) {
  const int rank = my_rank(c);
  const size_t chunk = (nvec + c.world - 1) / c.world;
  const size_t lo = min(nvec, chunk * rank), hi = min(nvec, lo + chunk);
  const size_t stride = static_cast<size_t>(gridDim.x) * blockDim.x;
  block_barrier(c, channel);
  for (size_t i = lo + static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < hi; i += stride) {
    uint4 v[kMaxRanks];
#pragma unroll
    for (int r = 0; r < kMaxRanks; ++r)
      if (r < c.world) v[r] = ld_peer16(c.heap[r] + g_off + i * 16);
    float a[8] = {0, 0, 0, 0, 0, 0, 0, 0};
#pragma unroll
    for (int r = 0; r < kMaxRanks; ++r)
      if (r < c.world) Vec<0>::add(a, v[r]);
    float4 w = *reinterpret_cast<const float4*>(c.heap[rank] + p_off + i * 16);
    w.x -= lr * (a[0] * scale + wd * w.x); w.y -= lr * (a[1] * scale + wd * w.y);
    w.z -= lr * (a[2] * scale + wd * w.z); w.w -= lr * (a[3] * scale + wd * w.w);
    const uint4 o = make_uint4(__float_as_uint(w.x), __float_as_uint(w.y), __float_as_uint(w.z), __float_as_uint(w.w));
    const uint4 z = make_uint4(0, 0, 0, 0);
#pragma unroll
    for (int r = 0; r < kMaxRanks; ++r)
      if (r < c.world) {
        st_peer16(c.heap[r] + p_off + i * 16, o);
        st_peer16(c.heap[r] + g_off + i * 16, z);
      }
  }
  block_barrier(c, channel);
}

// ---- low-latency all-reduce (fp32, <= kLLMaxFloats): data and flag travel together ------------------
// Every rank stores its vector into ITS slot of every rank's staging area as 8-byte {value, epoch} pairs
// (one NVLink one-way trip), then sums the P slots of its OWN staging area, spinning on each pair until its
// epoch matches: no barrier, no second round trip.  The staging area is double-buffered on the parity of the
// epoch: a peer can be at most one call ahead (it cannot finish call e+1 without our data for e+1, which we
// send only after we have read everything of call e), so call e+1 never overwrites pairs of call e that are
// still unread.  MODE 0: result written in place; MODE 1: fused SGD step on `p`, gradient cleared.
constexpr int kLLMaxFloats = TADNN_LL_MAX_FLOATS;
template <int MODE>
__global__ void __launch_bounds__(512) allreduce_ll_kernel(CommDev c, size_t ll_off, size_t data_off, size_t p_off, int n, float scale,
                                                           float lr, float wd, uint32_t epoch) {
  // programmatic dependent launch: the gradients were written by the kernel before this one; the kernel after it (the next
  // step's operand packing) waits for this grid's completion in its own griddepcontrol.wait
  asm volatile("griddepcontrol.wait;" ::: "memory");
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  const int rank = my_rank(c);
  const int stride = gridDim.x * blockDim.x;
  const int tid = blockIdx.x * blockDim.x + threadIdx.x;
  const size_t half = static_cast<size_t>(c.world) * kLLMaxFloats * 8;
  const size_t base = ll_off + (epoch & 1u) * half;
  float* data = reinterpret_cast<float*>(c.heap[rank] + data_off);
  for (int i = tid; i < n; i += stride) {
    const uint32_t v = __float_as_uint(data[i]);
    const size_t o = base + (static_cast<size_t>(rank) * kLLMaxFloats + i) * 8;
#pragma unroll
    for (int q = 0; q < kMaxRanks; ++q)
      if (q < c.world) asm volatile("st.relaxed.sys.global.v2.u32 [%0], {%1, %2};" ::"l"(c.heap[q] + o), "r"(v), "r"(epoch) : "memory");
  }
  const unsigned long long t0 = globaltimer_ns();
  for (int i = tid; i < n; i += stride) {
    float acc = 0.f;
    bool failed = false;
    for (int r = 0; r < c.world && !failed; ++r) {
      const uint8_t* src = c.heap[rank] + base + (static_cast<size_t>(r) * kLLMaxFloats + i) * 8;
      uint32_t v, e, spins = 0;
      for (;;) {
        asm volatile("ld.relaxed.sys.global.v2.u32 {%0, %1}, [%2];" : "=r"(v), "=r"(e) : "l"(src) : "memory");
        if (e == epoch) break;
        if (barrier_spin_check(c, spins, t0, rank, r, 7, 2)) { failed = true; break; }
      }
      acc += __uint_as_float(v);
    }
    const float g = acc * scale;
    if (MODE == 0) {
      data[i] = g;
    } else {
      float* pp = reinterpret_cast<float*>(c.heap[rank] + p_off) + i;
      const float w = *pp;
      *pp = w - lr * (g + wd * w);
      data[i] = 0.f;
    }
  }
}

// ---- tail (< 16 B) handled by a scalar one-shot on a single block -----------
template <int DT>
__global__ void allreduce_tail_kernel(CommDev c, size_t off_bytes, int n, float scale, int channel) {
  const int rank = my_rank(c);
  block_barrier(c, channel);
  float acc = 0.f;
  if (static_cast<int>(threadIdx.x) < n) {
    for (int r = 0; r < c.world; ++r) {
      if (DT == 0) acc += *reinterpret_cast<volatile float*>(c.heap[r] + off_bytes + threadIdx.x * 4);
      else acc += __bfloat162float(*reinterpret_cast<volatile __nv_bfloat16*>(c.heap[r] + off_bytes + threadIdx.x * 2));
    }
  }
  block_barrier(c, channel);
  if (static_cast<int>(threadIdx.x) < n) {
    if (DT == 0) *reinterpret_cast<float*>(c.heap[rank] + off_bytes + threadIdx.x * 4) = acc * scale;
    else *reinterpret_cast<__nv_bfloat16*>(c.heap[rank] + off_bytes + threadIdx.x * 2) = __float2bfloat16_rn(acc * scale);
  }
}

__global__ void barrier_kernel(CommDev c, int channel) { block_barrier(c, channel); }

// ---- K4: broadcast (root pushes) ------------------------------------------
__global__ void __launch_bounds__(512) broadcast_kernel(CommDev c, size_t off, size_t nvec, size_t tail_off, int tail_bytes, int root, int channel) {
  const int rank = my_rank(c);
  block_barrier(c, channel);
  if (rank == root) {
    const size_t stride = static_cast<size_t>(gridDim.x) * blockDim.x;
    for (size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < nvec; i += stride) {
      const uint4 v = *reinterpret_cast<const uint4*>(c.heap[rank] + off + i * 16);
      for (int r = 0; r < c.world; ++r)
        if (r != rank) st_peer16(c.heap[r] + off + i * 16, v);
    }
    if (blockIdx.x == 0 && static_cast<int>(threadIdx.x) < tail_bytes) {
      const uint8_t b = c.heap[rank][tail_off + threadIdx.x];
      for (int r = 0; r < c.world; ++r)
        if (r != rank) c.heap[r][tail_off + threadIdx.x] = b;
    }
  }
  block_barrier(c, channel);
}

// ---- K8: all-gather along the last dim (push) -----------------------------
template <typename V>
__global__ void __launch_bounds__(512) allgather_cols_kernel(CommDev c, const uint8_t* src, size_t src_pitch, size_t off, int rows, size_t width_v /* in V units */, int channel) {
  const int rank = my_rank(c);
  const size_t total = static_cast<size_t>(rows) * width_v;
  const size_t out_pitch = width_v * sizeof(V) * c.world;
  const size_t stride = static_cast<size_t>(gridDim.x) * blockDim.x;
  block_barrier(c, channel);
  for (size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < total; i += stride) {
    const size_t row = i / width_v, col = i - row * width_v;
    const V v = *reinterpret_cast<const V*>(src + row * src_pitch + col * sizeof(V));
    const size_t o = off + row * out_pitch + (static_cast<size_t>(rank) * width_v + col) * sizeof(V);
    for (int r = 0; r < c.world; ++r) *reinterpret_cast<V*>(c.heap[r] + o) = v;
  }
  block_barrier(c, channel);
}

// ---- K7: reduce-scatter along the last dim (pull) with fused bias + act ----
template <int DT>
__global__ void __launch_bounds__(512) reduce_scatter_cols_kernel(CommDev c, size_t off, void* dst, size_t dst_pitch, int rows, int width, float scale, const float* bias, int act, int channel) {
  const int rank = my_rank(c);
  const size_t es = DT == 0 ? 4 : 2;
  const size_t total = static_cast<size_t>(rows) * width;
  const size_t in_pitch = static_cast<size_t>(width) * c.world * es;
  const size_t stride = static_cast<size_t>(gridDim.x) * blockDim.x;
  block_barrier(c, channel);
  for (size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < total; i += stride) {
    const size_t row = i / width, col = i - row * width;
    const size_t o = off + row * in_pitch + (static_cast<size_t>(rank) * width + col) * es;
    float acc = 0.f;
    for (int r = 0; r < c.world; ++r) {
      if (DT == 0) acc += *reinterpret_cast<volatile float*>(c.heap[r] + o);
      else acc += __bfloat162float(*reinterpret_cast<volatile __nv_bfloat16*>(c.heap[r] + o));
    }
    acc *= scale;
    if (bias != nullptr) acc += bias[static_cast<size_t>(rank) * width + col];
    if (act == 1) acc = tanh_f32(acc);
    else if (act == 2) acc = fmaxf(acc, 0.f);
    if (DT == 0) reinterpret_cast<float*>(dst)[row * dst_pitch + col] = acc;
    else reinterpret_cast<__nv_bfloat16*>(dst)[row * dst_pitch + col] = __float2bfloat16_rn(acc);
  }
  block_barrier(c, channel);
}

// ---- finalize of the fused GEMM -> reduce-scatter: sum the P slots that the peers' GEMM epilogues
// pushed into THIS rank's staging buffer, add bias, activate; optionally push the result to every
// rank (all-gather), which turns the pair into a fused all-reduce.
__global__ void __launch_bounds__(512) reduce_slots_kernel(CommDev c, size_t slots_off, int rows, int width8 /* width/8 */,
                                                           float scale, const float* bias, int act,
                                                           __nv_bfloat16* out_local, size_t out_pitch,
                                                           size_t push_off, int do_push, int channel) {
  const int rank = my_rank(c);
  const size_t total = static_cast<size_t>(rows) * width8;
  const size_t slot_elems8 = total;  // uint4 units per slot
  const size_t stride = static_cast<size_t>(gridDim.x) * blockDim.x;
  const uint4* slots = reinterpret_cast<const uint4*>(c.heap[rank] + slots_off);
  block_barrier(c, channel);  // every rank's GEMM (and its remote stores) has completed
  for (size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < total; i += stride) {
    const size_t row = i / width8, c8 = i - row * width8;
    float a[8] = {0, 0, 0, 0, 0, 0, 0, 0};
#pragma unroll
    for (int q = 0; q < kMaxRanks; ++q)
      if (q < c.world) Vec<1>::add(a, slots[q * slot_elems8 + i]);
    float v[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      float x = a[e] * scale;
      if (bias != nullptr) x += bias[(static_cast<size_t>(rank) * width8 + c8) * 8 + e];
      if (act == 1) x = tanh_f32(x);
      else if (act == 2) x = fmaxf(x, 0.f);
      v[e] = x;
    }
    const uint4 o = make_uint4(pack_bf16x2(v[0], v[1]), pack_bf16x2(v[2], v[3]), pack_bf16x2(v[4], v[5]), pack_bf16x2(v[6], v[7]));
    if (out_local != nullptr) *reinterpret_cast<uint4*>(out_local + row * out_pitch + c8 * 8) = o;
    if (do_push) {
      const size_t off = push_off + (row * (static_cast<size_t>(width8) * c.world) + static_cast<size_t>(rank) * width8 + c8) * 16;
#pragma unroll
      for (int q = 0; q < kMaxRanks; ++q)
        if (q < c.world) st_peer16(c.heap[q] + off, o);
    }
  }
  block_barrier(c, channel);  // slots may be overwritten / pushed results are visible
}

// ---- reduce-scatter (optionally all-reduce) of column-split GEMM partials INSIDE the switch (NVLS).
// Every rank's GEMM wrote its full partial [rows, world*width] (bf16) at the same heap offset.  Rank r
// owns columns [r*width, (r+1)*width): one multimem.ld_reduce per 16 bytes returns the sum over all
// ranks (fp32 accumulation in the switch), so a rank RECEIVES rows*width*2 bytes instead of the
// (world-1) slots the push variant stores into it; bias/activation are applied on the way to the
// local output, and `bcast` multimem.st's the result back into every rank's buffer (all-reduce).
__global__ void __launch_bounds__(512) reduce_cols_nvls_kernel(CommDev c, size_t part_off, int rows, int width8, float scale,
                                                               const float* bias, int act, __nv_bfloat16* out_local, size_t out_pitch,
                                                               int bcast, int channel) {
  const int rank = my_rank(c);
  const size_t total = static_cast<size_t>(rows) * width8;
  const size_t stride = static_cast<size_t>(gridDim.x) * blockDim.x;
  const size_t row_vecs = static_cast<size_t>(width8) * c.world;
  block_barrier(c, channel);  // every rank's partial is complete and visible
  for (size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < total; i += stride) {
    const size_t row = i / width8, c8 = i - row * width8;
    uint8_t* p = c.mc_heap + part_off + (row * row_vecs + static_cast<size_t>(rank) * width8 + c8) * 16;
    uint32_t x, y, z, w;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0, %1, %2, %3}, [%4];"
                 : "=r"(x), "=r"(y), "=r"(z), "=r"(w) : "l"(p) : "memory");
    float v[8];
    float2 f;
    f = unpack_bf16x2(x); v[0] = f.x; v[1] = f.y;
    f = unpack_bf16x2(y); v[2] = f.x; v[3] = f.y;
    f = unpack_bf16x2(z); v[4] = f.x; v[5] = f.y;
    f = unpack_bf16x2(w); v[6] = f.x; v[7] = f.y;
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      float t = v[e] * scale;
      if (bias != nullptr) t += bias[(static_cast<size_t>(rank) * width8 + c8) * 8 + e];
      if (act == 1) t = tanh_f32(t);
      else if (act == 2) t = fmaxf(t, 0.f);
      v[e] = t;
    }
    const uint4 o = make_uint4(pack_bf16x2(v[0], v[1]), pack_bf16x2(v[2], v[3]), pack_bf16x2(v[4], v[5]), pack_bf16x2(v[6], v[7]));
    if (out_local != nullptr) *reinterpret_cast<uint4*>(out_local + row * out_pitch + c8 * 8) = o;
    if (bcast)
      asm volatile("multimem.st.relaxed.sys.global.v4.bf16x2 [%0], {%1, %2, %3, %4};" ::"l"(p), "r"(o.x), "r"(o.y), "r"(o.z), "r"(o.w) : "memory");
  }
  block_barrier(c, channel);  // partials may be overwritten / broadcast results are visible
}

// Slab-major staging (GEMM rs_slab_major): slot q holds ceil(rows/32) x (width/64) slabs of 4 KB, each a
// 32-row x 64-column tile in the SWIZZLE_128B order of the GEMM's staging buffer (16-byte chunk j of
// slab row rr sits at chunk position j ^ (rr & 7)).  Threads walk the slots linearly (fully
// coalesced) and undo the swizzle when they write the row-major result.
__global__ void __launch_bounds__(512) reduce_slots_slab_kernel(CommDev c, size_t slots_off, int rows, int width8, float scale,
                                                                const float* bias, int act, __nv_bfloat16* out_local,
                                                                size_t out_pitch, size_t push_off, int do_push, int channel) {
  const int rank = my_rank(c);
  const int slabs_per_rowblk = width8 >> 3;                       // width / 64
  const size_t rowblks = (static_cast<size_t>(rows) + 31) >> 5;
  const size_t total = rowblks * slabs_per_rowblk * 256;           // uint4 units per slot
  const size_t stride = static_cast<size_t>(gridDim.x) * blockDim.x;
  const uint4* slots = reinterpret_cast<const uint4*>(c.heap[rank] + slots_off);
  block_barrier(c, channel);
  for (size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < total; i += stride) {
    const size_t slab = i >> 8;
    const int within = static_cast<int>(i & 255), rr = within >> 3, pos = within & 7;
    const int j = pos ^ (rr & 7);
    const size_t row = (slab / slabs_per_rowblk) * 32 + rr;
    const size_t c8 = (slab % slabs_per_rowblk) * 8 + j;
    if (row >= static_cast<size_t>(rows)) continue;
    float a[8] = {0, 0, 0, 0, 0, 0, 0, 0};
#pragma unroll
    for (int q = 0; q < kMaxRanks; ++q)
      if (q < c.world) Vec<1>::add(a, slots[q * total + i]);
    float v[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      float x = a[e] * scale;
      if (bias != nullptr) x += bias[(static_cast<size_t>(rank) * width8 + c8) * 8 + e];
      if (act == 1) x = tanh_f32(x);
      else if (act == 2) x = fmaxf(x, 0.f);
      v[e] = x;
    }
    const uint4 o = make_uint4(pack_bf16x2(v[0], v[1]), pack_bf16x2(v[2], v[3]), pack_bf16x2(v[4], v[5]), pack_bf16x2(v[6], v[7]));
    if (out_local != nullptr) *reinterpret_cast<uint4*>(out_local + row * out_pitch + c8 * 8) = o;
    if (do_push) {
      const size_t off = push_off + (row * (static_cast<size_t>(width8) * c.world) + static_cast<size_t>(rank) * width8 + c8) * 16;
#pragma unroll
      for (int q = 0; q < kMaxRanks; ++q)
        if (q < c.world) st_peer16(c.heap[q] + off, o);
    }
  }
  block_barrier(c, channel);
}

int pick_blocks(size_t work_items_per_thread_1, int max_blocks, int cap_default) {
  int cap = max_blocks > 0 ? max_blocks : cap_default;
  if (cap > kMaxBlocks) cap = kMaxBlocks;
  long long need = static_cast<long long>((work_items_per_thread_1 + 511) / 512);
  if (need < 1) need = 1;
  return need < cap ? static_cast<int>(need) : cap;
}

int validate(const tadnn_comm* c, int channel, const char* what) {
  TADNN_REQUIRE(c != nullptr, "%s: null comm", what);
  TADNN_REQUIRE(channel >= 0 && channel < kChannels, "%s: channel %d out of range [0,%d)", what, channel, kChannels);
  TADNN_REQUIRE(!c->local_set || c->dev.rank < 0, "%s: a local (single-GPU) comm set can only be used through the *_emulated entry points", what);
  for (int r = 0; r < c->world; ++r)
    TADNN_REQUIRE(c->dev.heap[r] != nullptr && c->dev.pad[r] != nullptr,
                  "%s: the heap of rank %d is not mapped yet (tadnn_comm_open_peers / tadnn_comm_import_peer_fd first)", what, r);
  return 0;
}

template <typename K, typename... Args>
int launch_maybe_coop(const CommDev& dev, K kern, int blocks, int threads, cudaStream_t stream, const char* name, Args... args) {
  if (dev.rank >= 0) {
    kern<<<blocks, threads, 0, stream>>>(args...);
  } else {
    // emulation: all ranks in ONE cooperative launch (co-residency guaranteed or the launch fails)
    void* params[] = {(void*)&args...};
    dim3 grid(blocks, dev.world, 1);
    cudaError_t e = cudaLaunchCooperativeKernel((const void*)kern, grid, dim3(threads, 1, 1), params, 0, stream);
    if (e != cudaSuccess) {
      set_error("cooperative launch of %s (%d x %d blocks) failed: %s", name, blocks, dev.world, cudaGetErrorString(e));
      return 3;
    }
  }
  TADNN_LAUNCH_CHECK(name);
  return 0;
}

int do_allreduce(const CommDev& dev, bool has_mc, unsigned long long off_bytes, unsigned long long count, int dtype, float scale, int algo, int max_blocks, int channel, cudaStream_t stream) {
  TADNN_REQUIRE(dtype == 0 || dtype == 1, "allreduce: dtype must be 0 (fp32) or 1 (bf16)");
  TADNN_REQUIRE(off_bytes % 16 == 0, "allreduce: offset %llu is not 16-byte aligned", off_bytes);
  if (count == 0) return 0;
  const size_t es = dtype == 0 ? 4 : 2;
  const size_t per_vec = 16 / es;
  const size_t nvec = count / per_vec;
  const int tail = static_cast<int>(count - nvec * per_vec);
  const size_t bytes = count * es;
  if (algo == 0) {
    // one-shot moves (P-1)*S per rank but has the shortest critical path; two-shot / NVLS
    // move ~2(P-1)/P*S.  Crossover measured on NVSwitch sits at a few hundred KB.
    // (measured at P = 8: 64 KB one-shot 23 us vs two-shot 18 / NVLS 19; 1 MB 35 vs 26 / 24)
    if (bytes * static_cast<size_t>(dev.world - 1) <= (512u << 10)) algo = 1;
    else algo = has_mc ? 3 : 2;
  }
  TADNN_REQUIRE(algo >= 1 && algo <= 3, "allreduce: unknown algo %d", algo);
  TADNN_REQUIRE(algo != 3 || has_mc, "allreduce: NVLS requested but this context has no multicast mapping");
  const bool emu = dev.rank < 0;
  int rc = 0;
  if (nvec > 0) {
    if (algo == 1) {
      const int blocks = pick_blocks((nvec + 1) / 2, emu ? (max_blocks > 0 ? max_blocks : 8) : max_blocks, 64);
      rc = dtype == 0 ? launch_maybe_coop(dev, allreduce_oneshot_kernel<0>, blocks, 512, stream, "allreduce_oneshot_f32", dev, (size_t)off_bytes, nvec, scale, channel)
                      : launch_maybe_coop(dev, allreduce_oneshot_kernel<1>, blocks, 512, stream, "allreduce_oneshot_bf16", dev, (size_t)off_bytes, nvec, scale, channel);
    } else if (algo == 2) {
      const int blocks = pick_blocks((nvec / dev.world + 1) / 2 + 1, emu ? (max_blocks > 0 ? max_blocks : 8) : max_blocks, 128);
      rc = dtype == 0 ? launch_maybe_coop(dev, allreduce_twoshot_kernel<0>, blocks, 512, stream, "allreduce_twoshot_f32", dev, (size_t)off_bytes, nvec, scale, channel)
                      : launch_maybe_coop(dev, allreduce_twoshot_kernel<1>, blocks, 512, stream, "allreduce_twoshot_bf16", dev, (size_t)off_bytes, nvec, scale, channel);
    } else {
      const int blocks = pick_blocks(nvec / dev.world + 1, max_blocks, 64);
      rc = dtype == 0 ? launch_maybe_coop(dev, allreduce_nvls_kernel<0>, blocks, 512, stream, "allreduce_nvls_f32", dev, (size_t)off_bytes, nvec, scale, channel)
                      : launch_maybe_coop(dev, allreduce_nvls_kernel<1>, blocks, 512, stream, "allreduce_nvls_bf16", dev, (size_t)off_bytes, nvec, scale, channel);
    }
    if (rc) return rc;
  }
  if (tail > 0) {
    const size_t toff = off_bytes + nvec * 16;
    rc = dtype == 0 ? launch_maybe_coop(dev, allreduce_tail_kernel<0>, 1, 32, stream, "allreduce_tail_f32", dev, toff, tail, scale, channel)
                    : launch_maybe_coop(dev, allreduce_tail_kernel<1>, 1, 32, stream, "allreduce_tail_bf16", dev, toff, tail, scale, channel);
  }
  return rc;
}

}  // namespace

int comm_init_status(tadnn_comm* c) {
  c->dev.timeout_ns = kDefaultTimeoutNs;
  if (const char* e = getenv("TADNN_COMM_TIMEOUT_S")) {
    const double s = atof(e);
    if (s > 0) c->dev.timeout_ns = static_cast<unsigned long long>(s * 1e9);
  }
  void* h = nullptr;
  TADNN_CHECK_CUDA(cudaHostAlloc(&h, 64, cudaHostAllocMapped | cudaHostAllocPortable));
  memset(h, 0, 64);
  c->status_host = static_cast<uint32_t*>(h);
  void* hd = nullptr;
  TADNN_CHECK_CUDA(cudaHostGetDevicePointer(&hd, h, 0));
  c->dev.status = static_cast<uint32_t*>(hd);
  TADNN_CHECK_CUDA(cudaMalloc(&c->poison_dev, 16));
  TADNN_CHECK_CUDA(cudaMemset(c->poison_dev, 0, 16));
  c->dev.poison = static_cast<uint32_t*>(c->poison_dev);
  return 0;
}
void comm_free_status(tadnn_comm* c) {
  if (c->status_host) cudaFreeHost(c->status_host);
  if (c->poison_dev) cudaFree(c->poison_dev);
  c->status_host = nullptr;
  c->poison_dev = nullptr;
}

}  // namespace tadnn

using namespace tadnn;

extern "C" int tadnn_comm_create(int rank, int world, unsigned long long heap_bytes, tadnn_comm** out) {
  TADNN_REQUIRE(out != nullptr, "comm_create: null out");
  TADNN_REQUIRE(world >= 1 && world <= kMaxRanks, "comm_create: world %d not in [1,%d]", world, kMaxRanks);
  TADNN_REQUIRE(rank >= 0 && rank < world, "comm_create: rank %d not in [0,%d)", rank, world);
  *out = nullptr;
  heap_bytes = (heap_bytes + 1023ull) & ~1023ull;
  int device = 0;
  void* base = nullptr;
  TADNN_CHECK_CUDA(cudaGetDevice(&device));
  TADNN_CHECK_CUDA(cudaMalloc(&base, heap_bytes + kPadBytes));
  cudaError_t e = cudaMemset(static_cast<uint8_t*>(base) + heap_bytes, 0, kPadBytes);
  if (e == cudaSuccess) e = cudaDeviceSynchronize();
  if (e != cudaSuccess) {
    cudaFree(base);
    set_error("comm_create: zeroing the signal pad failed: %s", cudaGetErrorString(e));
    return 1;
  }
  tadnn_comm* c = new tadnn_comm();
  c->rank = rank; c->world = world; c->heap_bytes = heap_bytes;
  c->device = device;
  c->base = base;
  c->owns = true;
  memset(&c->dev, 0, sizeof(c->dev));
  c->dev.rank = rank; c->dev.world = world;
  c->dev.heap[rank] = static_cast<uint8_t*>(c->base);
  c->dev.pad[rank] = reinterpret_cast<uint32_t*>(static_cast<uint8_t*>(c->base) + heap_bytes);
  if (int rc = comm_init_status(c)) {
    cudaFree(base);
    delete c;
    return rc;
  }
  *out = c;
  return 0;
}

extern "C" int tadnn_comm_create_local(int world, unsigned long long heap_bytes, tadnn_comm** out) {
  TADNN_REQUIRE(out != nullptr, "comm_create_local: null out");
  TADNN_REQUIRE(world >= 1 && world <= kMaxRanks, "comm_create_local: world %d not in [1,%d]", world, kMaxRanks);
  for (int r = 0; r < world; ++r) {
    int rc = tadnn_comm_create(r, world, heap_bytes, &out[r]);
    if (rc) {  // nothing half-built is handed back
      for (int q = 0; q < r; ++q) { tadnn_comm_destroy(out[q]); out[q] = nullptr; }
      return rc;
    }
    out[r]->local_set = true;
  }
  for (int r = 0; r < world; ++r)
    for (int q = 0; q < world; ++q) {
      out[r]->dev.heap[q] = out[q]->dev.heap[q];
      out[r]->dev.pad[q] = out[q]->dev.pad[q];
    }
  return 0;
}

extern "C" int tadnn_comm_ipc_handle(tadnn_comm* c, void* host_out64) {
  TADNN_REQUIRE(c && host_out64, "comm_ipc_handle: null argument");
  TADNN_REQUIRE(c->owns && !c->vmm, "comm_ipc_handle: context does not own a cudaMalloc heap");
  static_assert(sizeof(cudaIpcMemHandle_t) == TADNN_IPC_HANDLE_BYTES, "IPC handle size");
  cudaIpcMemHandle_t h;
  TADNN_CHECK_CUDA(cudaIpcGetMemHandle(&h, c->base));
  memcpy(host_out64, &h, sizeof(h));
  return 0;
}

extern "C" int tadnn_comm_open_peers(tadnn_comm* c, const void* host_all_handles) {
  TADNN_REQUIRE(c && host_all_handles, "comm_open_peers: null argument");
  const uint8_t* hs = static_cast<const uint8_t*>(host_all_handles);
  for (int r = 0; r < c->world; ++r) {
    if (r == c->rank) continue;
    cudaIpcMemHandle_t h;
    memcpy(&h, hs + static_cast<size_t>(r) * TADNN_IPC_HANDLE_BYTES, sizeof(h));
    void* p = nullptr;
    TADNN_CHECK_CUDA(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
    c->opened.push_back(p);
    c->dev.heap[r] = static_cast<uint8_t*>(p);
    c->dev.pad[r] = reinterpret_cast<uint32_t*>(static_cast<uint8_t*>(p) + c->heap_bytes);
  }
  return 0;
}

extern "C" int tadnn_comm_attach(tadnn_comm* c, void* const* heaps, void* const* pads, void* multicast_heap) {
  TADNN_REQUIRE(c != nullptr, "comm_attach: null comm");
  for (int r = 0; r < c->world; ++r) {
    if (heaps && heaps[r]) c->dev.heap[r] = static_cast<uint8_t*>(heaps[r]);
    if (pads && pads[r]) c->dev.pad[r] = static_cast<uint32_t*>(pads[r]);
  }
  c->dev.mc_heap = static_cast<uint8_t*>(multicast_heap);
  return 0;
}

extern "C" int tadnn_comm_heap(tadnn_comm* c, int peer, void** out_ptr) {
  TADNN_REQUIRE(c && out_ptr && peer >= 0 && peer < c->world, "comm_heap: bad argument");
  *out_ptr = c->dev.heap[peer];
  return 0;
}

extern "C" int tadnn_comm_info(tadnn_comm* c, int* rank, int* world, unsigned long long* heap_bytes, int* has_mc) {
  TADNN_REQUIRE(c != nullptr, "comm_info: null comm");
  if (rank) *rank = c->rank;
  if (world) *world = c->world;
  if (heap_bytes) *heap_bytes = c->heap_bytes;
  if (has_mc) *has_mc = c->dev.mc_heap != nullptr;
  return 0;
}

extern "C" int tadnn_comm_destroy(tadnn_comm* c) {
  if (c == nullptr) return 0;
  cudaDeviceSynchronize();
  if (c->vmm) {
    tadnn_comm_vmm_release(c);
  } else {
    for (void* p : c->opened) cudaIpcCloseMemHandle(p);
    if (c->owns && c->base) cudaFree(c->base);
  }
  comm_free_status(c);
  delete c;
  return 0;
}

extern "C" int tadnn_comm_set_timeout(tadnn_comm* c, double seconds) {
  TADNN_REQUIRE(c != nullptr && seconds > 0, "comm_set_timeout: need a context and a positive number of seconds");
  c->dev.timeout_ns = static_cast<unsigned long long>(seconds * 1e9);
  return 0;
}

extern "C" int tadnn_comm_check(tadnn_comm* c) {
  TADNN_REQUIRE(c != nullptr, "comm_check: null comm");
  if (c->status_host == nullptr) return 0;
  const uint32_t code = *reinterpret_cast<volatile uint32_t*>(c->status_host);
  if (code == 0) return 0;
  const uint32_t v = code - 1u;
  set_error("collective timed out: rank %u waited more than %.1f s for rank %u (channel %u, %s) - that peer is dead, wedged or "
            "never entered the collective; this communicator is unusable from here on",
            (v >> 20) & 0xffu, static_cast<double>(c->dev.timeout_ns) * 1e-9, (v >> 12) & 0xffu, (v >> 4) & 0xffu,
            (v & 15u) ? "waiting for its arrival flag" : "signalling it: its previous flag was never consumed");
  return 4;
}

extern "C" int tadnn_barrier(tadnn_comm* c, int channel, tadnn_stream_t stream) {
  if (int rc = validate(c, channel, "barrier")) return rc;
  if (c->world == 1) return 0;
  barrier_kernel<<<1, 32, 0, static_cast<cudaStream_t>(stream)>>>(c->dev, channel);
  TADNN_LAUNCH_CHECK("device_barrier");
  return 0;
}

extern "C" int tadnn_allreduce(tadnn_comm* c, unsigned long long off_bytes, unsigned long long count, int dtype, float scale, int algo, int max_blocks, int channel, tadnn_stream_t stream) {
  if (int rc = validate(c, channel, "allreduce")) return rc;
  const unsigned long long es = dtype == 0 ? 4 : 2;
  TADNN_REQUIRE(off_bytes + count * es <= c->heap_bytes, "allreduce: [%llu, +%llu) exceeds the %llu-byte heap", off_bytes, count * es, c->heap_bytes);
  return do_allreduce(c->dev, c->dev.mc_heap != nullptr, off_bytes, count, dtype, scale, algo, max_blocks, channel, static_cast<cudaStream_t>(stream));
}

extern "C" int tadnn_allreduce_sgd(tadnn_comm* c, unsigned long long grad_off, unsigned long long param_off, unsigned long long count,
                                   float scale, float lr, float weight_decay, int algo, int max_blocks, int channel, tadnn_stream_t stream_) {
  if (int rc = validate(c, channel, "allreduce_sgd")) return rc;
  TADNN_REQUIRE(grad_off % 16 == 0 && param_off % 16 == 0 && count % 4 == 0, "allreduce_sgd: offsets must be 16-byte aligned and count a multiple of 4 floats");
  TADNN_REQUIRE(grad_off + count * 4 <= c->heap_bytes && param_off + count * 4 <= c->heap_bytes, "allreduce_sgd: range exceeds the heap");
  TADNN_REQUIRE(grad_off + count * 4 <= param_off || param_off + count * 4 <= grad_off, "allreduce_sgd: gradient and parameter ranges overlap");
  if (count == 0) return 0;
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  const size_t nvec = count / 4;
  const bool has_mc = c->dev.mc_heap != nullptr;
  if (algo == 0) algo = (count * 4 * static_cast<unsigned long long>(c->world - 1) <= (512u << 10)) ? 1 : (has_mc ? 3 : 2);
  TADNN_REQUIRE(algo >= 1 && algo <= 3 && (algo != 3 || has_mc), "allreduce_sgd: algo %d unavailable on this context", algo);
  if (algo == 1) {
    const int blocks = pick_blocks(nvec, max_blocks, 64);
    allreduce_sgd_oneshot_kernel<<<blocks, 512, 0, stream>>>(c->dev, grad_off, param_off, nvec, scale, lr, weight_decay, channel);
    TADNN_LAUNCH_CHECK("allreduce_sgd_oneshot");
  } else if (algo == 3) {
    const int blocks = pick_blocks(nvec / c->world / 4 + 1, max_blocks, 64);
    allreduce_sgd_nvls_kernel<<<blocks, 512, 0, stream>>>(c->dev, grad_off, param_off, nvec, scale, lr, weight_decay, channel);
    TADNN_LAUNCH_CHECK("allreduce_sgd_nvls");
  } else {
    const int blocks = pick_blocks(nvec / c->world + 1, max_blocks, 128);
    allreduce_sgd_twoshot_kernel<<<blocks, 512, 0, stream>>>(c->dev, grad_off, param_off, nvec, scale, lr, weight_decay, channel);
    TADNN_LAUNCH_CHECK("allreduce_sgd_twoshot");
  }
  return 0;
}

extern "C" int tadnn_allreduce_sgd_emulated(tadnn_comm* const* comms, int world, unsigned long long grad_off, unsigned long long param_off,
                                            unsigned long long count, float scale, float lr, float weight_decay, int algo, tadnn_stream_t stream_) {
  TADNN_REQUIRE(comms && world >= 1 && world <= kMaxRanks, "allreduce_sgd_emulated: bad arguments");
  TADNN_REQUIRE(algo == 1 || algo == 2, "allreduce_sgd_emulated: algo must be 1 (one-shot) or 2 (two-shot)");
  TADNN_REQUIRE(grad_off % 16 == 0 && param_off % 16 == 0 && count % 4 == 0, "allreduce_sgd_emulated: alignment");
  CommDev dev = comms[0]->dev;
  dev.rank = -1;
  dev.world = world;
  const size_t nvec = count / 4;
  const size_t g = grad_off, p = param_off;
  if (algo == 1) return launch_maybe_coop(dev, allreduce_sgd_oneshot_kernel, pick_blocks(nvec, 8, 8), 512, static_cast<cudaStream_t>(stream_), "allreduce_sgd_oneshot", dev, g, p, nvec, scale, lr, weight_decay, 0);
  return launch_maybe_coop(dev, allreduce_sgd_twoshot_kernel, pick_blocks(nvec / world + 1, 8, 8), 512, static_cast<cudaStream_t>(stream_), "allreduce_sgd_twoshot", dev, g, p, nvec, scale, lr, weight_decay, 0);
}

extern "C" int tadnn_allreduce_ll(tadnn_comm* c, unsigned long long ll_off, unsigned long long off_bytes, unsigned long long param_off,
                                  unsigned long long count, float scale, int sgd, float lr, float weight_decay, unsigned int epoch,
                                  tadnn_stream_t stream_) {
  if (int rc = validate(c, 0, "allreduce_ll")) return rc;
  TADNN_REQUIRE(count > 0 && count <= static_cast<unsigned long long>(kLLMaxFloats), "allreduce_ll: %llu floats exceed the %d the staging area holds", count, kLLMaxFloats);
  TADNN_REQUIRE(epoch != 0, "allreduce_ll: epoch 0 is reserved (freshly zeroed staging)");
  const unsigned long long need = 2ull * c->world * kLLMaxFloats * 8ull;
  TADNN_REQUIRE(ll_off % 8 == 0 && ll_off + need <= c->heap_bytes && off_bytes % 4 == 0 && off_bytes + count * 4 <= c->heap_bytes,
                "allreduce_ll: staging area or data range outside the heap");
  const int n = static_cast<int>(count);
  int blocks = (n + 511) / 512;
  if (blocks > 64) blocks = 64;
  const size_t zero = 0;
  if (sgd) TADNN_CHECK_CUDA(launch_pdl(allreduce_ll_kernel<1>, dim3(blocks), dim3(512), 0, static_cast<cudaStream_t>(stream_), c->dev, static_cast<size_t>(ll_off),
                                       static_cast<size_t>(off_bytes), static_cast<size_t>(param_off), n, scale, lr, weight_decay, epoch));
  else TADNN_CHECK_CUDA(launch_pdl(allreduce_ll_kernel<0>, dim3(blocks), dim3(512), 0, static_cast<cudaStream_t>(stream_), c->dev, static_cast<size_t>(ll_off),
                                   static_cast<size_t>(off_bytes), zero, n, scale, 0.f, 0.f, epoch));
  TADNN_LAUNCH_CHECK(sgd ? "allreduce_sgd_ll" : "allreduce_ll");
  return 0;
}

extern "C" int tadnn_allreduce_ll_emulated(tadnn_comm* const* comms, int world, unsigned long long ll_off, unsigned long long off_bytes,
                                           unsigned long long param_off, unsigned long long count, float scale, int sgd, float lr,
                                           float weight_decay, unsigned int epoch, tadnn_stream_t stream_) {
  TADNN_REQUIRE(comms && world >= 1 && world <= kMaxRanks && count > 0 && count <= static_cast<unsigned long long>(kLLMaxFloats) && epoch != 0,
                "allreduce_ll_emulated: bad arguments");
  CommDev dev = comms[0]->dev;
  dev.rank = -1;
  dev.world = world;
  const int n = static_cast<int>(count);
  int blocks = (n + 511) / 512;
  if (blocks > 8) blocks = 8;
  const size_t lo = ll_off, d = off_bytes, po = param_off;
  if (sgd) return launch_maybe_coop(dev, allreduce_ll_kernel<1>, blocks, 512, static_cast<cudaStream_t>(stream_), "allreduce_sgd_ll", dev, lo, d, po, n, scale, lr, weight_decay, epoch);
  return launch_maybe_coop(dev, allreduce_ll_kernel<0>, blocks, 512, static_cast<cudaStream_t>(stream_), "allreduce_ll", dev, lo, d, po, n, scale, lr, weight_decay, epoch);
}

extern "C" int tadnn_allreduce_emulated(tadnn_comm* const* comms, int world, unsigned long long off_bytes, unsigned long long count, int dtype, float scale, int algo, tadnn_stream_t stream) {
  TADNN_REQUIRE(comms && world >= 1 && world <= kMaxRanks, "allreduce_emulated: bad arguments");
  TADNN_REQUIRE(algo == 1 || algo == 2, "allreduce_emulated: algo must be 1 (one-shot) or 2 (two-shot)");
  CommDev dev = comms[0]->dev;
  dev.rank = -1;
  dev.world = world;
  return do_allreduce(dev, false, off_bytes, count, dtype, scale, algo, 8, 0, static_cast<cudaStream_t>(stream));
}

extern "C" int tadnn_broadcast(tadnn_comm* c, unsigned long long off_bytes, unsigned long long bytes, int root, int max_blocks, int channel, tadnn_stream_t stream) {
  if (int rc = validate(c, channel, "broadcast")) return rc;
  TADNN_REQUIRE(root >= 0 && root < c->world, "broadcast: root %d out of range", root);
  TADNN_REQUIRE(off_bytes % 16 == 0, "broadcast: offset not 16-byte aligned");
  TADNN_REQUIRE(off_bytes + bytes <= c->heap_bytes, "broadcast: range exceeds the heap");
  if (c->world == 1 || bytes == 0) return 0;
  const size_t nvec = bytes / 16;
  const int tail = static_cast<int>(bytes - nvec * 16);
  const int blocks = pick_blocks(nvec / 4 + 1, max_blocks, 64);
  broadcast_kernel<<<blocks, 512, 0, static_cast<cudaStream_t>(stream)>>>(c->dev, off_bytes, nvec, off_bytes + nvec * 16, tail, root, channel);
  TADNN_LAUNCH_CHECK("broadcast_flat");
  return 0;
}

extern "C" int tadnn_allgather_cols(tadnn_comm* c, const void* src, unsigned long long src_pitch, unsigned long long off_bytes, int rows, unsigned long long width_bytes, int max_blocks, int channel, tadnn_stream_t stream_) {
  if (int rc = validate(c, channel, "allgather_cols")) return rc;
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  TADNN_REQUIRE(rows > 0 && width_bytes > 0, "allgather_cols: empty");
  TADNN_REQUIRE(off_bytes + static_cast<unsigned long long>(rows) * width_bytes * c->world <= c->heap_bytes, "allgather_cols: output exceeds the heap");
  const uintptr_t a = reinterpret_cast<uintptr_t>(src) | src_pitch | off_bytes | width_bytes;
  const uint8_t* s = static_cast<const uint8_t*>(src);
  if (a % 16 == 0) {
    const size_t wv = width_bytes / 16;
    allgather_cols_kernel<uint4><<<pick_blocks(rows * wv / 2 + 1, max_blocks, 64), 512, 0, stream>>>(c->dev, s, src_pitch, off_bytes, rows, wv, channel);
  } else if (a % 4 == 0) {
    const size_t wv = width_bytes / 4;
    allgather_cols_kernel<uint32_t><<<pick_blocks(rows * wv / 2 + 1, max_blocks, 64), 512, 0, stream>>>(c->dev, s, src_pitch, off_bytes, rows, wv, channel);
  } else {
    TADNN_REQUIRE(a % 2 == 0, "allgather_cols: operands must be at least 2-byte aligned");
    const size_t wv = width_bytes / 2;
    allgather_cols_kernel<uint16_t><<<pick_blocks(rows * wv / 2 + 1, max_blocks, 64), 512, 0, stream>>>(c->dev, s, src_pitch, off_bytes, rows, wv, channel);
  }
  TADNN_LAUNCH_CHECK("allgather_cols");
  return 0;
}

extern "C" int tadnn_reduce_scatter_cols(tadnn_comm* c, unsigned long long off_bytes, void* dst, unsigned long long dst_pitch, int rows, int width, int dtype, float scale, const float* bias, int act, int max_blocks, int channel, tadnn_stream_t stream_) {
  if (int rc = validate(c, channel, "reduce_scatter_cols")) return rc;
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  TADNN_REQUIRE(dtype == 0 || dtype == 1, "reduce_scatter_cols: dtype must be 0 or 1");
  TADNN_REQUIRE(rows > 0 && width > 0, "reduce_scatter_cols: empty");
  const unsigned long long es = dtype == 0 ? 4 : 2;
  TADNN_REQUIRE(off_bytes + static_cast<unsigned long long>(rows) * width * c->world * es <= c->heap_bytes, "reduce_scatter_cols: input exceeds the heap");
  const int blocks = pick_blocks(static_cast<size_t>(rows) * width / 2 + 1, max_blocks, 128);
  if (dtype == 0)
    reduce_scatter_cols_kernel<0><<<blocks, 512, 0, stream>>>(c->dev, off_bytes, dst, dst_pitch, rows, width, scale, bias, act, channel);
  else
    reduce_scatter_cols_kernel<1><<<blocks, 512, 0, stream>>>(c->dev, off_bytes, dst, dst_pitch, rows, width, scale, bias, act, channel);
  TADNN_LAUNCH_CHECK("reduce_scatter_cols");
  return 0;
}

extern "C" int tadnn_reduce_cols_nvls(tadnn_comm* c, unsigned long long part_off, int rows, int width, float scale, const float* bias,
                                      int act, void* out_local, unsigned long long out_pitch, int bcast, int max_blocks, int channel,
                                      tadnn_stream_t stream) {
  if (int rc = validate(c, channel, "reduce_cols_nvls")) return rc;
  TADNN_REQUIRE(c->dev.mc_heap != nullptr, "reduce_cols_nvls: this context has no multicast mapping (NVLS unavailable)");
  TADNN_REQUIRE(rows > 0 && width > 0 && width % 8 == 0, "reduce_cols_nvls: width must be a positive multiple of 8");
  TADNN_REQUIRE(part_off % 16 == 0, "reduce_cols_nvls: offset must be 16-byte aligned");
  TADNN_REQUIRE(part_off + 2ull * rows * width * c->world <= c->heap_bytes, "reduce_cols_nvls: partials exceed the heap");
  TADNN_REQUIRE(out_local == nullptr || ((reinterpret_cast<uintptr_t>(out_local) & 15) == 0 && out_pitch % 8 == 0),
                "reduce_cols_nvls: out_local must be 16-byte aligned with a pitch multiple of 8");
  const int blocks = pick_blocks(static_cast<size_t>(rows) * (width / 8), max_blocks, 148);
  reduce_cols_nvls_kernel<<<blocks, 512, 0, static_cast<cudaStream_t>(stream)>>>(
      c->dev, part_off, rows, width / 8, scale, bias, act, static_cast<__nv_bfloat16*>(out_local), out_pitch, bcast, channel);
  TADNN_LAUNCH_CHECK("reduce_cols_nvls");
  return 0;
}

extern "C" int tadnn_reduce_slots(tadnn_comm* c, unsigned long long slots_off, int rows, int width, float scale,
                                  const float* bias, int act, void* out_local, unsigned long long out_pitch,
                                  unsigned long long push_off, int do_push, int slab_major, int max_blocks, int channel,
                                  tadnn_stream_t stream) {
  if (int rc = validate(c, channel, "reduce_slots")) return rc;
  TADNN_REQUIRE(rows > 0 && width > 0 && width % 8 == 0, "reduce_slots: width must be a positive multiple of 8");
  TADNN_REQUIRE(slots_off % 16 == 0 && (!do_push || push_off % 16 == 0), "reduce_slots: offsets must be 16-byte aligned");
  const unsigned long long rows_pad = slab_major ? ((static_cast<unsigned long long>(rows) + 31) / 32) * 32 : rows;
  TADNN_REQUIRE(!slab_major || width % 64 == 0, "reduce_slots: slab-major staging needs width %% 64 == 0");
  TADNN_REQUIRE(slots_off + 2ull * rows_pad * width * c->world <= c->heap_bytes, "reduce_slots: slots exceed the heap");
  TADNN_REQUIRE(out_local == nullptr || ((reinterpret_cast<uintptr_t>(out_local) & 15) == 0 && out_pitch % 8 == 0),
                "reduce_slots: out_local must be 16-byte aligned with a pitch multiple of 8");
  const int blocks = pick_blocks(static_cast<size_t>(rows) * (width / 8), max_blocks, 148);
  if (slab_major)
    reduce_slots_slab_kernel<<<blocks, 512, 0, static_cast<cudaStream_t>(stream)>>>(
        c->dev, slots_off, rows, width / 8, scale, bias, act, static_cast<__nv_bfloat16*>(out_local), out_pitch, push_off,
        do_push, channel);
  else
    reduce_slots_kernel<<<blocks, 512, 0, static_cast<cudaStream_t>(stream)>>>(
        c->dev, slots_off, rows, width / 8, scale, bias, act, static_cast<__nv_bfloat16*>(out_local), out_pitch, push_off,
        do_push, channel);
  TADNN_LAUNCH_CHECK("reduce_slots");
  return 0;
}
