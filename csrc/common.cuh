// tadnn native core — shared host/device helpers.
// B200 (sm_100a) only. No multi-arch dispatch.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include <atomic>

#include "../include/tadnn.h"

namespace tadnn {

// ---- error plumbing ---------------------------------------------------------
// Every extern "C" entry point returns 0 on success, non-zero on failure and
// leaves a human-readable message retrievable with tadnn_last_error().
void set_error(const char* fmt, ...);
int  note_launch(const char* kernel_name);   // bumps the launch counter; returns 0

#define TADNN_CHECK_CUDA(expr)                                                   \
  do {                                                                           \
    cudaError_t _e = (expr);                                                     \
    if (_e != cudaSuccess) {                                                     \
      ::tadnn::set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #expr,           \
                         cudaGetErrorString(_e));                                \
      return 1;                                                                  \
    }                                                                            \
  } while (0)

#define TADNN_REQUIRE(cond, ...)                                                 \
  do {                                                                           \
    if (!(cond)) {                                                               \
      ::tadnn::set_error(__VA_ARGS__);                                           \
      return 2;                                                                  \
    }                                                                            \
  } while (0)

#define TADNN_LAUNCH_CHECK(name)                                                 \
  do {                                                                           \
    cudaError_t _e = cudaGetLastError();                                         \
    if (_e != cudaSuccess) {                                                     \
      ::tadnn::set_error("launch of %s failed: %s", name, cudaGetErrorString(_e)); \
      return 3;                                                                  \
    }                                                                            \
    ::tadnn::note_launch(name);                                                  \
  } while (0)

inline int ceil_div(long long a, long long b) { return (int)((a + b - 1) / b); }

int sm_count();  // cached multiProcessorCount of the current device

// Tiled bf16 tensor map of rank 2..5, SWIZZLE_128B; dimensions innermost-first (core.cu).
int encode_tmap_bf16(CUtensorMap* out, const void* ptr, int rank, const unsigned long long* gdim,
                     const unsigned long long* gstride_bytes, const unsigned int* box);

// Opt a kernel into more than 48 KB of dynamic shared memory, once per (call site, device): `done` is a
// call-site static holding one bit per device ordinal, so a process that drives several GPUs sets the
// attribute on each of them, and concurrent first calls are harmless (the attribute is idempotent).
template <typename K>
inline int ensure_dyn_smem(K kern, int bytes, std::atomic<unsigned long long>& done) {
  int dev = 0;
  TADNN_CHECK_CUDA(cudaGetDevice(&dev));
  const unsigned long long bit = 1ull << (dev & 63);
  if (!(done.load(std::memory_order_acquire) & bit)) {
    TADNN_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes));
    done.fetch_or(bit, std::memory_order_release);
  }
  return 0;
}

// Launch `kern` so that it may overlap the tail of the previous kernel in the stream (programmatic dependent launch):
// ONLY for kernels that call sm100::pdl_wait() before touching memory written by earlier kernels.  Falls back to a normal
// launch when TADNN_PDL=0 is set in the environment (TADNN_PDL_GRAPH=0: only inside CUDA-graph capture).
bool pdl_enabled();
bool pdl_in_graphs();   // the attribute is kept while a stream is being captured (programmatic edges in the graph) unless TADNN_PDL_GRAPH=0
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream, Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
  if (pdl_enabled() && cudaStreamIsCapturing(stream, &cap) == cudaSuccess && (cap == cudaStreamCaptureStatusNone || pdl_in_graphs())) {
    cfg.attrs = attr;
    cfg.numAttrs = 1;
  }
  return cudaLaunchKernelEx(&cfg, kern, KArgs(args)...);
}

// ---- device helpers ---------------------------------------------------------
__device__ __forceinline__ float bf16_to_f32(__nv_bfloat16 v) { return __bfloat162float(v); }

__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
  __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&v);
}
__device__ __forceinline__ float2 unpack_bf16x2(uint32_t u) {
  __nv_bfloat162 v = *reinterpret_cast<__nv_bfloat162*>(&u);
  return __bfloat1622float2(v);
}

// Accurate-enough tanh for the epilogues: tanh(x) = 1 - 2/(exp(2x)+1).
// (tanh.approx.f32 has ~1e-3 abs error near 0.5 which is visible in fp32 tests;
//  this form is ~1e-6 and still two SFU ops.)
__device__ __forceinline__ float tanh_f32(float x) {
  float e = __expf(2.0f * x);
  return 1.0f - __fdividef(2.0f, e + 1.0f);
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

}  // namespace tadnn
