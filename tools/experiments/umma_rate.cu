// Cycles per tcgen05.mma (M = 128, K = 16, bf16) as a function of N, of where A comes from (shared memory / tensor
// memory) and of how many independent accumulators the issue stream rotates over.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I csrc -o /tmp/umma_rate tools/experiments/umma_rate.cu && /tmp/umma_rate
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
#include "../../csrc/sm100.cuh"
using namespace tadnn::sm100;

__device__ __forceinline__ uint64_t desc_nosw(uint32_t a, uint32_t lbo, uint32_t sbo) {
  return (uint64_t)((a >> 4) & 0x3FFF) | ((uint64_t)((lbo >> 4) & 0x3FFF) << 16) | ((uint64_t)((sbo >> 4) & 0x3FFF) << 32) | (1ull << 46);
}
__device__ __forceinline__ void umma_ts(uint32_t d, uint32_t a, uint64_t b, uint32_t idesc, uint32_t acc) {
  asm volatile("{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\ntcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n}\n" ::"r"(d), "r"(a), "l"(b), "r"(idesc), "r"(acc) : "memory");
}

// MODE 0: A from smem (SWIZZLE_128B K-major), 1: A from TMEM.  CHAINS: accumulators rotated over.  bsw: B swizzled (1) or no-swizzle (0)
template <int MODE, int CHAINS>
__global__ void __launch_bounds__(128, 1) rate_kernel(int N, int bsw, int rounds, long long* out) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
  __shared__ __align__(8) uint64_t bar;
  __shared__ uint32_t tbase;
  for (int i = threadIdx.x; i < 64 * 1024 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0x3c003c00u;
  if (threadIdx.x == 0) { mbar_init(&bar, 1); fence_mbar_init(); }
  if (threadIdx.x < 32) tmem_alloc<512>(&tbase);
  fence_proxy_async_smem();
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  const uint32_t tb = tbase;
  if (threadIdx.x < 32) {
    const uint32_t idesc = make_idesc_bf16(128, N, 0, bsw == 2 ? 1 : 0);
    const uint32_t sa = smem_u32(smem), sb = smem_u32(smem + 16384);
    const uint64_t adesc = make_smem_desc_sw128(sa, 16, 1024);
    // bsw 0: K-major no-swizzle (rows 16 B apart), 1: K-major SWIZZLE_128B, 2: MN-major no-swizzle (8-K groups 128 B, 8-N groups 256 B apart)
    const uint64_t bdesc = bsw == 1 ? make_smem_desc_sw128(sb, 16, 1024) : (bsw == 2 ? desc_nosw(sb, 128, 256) : desc_nosw(sb, 16, 128));
    constexpr int stride = 512 / (CHAINS + (MODE == 1 ? 1 : 0)) / 32 * 32;   // accumulator column stride; last slot = A when in TMEM
    const uint32_t ta = tb + CHAINS * stride;
    long long t0 = 0, t1 = 0;
    if (elect_one()) {
      t0 = clock64();
#pragma unroll 1
      for (int r = 0; r < rounds; r += 8) {
#pragma unroll
        for (int u = 0; u < 8; ++u)
#pragma unroll
          for (int c = 0; c < CHAINS; ++c) {
            const uint32_t d = tb + c * stride;
            if (MODE == 0) umma_f16(d, adesc + (u & 3) * 2, bdesc + u, idesc, 1u);
            else umma_ts(d, ta + (u & 3) * 8, bdesc + u, idesc, 1u);
          }
      }
      umma_commit(&bar);
      t1 = clock64();
    }
    __syncwarp();
    mbar_wait(&bar, 0);
    const long long t2 = clock64();
    if (blockIdx.x == 0 && t0 != 0) { out[0] = t1 - t0; out[1] = t2 - t0; }
  }
  tc_fence_before_sync();
  __syncthreads();
  if (threadIdx.x < 32) { tc_fence_after_sync(); tmem_dealloc<512>(tb); }
}

template <int MODE, int CHAINS>
void run(int N, int bsw, long long* out) {
  if ((CHAINS + MODE) * N > 512) return;
  const int total = 4096, rounds = total / CHAINS;
  cudaFuncSetAttribute(rate_kernel<MODE, CHAINS>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
  for (int rep = 0; rep < 2; ++rep) {
    rate_kernel<MODE, CHAINS><<<148, 128, 100 * 1024>>>(N, bsw, rounds, out);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); exit(1); }
  }
  printf("%-4s %-5d %-6d %-4d %10.1f %10.1f\n", MODE ? "tmem" : "smem", N, CHAINS, bsw, (double)out[0] / total, (double)out[1] / total);
}

int main() {
  long long* out;
  cudaMallocManaged(&out, 16);
  printf("%-4s %-5s %-6s %-4s %10s %10s\n", "A", "N", "chains", "Bsw", "issue/mma", "cyc/mma");
  for (int bsw = 0; bsw < 3; ++bsw)
    for (int N : {48, 96, 128, 192, 256}) {
      run<0, 1>(N, bsw, out); run<0, 2>(N, bsw, out); run<0, 4>(N, bsw, out);
      run<1, 1>(N, bsw, out); run<1, 2>(N, bsw, out); run<1, 4>(N, bsw, out);
    }
  return 0;
}
