/* tadnn — C ABI of the B200-native core (libtadnn.so).
 *
 * Every entry point returns 0 on success and non-zero on failure; the failure
 * text is retrievable with tadnn_last_error().  Pointers are device pointers
 * unless a comment says "host".  All kernels are compiled for sm_100a only.
 *
 * Reference citations are file:line into the upstream Lua tree
 * (ngrabaskas/Torch-Automatic-Distributed-Neural-Network); the reference ships no
 * native code, so each entry point names the Lua call site whose library call
 * (THNN / cuBLAS / TorchMPI) it replaces.
 */
#ifndef TADNN_H_
#define TADNN_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef void* tadnn_stream_t; /* cudaStream_t */

/* ---- library-wide -------------------------------------------------------- */
const char* tadnn_last_error(void);
int tadnn_version(void);
/* Number of tadnn kernels launched by this process since load (or since the last
 * tadnn_reset_launch_count); bench.py reports it as "gpu_launches". */
long long tadnn_launch_count(void);
void tadnn_reset_launch_count(void);
/* Copies up to `cap` bytes of a '\n'-separated "name count" table (host buffer). */
int tadnn_launch_table(char* host_buf, int cap);

/* ---- tcgen05 GEMM (csrc/gemm_tcgen05.cu) ---------------------------------
 * C[M,N] = act(alpha * A·Bᵀ + bias), bf16 operands, fp32 accumulation in TMEM.
 * Replaces the cuBLAS SGEMM behind nn.Linear.updateOutput / updateGradInput /
 * accGradParameters (nodemodule.lua:47,65,72,95,116) and behind
 * nn.SpatialConvolutionMM (cifar_example/sgd-torchad_nn-cifar.lua:91,96).
 *
 * Operand storage ("major"):
 *   a_mn_major = 0: A stored [M, K] row-major (K contiguous), lda = row pitch (elements)
 *   a_mn_major = 1: A stored [K, M] row-major (M contiguous), lda = row pitch
 *   b_mn_major = 0: B stored [N, K] row-major;  b_mn_major = 1: B stored [K, N] row-major
 * So   forward  y = x·Wᵀ   is (0,0);  dgrad  dx = g·W  is (0,1);  wgrad  dW = gᵀ·x  is (1,1).
 * Pitches must be multiples of 8 elements and base pointers 16-byte aligned (TMA).
 */
typedef struct tadnn_gemm_args {
  int M, N, K;
  const void* A;
  int lda;
  int a_mn_major;
  const void* B;
  int ldb;
  int b_mn_major;
  void* C;
  int ldc;
  int c_fp32;        /* 0: C is bf16, 1: C is fp32 */
  const float* bias; /* [N] fp32 or NULL */
  int act;           /* 0 none, 1 tanh, 2 relu */
  float alpha;
  int accumulate;    /* 1: C += result (fp32 C only; implemented with red.global.add) */
  int split_k;       /* 0/1: none; >1: split the K loop over that many CTAs (fp32 C, atomics) */
  /* Fused all-gather: every finished output tile is additionally stored to these
   * buffers (peer-mapped pointers), same dtype as C, at [row, push_col_offset + col]. */
  int n_push;
  void* push[8];
  int push_ldc;
  int push_col_offset;
  int tile_n;        /* 0: auto; else 64, 128 or 256 */
} tadnn_gemm_args;

int tadnn_gemm_bf16(const tadnn_gemm_args* args, tadnn_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* TADNN_H_ */
