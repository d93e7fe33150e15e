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

/* ---- symmetric-memory communication context (csrc/comm.cu) -----------------
 * One process per GPU.  Every rank allocates the same-sized heap, exchanges a 64-byte
 * handle with its peers (any out-of-band channel; tadnn uses torch.distributed) and maps
 * the peers' heaps.  Kernels then address peer memory directly over NVLink/NVSwitch.
 * Replaces the TorchMPI C++ core the reference links against (communicators, CUDA-IPC
 * rings, NCCL-1 wrappers; reached from datamodule.lua:25-33,50,199-224 and
 * nodemodule.lua:52,66,103,117,166,266).
 *
 * The heap is followed by a signal pad (flag words for the device-side barrier).
 * `channel` (0..TADNN_COMM_CHANNELS-1) selects an independent set of flags: collectives
 * that may be in flight at the same time (different streams) must use different channels.
 */
#define TADNN_COMM_MAX_RANKS 8
#define TADNN_COMM_CHANNELS 8
#define TADNN_COMM_MAX_BLOCKS 512
#define TADNN_IPC_HANDLE_BYTES 64

typedef struct tadnn_comm tadnn_comm;

int tadnn_comm_create(int rank, int world, unsigned long long heap_bytes, tadnn_comm** out);
/* Single-process variant for tests on ONE GPU: `world` heaps in this process, no IPC.
 * Kernels launched on such a context must not wait on other launches; use the *_emulated
 * entry points or world-as-data kernels only. */
int tadnn_comm_create_local(int world, unsigned long long heap_bytes, tadnn_comm** out /* [world] */);
int tadnn_comm_ipc_handle(tadnn_comm* c, void* host_out64);
int tadnn_comm_open_peers(tadnn_comm* c, const void* host_all_handles /* world x 64 B */);
/* Attach externally mapped peer heaps (e.g. torch symmetric memory, or a multicast VA). */
int tadnn_comm_attach(tadnn_comm* c, void* const* host_peer_heaps, void* const* host_peer_pads,
                      void* multicast_heap);
int tadnn_comm_heap(tadnn_comm* c, int peer, void** out_ptr);
int tadnn_comm_info(tadnn_comm* c, int* rank, int* world, unsigned long long* heap_bytes, int* has_multicast);
int tadnn_comm_destroy(tadnn_comm* c);

/* K5 device barrier across all ranks (signal-pad flags, release/acquire at system scope).
 * Replaces mpi.barrier() (datamodule.lua:50,283). */
int tadnn_barrier(tadnn_comm* c, int channel, tadnn_stream_t stream);

/* K1/K2/K3 fused all-reduce: heap[off .. off+count) = scale * sum_over_ranks, in place, on
 * every rank; dtype 0 = fp32, 1 = bf16 (fp32 accumulation, fixed rank order => bit-identical
 * results on all ranks).  algo 0 = auto (size-based), 1 = one-shot, 2 = two-shot,
 * 3 = NVLS multimem (needs a multicast mapping).  Replaces the per-tensor
 * allreduceTensor + div(size) loop of synchronizeModel (datamodule.lua:211-224). */
int tadnn_allreduce(tadnn_comm* c, unsigned long long off_bytes, unsigned long long count, int dtype,
                    float scale, int algo, int max_blocks, int channel, tadnn_stream_t stream);

/* K4 broadcast of heap[off .. off+bytes) from `root` to every rank (push).  Replaces
 * mpinn.synchronizeParameters (datamodule.lua:33). */
int tadnn_broadcast(tadnn_comm* c, unsigned long long off_bytes, unsigned long long bytes, int root,
                    int max_blocks, int channel, tadnn_stream_t stream);

/* K8 all-gather along the last dim (push): every rank stores its local src[rows, width_bytes]
 * into heap[off] of ALL ranks, viewed as [rows, world*width_bytes], at column rank*width_bytes.
 * src is any local device pointer with row pitch src_pitch_bytes.  Replaces
 * mpi.allgatherTensor (nodemodule.lua:166,266). */
int tadnn_allgather_cols(tadnn_comm* c, const void* src, unsigned long long src_pitch_bytes,
                         unsigned long long off_bytes, int rows, unsigned long long width_bytes,
                         int max_blocks, int channel, tadnn_stream_t stream);

/* K7 reduce-scatter along the last dim (pull): dst[rows, width] (local, fp32 or bf16) =
 * scale * sum over ranks of heap_r[off] viewed as [rows, world*width] at column rank*width.
 * Optional fused epilogue: + bias[rank*width + col] (fp32, may be NULL), then act (0/1 tanh/2 relu). */
int tadnn_reduce_scatter_cols(tadnn_comm* c, unsigned long long off_bytes, void* dst,
                              unsigned long long dst_pitch_elems, int rows, int width, int dtype,
                              float scale, const float* bias, int act, int max_blocks, int channel,
                              tadnn_stream_t stream);

/* Emulated (single launch, no cross-launch waiting) variants for 1-GPU numerics tests on a
 * tadnn_comm_create_local() set: performs rank `as_rank`'s share of the math with all
 * barriers skipped. */
int tadnn_allreduce_emulated(tadnn_comm* const* comms, int world, unsigned long long off_bytes,
                             unsigned long long count, int dtype, float scale, int algo,
                             tadnn_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* TADNN_H_ */
