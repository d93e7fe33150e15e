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
  /* Fused reduce-scatter push (K-split partial products): rs_width > 0 means output column
   * block q = col / rs_width is stored ONLY to push[q] (this rank's slot in rank q's staging
   * buffer, [M, push_ldc] bf16) at column col - q*rs_width; C is not written. */
  int rs_width;
  /* Fused tanh backward: multiply the result by (1 - y^2), y = aux_tanh_out[M, ld_aux] bf16. */
  const void* aux_tanh_out;
  int ld_aux;
  /* Start the tile sweep at this n-block (wraps around).  With rs_width each rank passes a
   * different value so that the ranks never all push to the same owner at the same time. */
  int n_rotate;
  /* With rs_width: store each 32-row x 64-column slab as one contiguous 4 KB block (slab-major
   * staging, rows padded to a multiple of 32; slab s = (row/32)*(rs_width/64) + col/64, contents in
   * the SWIZZLE_128B order of the staging tile) -- tadnn_reduce_slots(slab_major=1) reads it. */
  int rs_slab_major;
  /* Fused all-gather -> GEMM (K13: the backward of the model-parallel linears, nodemodule.lua:56-68,107-119,
   * 165-167).  PRODUCER side (n_push > 0, bf16 TMA-store epilogue): after the 64-column slabs of a 32-row block
   * have landed in every push target, the kernel adds their number to flag word
   * sig_flags[q][(row/32)*flag_stride + sig_src] of each target q (release, system scope).  CONSUMER side: the A
   * operand is such a gathered matrix whose column block c belongs to source rank c / wait_cols_per_src; the
   * TMA producer warp spins (acquire, system scope) until wait_flags[(row/32)*flag_stride + src] >= wait_target
   * before it loads a tile of those rows from that source.  No separate all-gather kernel and no rank barrier:
   * tiles are consumed as they arrive.  Flags only ever grow; the caller keeps the running target. */
  void* sig_flags[8];
  int n_sig;
  int sig_src;
  int flag_stride;
  const void* wait_flags;
  unsigned int wait_target;
  int wait_cols_per_src;
  /* Structurally sparse B (e.g. the Toeplitz matrix of a transposed convolution): HOST array with one word per column
   * tile of width tile_n (which must be given); bit kb clear = the 64-deep k-block kb of that column tile is all zeros
   * and is neither loaded nor multiplied.  <= 32 column tiles, <= 32 k-blocks, no split-K.  NULL: dense. */
  const unsigned int* kblock_mask;
  int n_kblock_mask;
} tadnn_gemm_args;

int tadnn_gemm_bf16(const tadnn_gemm_args* args, tadnn_stream_t stream);

/* 2-CTA variant (csrc/gemm_2cta.cu): thread-block clusters of two CTAs, tcgen05.mma.cta_group::2,
 * 256x256 tiles, each CTA stages its 128 A rows and half of the B tile.  K-major bf16 operands,
 * bf16 output, bias/act/alpha epilogue.  Same call sites as tadnn_gemm_bf16 (nodemodule.lua:47,95). */
int tadnn_gemm_bf16_2cta(const void* A, int lda, const void* B, int ldb, void* C, int ldc, int M, int N,
                         int K, const float* bias, int act, float alpha, tadnn_stream_t stream);

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
#define TADNN_LL_MAX_FLOATS 131072 /* largest message of the low-latency all-reduce (512 KB) */

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
/* Failure detection.  The reference hangs for ever when a rank dies (every MPI call it makes blocks:
 * datamodule.lua:50,216,221; nodemodule.lua:52).  Here a device-side barrier that has waited `seconds`
 * (default 120, or TADNN_COMM_TIMEOUT_S) gives up WITHOUT trapping: it records which rank was waiting for
 * which peer, marks the context dead and lets the kernel run to completion (its output is garbage), and so
 * does every later collective on the context.  tadnn_comm_check returns non-zero once that has happened,
 * with a message naming the missing rank in tadnn_last_error(); it reads one word of mapped host memory and
 * is safe to call at any time, including while kernels are running. */
int tadnn_comm_set_timeout(tadnn_comm* c, double seconds);
int tadnn_comm_check(tadnn_comm* c);

/* VMM + multicast backend (csrc/comm_vmm.cu): the heap is a cuMemCreate allocation shared as a
 * POSIX file descriptor; where the driver exposes NVLS a multicast object over all heaps is mapped
 * and the multimem all-reduce (algo 3) becomes available.  Sequence per rank: create_vmm ->
 * export_fd (send the fd to every peer) -> import_peer_fd for every peer; optional multicast:
 * rank 0 mc_create (send fd) / others mc_import -> mc_add_device -> host barrier -> mc_bind.
 * Descriptor ownership: export_fd / mc_create return a descriptor the caller closes after sending it;
 * import_peer_fd / mc_import take ownership of theirs and close it whether or not the import succeeds.
 * A failed create_vmm leaves nothing allocated or mapped. */
int tadnn_multicast_supported(int* out);
int tadnn_comm_create_vmm(int rank, int world, unsigned long long heap_bytes, tadnn_comm** out);
int tadnn_comm_export_fd(tadnn_comm* c, int* fd);
int tadnn_comm_import_peer_fd(tadnn_comm* c, int peer, int fd);
int tadnn_comm_mc_create(tadnn_comm* c, int* fd);
int tadnn_comm_mc_import(tadnn_comm* c, int fd);
int tadnn_comm_mc_add_device(tadnn_comm* c);
int tadnn_comm_mc_bind(tadnn_comm* c);
int tadnn_comm_vmm_release(tadnn_comm* c);

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

/* Gradient all-reduce-average FUSED with the SGD step (fp32): for count floats,
 *   g = scale * sum_ranks heap_r[grad_off ..];  heap[param_off ..] -= lr * (g + weight_decay * p);  heap[grad_off ..] = 0
 * on every rank, in one kernel (one-shot / two-shot / NVLS as for tadnn_allreduce).  Replaces synchronizeModel
 * (datamodule.lua:211-224) plus the optimiser pass when the replicas hold identical parameters (synchronised
 * every step from a broadcast start): each rank then computes the same update in the same rank order, so the
 * parameters stay bit-identical without being communicated.  *_emulated: all ranks in one cooperative launch
 * on a tadnn_comm_create_local set (single-GPU tests). */
int tadnn_allreduce_sgd(tadnn_comm* c, unsigned long long grad_off_bytes, unsigned long long param_off_bytes,
                        unsigned long long count, float scale, float lr, float weight_decay, int algo,
                        int max_blocks, int channel, tadnn_stream_t stream);
int tadnn_allreduce_sgd_emulated(tadnn_comm* const* comms, int world, unsigned long long grad_off_bytes,
                                 unsigned long long param_off_bytes, unsigned long long count, float scale,
                                 float lr, float weight_decay, int algo, tadnn_stream_t stream);

/* Low-latency all-reduce for small fp32 messages (<= TADNN_LL_MAX_FLOATS): every rank writes {value, epoch}
 * pairs into its slot of every rank's staging area and sums its own staging area as the pairs arrive: one
 * NVLink one-way trip, no barrier.  ll_off: a zero-initialised staging area of
 * 2 * world * TADNN_LL_MAX_FLOATS * 8 bytes at the same heap offset on every rank; epoch: non-zero, +1 on every
 * call that uses this staging area (all ranks make the same calls in the same order).  sgd != 0: the fused form
 * of tadnn_allreduce_sgd (param_off = parameters, off = gradients, cleared).  This is the path the reference's
 * own workload lives on: one sample per step, 40-byte to 8 KB messages (nodemodule.lua:52,103). */
int tadnn_allreduce_ll(tadnn_comm* c, unsigned long long ll_off, unsigned long long off_bytes,
                       unsigned long long param_off, unsigned long long count, float scale, int sgd, float lr,
                       float weight_decay, unsigned int epoch, tadnn_stream_t stream);
int tadnn_allreduce_ll_emulated(tadnn_comm* const* comms, int world, unsigned long long ll_off,
                                unsigned long long off_bytes, unsigned long long param_off, unsigned long long count,
                                float scale, int sgd, float lr, float weight_decay, unsigned int epoch,
                                tadnn_stream_t stream);

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

/* Second half of the fused K-split GEMM -> reduce-scatter (K11): the peers' GEMM epilogues have
 * pushed their partial [rows, width] bf16 tiles into this rank's staging slots (heap[slots_off],
 * layout [world][rows][width]); this kernel barriers, sums the slots in fp32, scales, adds
 * bias[rank*width + col], applies act and writes bf16 to out_local (may be NULL) and, if do_push,
 * to heap[push_off] of EVERY rank viewed as [rows, world*width] (all-gather => all-reduce).
 * Replaces the allreduceTensor after nn.Linear.updateOutput (nodemodule.lua:52,103) and the
 * following Tanh (nodemodule.lua:134). */
int tadnn_reduce_slots(tadnn_comm* c, unsigned long long slots_off, int rows, int width, float scale,
                       const float* bias, int act, void* out_local, unsigned long long out_pitch,
                       unsigned long long push_off, int do_push, int slab_major, int max_blocks,
                       int channel, tadnn_stream_t stream);

/* NVLS form of the GEMM -> reduce-scatter finalize: every rank's GEMM wrote its full partial
 * [rows, world*width] bf16 at heap offset part_off; this rank reads the switch-reduced sum of ITS
 * column block (multimem.ld_reduce, fp32 accumulation), applies scale/bias/act and writes out_local
 * [rows, width]; bcast != 0 also multimem.st's the result into every rank's buffer (all-reduce).
 * Needs the VMM + multicast heap (tadnn_comm_mc_bind).  Replaces the blocking allreduceTensor of the
 * layer output (nodemodule.lua:49,97). */
int tadnn_reduce_cols_nvls(tadnn_comm* c, unsigned long long part_off, int rows, int width, float scale,
                           const float* bias, int act, void* out_local, unsigned long long out_pitch,
                           int bcast, int max_blocks, int channel, tadnn_stream_t stream);

/* Emulated (single launch, no cross-launch waiting) variants for 1-GPU numerics tests on a
 * tadnn_comm_create_local() set: performs rank `as_rank`'s share of the math with all
 * barriers skipped. */
int tadnn_allreduce_emulated(tadnn_comm* const* comms, int world, unsigned long long off_bytes,
                             unsigned long long count, int dtype, float scale, int algo,
                             tadnn_stream_t stream);

/* ---- CIFAR-CNN kernels (csrc/cnn.cu) ----------------------------------------
 * Replace the cunn library calls behind the example model
 * (cifar_example/sgd-torchad_nn-cifar.lua:88-118,237): SpatialConvolutionMM's im2col,
 * ReLU, SpatialMaxPooling, Dropout, the two nn.Linear, LogSoftMax and the criterion.
 * Activations are bf16; NHWC between layers; valid convolution, stride 1.
 */
/* col[row, (kh*KW+kw)*C + c] with row = (b*OH+oy)*OW+ox; Kpad >= C*KH*KW, multiple of 8,
 * zero padded.  nchw=1: x is [B,C,H,W] (the input image); nchw=0: x is [B,H,W,C], C%8==0. */
int tadnn_im2col(const void* x, void* col, int B, int C, int H, int W, int KH, int KW, int Kpad,
                 int nchw, tadnn_stream_t stream);
/* im2col (NHWC) whose rows put the win x win output pixels read by the following max-pool first
 * (rows [0, B*win*win), pixel-major per image), then all other pixels; only_window=1 writes just
 * the first block.  The backward pass uses only that block (all other gradients are zero). */
int tadnn_im2col_winfirst(const void* x, void* col, int B, int C, int H, int W, int KH, int KW, int win,
                          int only_window, tadnn_stream_t stream);
/* T[(y*FW+x)*C + c, (oy*win+ox)*O + co] = Wp[co, ((y-oy)*KW+(x-ox))*C + c] (0 outside the kernel),
 * FH = win+KH-1: turns the transposed convolution of a gradient that is non-zero only inside the
 * pooled window into one dense GEMM dX[B, FH*FW*C] = dYw[B, win*win*O] * T^T. */
int tadnn_conv_toeplitz_pack(const void* wp, void* T, int O, int C, int KH, int KW, int win,
                             tadnn_stream_t stream);
/* dst[B,H,W,C] = src[B,FH,FW,C] in the top-left corner, zeros elsewhere. */
int tadnn_expand_corner(const void* src, void* dst, int B, int H, int W, int FH, int FW, int C,
                        tadnn_stream_t stream);
/* dx[B,H,W,C] = col2im(dcol[B*OH*OW, C*KH*KW]) (gather form, no atomics). */
int tadnn_col2im(const void* dcol, void* dx, int B, int C, int H, int W, int KH, int KW,
                 tadnn_stream_t stream);
/* k x k / stride k max pooling, floor mode, NHWC; idx holds the argmax (one byte per element). */
int tadnn_maxpool_fwd(const void* x, void* y, void* idx, int B, int H, int W, int C, int k,
                      tadnn_stream_t stream);
/* Backward of ReLU followed by that pooling: dx = dy routed to the argmax where y > 0. */
/* tadnn_maxpool_bwd: plain max-pool backward (gradient to the arg-max); tadnn_maxpool_relu_bwd additionally
 * masks windows whose pooled value is <= 0 (a ReLU directly before the pool, folded in). */
int tadnn_maxpool_bwd(const void* dy, const void* y, const void* idx, void* dx, int B, int H, int W, int C, int k,
                      tadnn_stream_t stream);
int tadnn_maxpool_relu_bwd(const void* dy, const void* y, const void* idx, void* dx, int B, int H,
                           int W, int C, int k, tadnn_stream_t stream);
/* out[c] += sum_r x[r,c]  (bias gradient; x bf16 [R, ld], out fp32). */
int tadnn_colsum_bf16(const void* x, float* out, long long R, int C, int ld, tadnn_stream_t stream);
/* out[c] += sum_r dy[r,c] * (y[r,c] > 0): bias gradient of conv+ReLU+maxpool evaluated on the
 * pooled tensors (dy, y: bf16 [R, C] contiguous). */
int tadnn_colsum_relu_mask(const void* dy, const void* y, float* out, long long R, int C,
                           tadnn_stream_t stream);
/* Same, dy stored compactly as [B, FH, FW, C] (top-left corner of y's [B, H, W, C] grid). */
int tadnn_colsum_relu_mask_corner(const void* dy, const void* y, float* out, int B, int H, int W, int FH,
                                  int FW, int C, tadnn_stream_t stream);
/* Torch7 conv weight [O, C*KH*KW] fp32 (k = (c*KH+kh)*KW+kw)  <->  GEMM operand [O, Kpad] bf16
 * (k' = (kh*KW+kw)*C+c); unpack accumulates: grad += scale * dwp. */
int tadnn_conv_weight_pack(const float* w, void* wp, int O, int C, int KH, int KW, int Kpad,
                           tadnn_stream_t stream);
int tadnn_conv_wgrad_unpack(const float* dwp, float* grad, int O, int C, int KH, int KW, int Kpad,
                            float scale, tadnn_stream_t stream);

/* Fused first stage of the example CNN (csrc/conv1_fused.cu): implicit-GEMM conv5x5(3->64) on
 * tcgen05 with the im2col rows generated in shared memory, bias + ReLU + 3x3/3 max-pool in the
 * epilogue; backward regenerates both GEMM operands on the fly from (x, dp1, p1, idx1).  Replaces
 * SpatialConvolutionMM + ReLU + SpatialMaxPooling of cifar_example/sgd-torchad_nn-cifar.lua:91-93
 * and their backward.  x: [B,3,32,32] bf16; p1/dp1: [B,9,9,64] bf16; idx1: [B,9,9,64] bytes.
 * w_packed: bf16 [64,80] produced by tadnn_conv1_weight_pack from the Torch7 fp32 [64,75] weight;
 * grad_weight: fp32 [64,75] in the Torch7 order, accumulated (+=). */
int tadnn_conv1_weight_pack(const float* w, void* w_packed, tadnn_stream_t stream);
/* Second stage (csrc/conv2_fused.cu; cifar_example/sgd-torchad_nn-cifar.lua:96-98): conv5x5(64->64) + bias +
 * ReLU + maxpool3/3 on p1 [B,9,9,64] as ONE implicit-GEMM kernel whose A tiles are 4-D TMA boxes of p1 (no
 * im2col matrix); writes p2 [B,64] bf16 and the arg-max bytes idx2 [B,64] (and the un-pooled activation
 * a2 [B,5,5,64] if non-NULL).  w2_packed: bf16 [64, 1600], k = tap*64 + c (tadnn_cnn_pack_shadows).
 * tadnn_conv2_wgrad: dw_packed[o, tap*64 + c] += sum over images and the 3x3 window of
 * da2w[n*9+pos, o] * p1[n, py+kh, px+kw, c], the operands again TMA boxes of p1 (MN-major);
 * tadnn_conv2_wgrad_unpack adds dw_packed into the Torch7-layout gradient [64, 1600] and zeroes it.
 * tadnn_cnn_pack_shadows rebuilds, in one launch, the bf16 operand copies of the fp32 master weights:
 * w1_packed [64,80], w2_packed [64,1600] and the Toeplitz matrix [576, 4032] of the transposed conv. */
int tadnn_conv2_pool_forward(const void* p1, const void* w2_packed, const float* bias, void* p2, void* idx2,
                             void* a2, int B, tadnn_stream_t stream);
int tadnn_conv2_wgrad(const void* p1, const void* da2w, float* dw_packed, int B, tadnn_stream_t stream);
int tadnn_conv2_wgrad_unpack(float* dw_packed, float* grad_weight, tadnn_stream_t stream);
int tadnn_cnn_pack_shadows(const float* w1, const float* w2, void* w1_packed, void* w2_packed,
                           void* toeplitz_t, tadnn_stream_t stream);
/* Same with the conv1 operand of the sliding-operand kernels (bf16 [128, 128], bias folded in) instead of w1_packed. */
int tadnn_cnn_pack_shadows_sliding(const float* w1, const float* b1, const float* w2, void* w1x_packed,
                                   void* w2_packed, void* toeplitz_t, tadnn_stream_t stream);
int tadnn_conv1_pool_forward(const void* x, const void* w_packed, const float* bias, void* p1,
                             void* idx1, int B, tadnn_stream_t stream);
/* dp1 may be stored compactly as [B, dp1_h, dp1_w, 64] (its non-zero top-left corner); pass 9, 9
 * for the full tensor. */
int tadnn_conv1_pool_backward(const void* x, const void* dp1, int dp1_h, int dp1_w, const void* p1,
                              const void* idx1, float* grad_weight, int B, tadnn_stream_t stream);
/* Same, and grad_bias[64] += sum over pooling windows of dp1 * (p1 > 0) in the same pass (the kernel
 * already loads both for every window); grad_bias may be NULL. */
int tadnn_conv1_pool_backward_bias(const void* x, const void* dp1, int dp1_h, int dp1_w, const void* p1,
                                   const void* idx1, float* grad_weight, float* grad_bias, int B,
                                   tadnn_stream_t stream);

/* conv1 + bias + ReLU + maxpool 3x3/3 with the staged image itself as the tcgen05 operand (csrc/conv1_sliding.cu).
 * Input layout NHWC4: bf16 [B, 32, 32, 4] with channel 3 == 1.0 (tadnn_nhwc4_pack writes it from NCHW uint8
 * (src_dtype 0), bf16 (1) or fp32 (2), y = x * host_scale[c] + host_shift[c]; NULL = identity).  w_packed is bf16
 * [128, 128] from tadnn_conv1s_weight_pack (Torch7 weight [64, 75] + bias [64]).  idx1 holds 8 - (window position
 * 0..8 of the arg-max), or 15 where the pooled ReLU output is zero (the backward then needs no p1). */
int tadnn_nhwc4_pack(const void* x, int src_dtype, void* y, long long B, const float* host_scale,
                     const float* host_shift, tadnn_stream_t stream);
int tadnn_conv1s_weight_pack(const float* w, const float* bias, void* w_packed, tadnn_stream_t stream);
int tadnn_conv1s_pool_forward(const void* x_nhwc4, const void* w_packed, void* p1, void* idx1, int B,
                              tadnn_stream_t stream);
/* grad_weight[64,75] (Torch7 layout) and grad_bias[64] (may be NULL) += the conv1 gradients for the pooled gradient
 * dp1 [B,9,9,64] bf16 and the idx1 written by tadnn_conv1s_pool_forward; only pooled rows 0..rows-1 of dp1 are read
 * (7 for the example CNN: conv2 + pool2 never touch the last two).  d(conv out) is rebuilt in tensor memory and is the
 * A operand of the MMAs; the B operand is the staged image. */
int tadnn_conv1s_pool_backward(const void* x_nhwc4, const void* dp1, const void* idx1, float* grad_weight,
                               float* grad_bias, int rows, int B, tadnn_stream_t stream);

/* Fused classifier head: dropout -> Linear(D,H)+ReLU -> Linear(H,NC) -> log-softmax -> NLL.
 * Limits: D <= 64, H <= 128, NC <= 32.  fc weights in Torch7 nn.Linear layout, fp32. */
typedef struct tadnn_head_args {
  const void* p2;            /* [B, D] bf16 input (second pooling output) */
  const float* w1; const float* b1;   /* [H, D], [H] */
  const float* w2; const float* b2;   /* [NC, H], [NC] */
  const long long* target;   /* [B] 0-based class indices */
  int B, D, H, NC;
  float dropout_p;           /* 0 disables dropout (evaluation) */
  unsigned int seed;         /* dropout mask = hash(seed, element); same seed in forward and backward */
  float* hidden;             /* [B, H] fp32, written by forward, read by backward */
  float* logp;               /* [B, NC] fp32 log-probabilities */
  float* loss;               /* [1] fp32 mean NLL (forward) */
  float* gw1; float* gb1; float* gw2; float* gb2;  /* backward: fp32 gradients, accumulated (+=) */
  void* dp2;                 /* backward: [B, D] bf16 gradient w.r.t. p2 */
  float grad_scale;          /* backward: upstream dLoss (1 for a plain mean loss) */
  const float* grad_scale_dev; /* backward: optional device scalar multiplied into grad_scale */
  /* backward, optional (all three NULL to skip): the head's input is the output of ReLU + 3x3 max-pooling
   * with saved arg-max bytes (0..8); route the input gradient through both in the same pass.  dpre is
   * [B*9, D] bf16 (row s*9+pos = gradient at window pixel pos of sample s), gbias_pre [D] fp32 (+=) is the
   * bias gradient of the layer feeding the pool.  dp2 may then be NULL. */
  const void* pool_idx;
  void* dpre;
  float* gbias_pre;
  /* optional device word XOR-mixed into `seed` (forward and backward alike): a captured CUDA graph bakes `seed`
   * in, so a per-replay counter in device memory is what gives every replay its own dropout mask */
  const unsigned int* seed_dev;
} tadnn_head_args;
/* Scratch needed by tadnn_head_forward (per-block loss partials + a completion counter).  The
 * caller allocates it ONCE, zero-filled; the kernel leaves the counter at zero again. */
int tadnn_head_workspace_bytes(int B);
int tadnn_head_forward(const tadnn_head_args* a, void* workspace, tadnn_stream_t stream);
int tadnn_head_backward(const tadnn_head_args* a, tadnn_stream_t stream);
/* Same result as tadnn_head_forward, for callers that cannot give it a workspace because they
 * cannot allocate device memory at the point of the call (a latency-bound loop that must stay
 * clear of cudaMalloc, a host that has already committed all device memory to its own pool, a C
 * client with no allocator).  This entry point has no counterpart in the reference.
 * The scratch is a fixed 2 KB array inside the library; the call allocates nothing (no
 * cudaMalloc / cudaFree, no allocating library routine).
 *  - Limit: a->B <= 65536 samples; larger batches are refused with an error.
 *  - The first call on each device looks up the array's device address and sets a kernel
 *    attribute; make that first call before any stream capture is begun on that device.
 *  - All calls on one device share the array: two calls whose kernels can overlap on the GPU
 *    (different streams) race on it.  Calls on one stream are safe back to back.
 *  - The looked-up address is cached for the life of the process: after a cudaDeviceReset() the
 *    cached value is stale, so a process that resets a device must not call this again on it. */
int tadnn_head_forward_noalloc(const tadnn_head_args* a, tadnn_stream_t stream);

/* The whole fused step of the example CNN (sliding-operand conv1 path) in one host call (csrc/cnn_step.cu):
 * phase 1 = pack the bf16 operand copies, forward, loss, and the backward pass down to dp1 (everything but conv1's
 * gradients: nine launches); phase 2 = conv1's weight and bias gradients (one launch); 3 = both.  The split lets a
 * data-parallel caller start synchronising the finished gradients before conv1's backward runs.  Every buffer is the
 * caller's: xq NHWC4 bf16 [B,32,32,4]; parameters and gradients fp32 in Torch7 layouts (gradients are += ); w1x [128,128],
 * w2p [64,1600], Tt [576,4032] bf16 and dwp [64,1600] fp32 (zero on entry, zero again on return) are the operand copies;
 * p1/idx1 [B,9,9,64] bf16/u8, p2/idx2 [B,64], hidden [B,100], logp [B,10], loss [1] fp32, da2w [B*9,64] bf16,
 * dp1 [B,9,9,64] bf16 (rows 7, 8 zero and left alone), head_ws from tadnn_head_workspace_bytes (zero-filled once).
 * tt_mask: HOST k-block mask of Tt for 256-wide column tiles (16 words) or NULL.  forward_only: stop after the loss. */
typedef struct tadnn_cnn_step_args {
  const void* xq;
  const long long* target;
  int B;
  const float *w1, *b1, *w2, *b2, *fw1, *fb1, *fw2, *fb2;
  float *gw1, *gb1, *gw2, *gb2, *gfw1, *gfb1, *gfw2, *gfb2;
  void *w1x, *w2p, *Tt;
  float* dwp;
  void *p1, *idx1, *p2, *idx2;
  float *hidden, *logp, *loss;
  void *da2w, *dp1, *head_ws;
  float dropout_p;
  unsigned int seed;
  const unsigned int* seed_dev;
  float grad_scale;
  const unsigned int* tt_mask;
  int n_tt_mask;
  int phase;
  int forward_only;
} tadnn_cnn_step_args;
int tadnn_cnn_step(const tadnn_cnn_step_args* args, tadnn_stream_t stream);

/* p -= lr * (g + wd * p); optionally g = 0.  Replaces accUpdateGradParameters' in-place step
 * (datamodule.lua:142) for batched training on the flat buffer. */
int tadnn_sgd_step(float* p, float* g, long long n, float lr, float weight_decay, int zero_grad,
                   tadnn_stream_t stream);
/* SGD step (as tadnn_sgd_step, gradients left untouched) that also writes the bf16 copy of the updated weights
 * the GEMMs read, so that no forward pass has to re-cast them. */
int tadnn_sgd_step_shadow(float* p, const float* g, void* shadow_bf16, long long n, float lr, float weight_decay,
                          tadnn_stream_t stream);
int tadnn_cast_f32_bf16(const float* x, void* y, long long n, tadnn_stream_t stream);
/* 2-D form with row pitches (elements): lets a layer of any width become a TMA operand (pitch rounded up to 8). */
int tadnn_cast_f32_bf16_2d(const float* x, long long ldx, void* y, long long ldy, long long rows, int cols,
                           tadnn_stream_t stream);

/* ---- linear layers with a tiny output width (csrc/smalln.cu) -----------------------------------
 * The classifier layer of the README MLP is MPBaseLinear(i, 10) (nodemodule.lua:86-119): one 128-wide
 * MMA tile would waste >90 % of the tensor core and leave most SMs without work, so N <= 16 gets
 * bandwidth-shaped CUDA-core kernels.  x, W, g, dx are bf16 with row pitches in elements (K-wide
 * operands: pitch % 8 == 0 and 16-byte aligned); out, dW, db are fp32.
 *   fwd   : out[m, n] = sum_k x[m, k] W[n, k] (+ bias[n])
 *   dgrad : dx[m, k]  = (sum_n g[m, n] W[n, k]) * (aux ? 1 - aux[m, k]^2 : 1)
 *   wgrad : dW[n, k] (+)= sum_m g[m, n] x[m, k];   db[n] (+)= sum_m g[m, n]   (db may be NULL) */
int tadnn_linear_smalln_fwd(const void* x, int ldx, const void* W, int ldw, const float* bias, float* out,
                            int ldo, int M, int N, int K, tadnn_stream_t stream);
int tadnn_linear_smalln_dgrad(const void* g, int ldg, const void* W, int ldw, const void* aux_tanh_out,
                              int ld_aux, void* dx, int lddx, int M, int N, int K, tadnn_stream_t stream);
int tadnn_linear_smalln_wgrad(const void* g, int ldg, const void* x, int ldx, float* dW, int lddw, float* db,
                              int M, int N, int K, int accumulate, tadnn_stream_t stream);

/* ---- fused MSE criterion (csrc/loss.cu) ----------------------------------------------------------
 * nn.MSECriterion forward + updateGradInput (datamodule.lua:139-141) in one pass over a bf16 output:
 * loss[0] = scale_loss * sum (y - t)^2, grad = scale_grad * (y - t) as bf16 (grad may be NULL).
 * target is bf16 (target_fp32 = 0) or fp32 (1).  All pointers 16-byte aligned. */
int tadnn_mse_loss_grad(const void* y, const void* target, int target_fp32, void* grad, float* loss,
                        long long n, float scale_loss, float scale_grad, tadnn_stream_t stream);

/* ---- dataset preprocessing (csrc/preprocess.cu) -------------------------------------------------
 * cifar_example/sgd-torchad_nn-cifar.lua:182-216: rgb2yuv, U/V standardised with the training
 * shard's mean/std, Y through SpatialContrastiveNormalization(1, gaussian1D(7)).  rgb is fp32
 * [n, 3, H, W]; stats4 receives (sum u, sum u^2, sum v, sum v^2) in fp64 on the device; out is fp32 or
 * bf16 [n, 3, H, W].  H*W <= 5120, odd tap count <= 15.  threshold / thresval are those of
 * nn.SpatialDivisiveNormalization (defaults 1e-4 / threshold).  Pass a negative std for the reference's
 * sign quirk (lines 195, 200, 213, 216). */
int tadnn_preprocess_uv_stats(const float* rgb, double* stats4, long long n_img, int H, int W,
                              tadnn_stream_t stream);
int tadnn_preprocess_yuv_lcn(const float* rgb, void* out, int out_bf16, long long n_img, int H, int W,
                             const float* taps1d, int n_taps, float mean_u, float std_u, float mean_v,
                             float std_v, float threshold, float thresval, tadnn_stream_t stream);

/* uint8 -> bf16 widening with a per-channel affine map, y = x * scale[c] + shift[c], over an NCHW-style
 * buffer (`plane` = H*W values per channel, C <= 4 channels; scale/shift are HOST arrays of C floats, NULL =
 * identity).  The reference keeps its shard device-resident as preprocessed floats
 * (cifar_example/sgd-torchad_nn-cifar.lua:239-245); a streaming input pipeline instead ships the raw bytes
 * (3 KB per CIFAR image) over PCIe every step and widens them here. */
int tadnn_u8_to_bf16(const void* x, void* y, long long n, int plane, int C, const float* host_scale,
                     const float* host_shift, tadnn_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* TADNN_H_ */
