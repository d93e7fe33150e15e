"""One launch (after a warm-up) of every kernel family outside the CNN step, for `ncu --set full`:
    python tools/prof_kernels.py
GEMM 8192^3 (1-CTA rasterised and 2-CTA), the model-parallel layer shapes on one rank (reduce-scatter-push epilogue
into the own staging slots, fused-tanh' dgrad, split-K wgrad, all-gather-push with arrival counters), reduce_slots,
the emulated all-reduces (one-shot / two-shot / fused SGD / low-latency), small-N kernels, u8 widening, MSE criterion."""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from tadnn import _native  # noqa: E402
from tadnn.ops import cnn as K  # noqa: E402
from tadnn.ops import smalln  # noqa: E402
from tadnn.ops.gemm import gemm_bf16, gemm_bf16_2cta  # noqa: E402

torch.manual_seed(0)
lib = _native.lib()
dev = "cuda"
s = C.c_void_p(_native.stream_ptr())

for rep in range(2):  # first pass = warm-up (ncu is told to skip it)
    a = torch.randn(8192, 8192, device=dev).bfloat16()
    b = torch.randn(8192, 8192, device=dev).bfloat16()
    gemm_bf16(a, b, tile_n=256)                 # 1-CTA persistent kernel, rasterised
    gemm_bf16_2cta(a, b)                         # cta_group::2
    del a, b
    # wide model-parallel layer on one of 8 ranks: x [4096, 2048] (K-slice), W_r [16384, 2048]
    x = torch.randn(4096, 2048, device=dev).bfloat16()
    w = torch.randn(16384, 2048, device=dev).bfloat16()
    slots = torch.empty(8, 4096, 2048, device=dev, dtype=torch.bfloat16)
    gemm_bf16(x, w, push=[slots[q].data_ptr() for q in range(8)], push_ldc=2048, rs_width=2048, n_rotate=8)   # fwd: reduce-scatter push
    g = torch.randn(4096, 16384, device=dev).bfloat16()
    y = torch.randn(4096, 2048, device=dev).bfloat16()
    gath = torch.empty(4096, 16384, device=dev, dtype=torch.bfloat16)
    flags = torch.zeros(128 * 8, device=dev, dtype=torch.int32)
    gemm_bf16(g, w, b_mn_major=True, aux_tanh_out=y, tile_n=256, out=gath[:, :2048], push=[gath.data_ptr()], push_ldc=16384,
              push_col_offset=2048, sig_flags=[flags.data_ptr()], sig_src=1, flag_stride=8)                       # dgrad + tanh' + gather push
    gemm_bf16(g, x, a_mn_major=True, b_mn_major=True, out_dtype=torch.float32, split_k=1, tile_n=256)             # wgrad
    del slots, g, gath, w, x, y
    xs = torch.randn(8192, 256, device=dev).bfloat16()
    ws = torch.randn(10, 256, device=dev).bfloat16()
    gs = torch.randn(8192, 16, device=dev).bfloat16()[:, :10]
    smalln.linear_smalln_fwd(xs, ws, torch.zeros(10, device=dev))
    smalln.linear_smalln_dgrad(gs, ws)
    smalln.linear_smalln_wgrad(gs, xs)
    u8 = torch.randint(0, 256, (4096, 3, 32, 32), device=dev, dtype=torch.uint8)
    K.u8_to_bf16(u8, 1024, 3, [1 / 255.] * 3, [0.] * 3)
    from tadnn.ops.loss import mse_fused
    mse_fused(torch.randn(4096, 16384, device=dev).bfloat16(), torch.randn(4096, 16384, device=dev).bfloat16())
    torch.cuda.synchronize()

# collectives: 8 emulated ranks in one cooperative launch, 1 MB of fp32 per rank
from tadnn.parallel.symm import _RawCuda  # noqa: E402
world, heap = 8, 8 << 20
comms = (C.c_void_p * world)()
_native.check(lib.tadnn_comm_create_local(world, C.c_ulonglong(heap), comms), "create_local")
views = []
for r in range(world):
    p = C.c_void_p()
    lib.tadnn_comm_heap(C.c_void_p(comms[r]), r, C.byref(p))
    views.append(torch.as_tensor(_RawCuda(int(p.value), heap, comms), device=dev))
for v in views:
    v.zero_()
LLB = 2 * world * 16384 * 8
n = (1 << 20) // 4
for rep in range(2):
    for algo in (1, 2):
        _native.check(lib.tadnn_allreduce_emulated(comms, world, C.c_ulonglong(LLB), C.c_ulonglong(n), 0, C.c_float(0.125), algo, s), "ar")
        _native.check(lib.tadnn_allreduce_sgd_emulated(comms, world, C.c_ulonglong(LLB + 4 * n), C.c_ulonglong(LLB), C.c_ulonglong(n), C.c_float(0.125),
                                                       C.c_float(0.1), C.c_float(0.0), algo, s), "ar sgd")
    _native.check(lib.tadnn_allreduce_ll_emulated(comms, world, C.c_ulonglong(0), C.c_ulonglong(LLB), C.c_ulonglong(0), C.c_ulonglong(4864),
                                                  C.c_float(0.125), 0, C.c_float(0), C.c_float(0), C.c_uint(rep + 1), s), "ll")
    torch.cuda.synchronize()
print("prof_kernels: done")
