"""Small-shape pass over every hand-written kernel family, meant to run under compute-sanitizer:

    compute-sanitizer --tool memcheck|racecheck|synccheck python tools/sanitize_small.py

One tool per gpurun call (B200_PROFILING.md).  Prints one line per family and "sanitize_small: done"."""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from tadnn import _native  # noqa: E402
from tadnn.ops import cnn as K  # noqa: E402
from tadnn.ops import smalln  # noqa: E402
from tadnn.ops.gemm import gemm_bf16  # noqa: E402

torch.manual_seed(0)
lib = _native.lib()
dev = "cuda"


def done(name):
    torch.cuda.synchronize()
    print("ok", name, flush=True)


# ---- collectives, all ranks emulated in one cooperative launch
from tadnn.parallel.symm import _RawCuda  # noqa: E402
world, heap = 4, 4 << 20
comms = (C.c_void_p * world)()
_native.check(lib.tadnn_comm_create_local(world, C.c_ulonglong(heap), comms), "create_local")
views = []
for r in range(world):
    p = C.c_void_p()
    lib.tadnn_comm_heap(C.c_void_p(comms[r]), r, C.byref(p))
    views.append(torch.as_tensor(_RawCuda(int(p.value), heap, comms), device=dev))
for v in views:
    v.zero_()
s = C.c_void_p(_native.stream_ptr())
LLB = 2 * world * 16384 * 8
for algo in (1, 2):
    for dt, code in ((torch.float32, 0), (torch.bfloat16, 1)):
        for v in views:
            v[LLB:].view(dt)[:1003].normal_()
        _native.check(lib.tadnn_allreduce_emulated(comms, world, C.c_ulonglong(LLB), C.c_ulonglong(1003), code, C.c_float(0.25), algo, s), "ar")
    _native.check(lib.tadnn_allreduce_sgd_emulated(comms, world, C.c_ulonglong(LLB + 8192), C.c_ulonglong(LLB), C.c_ulonglong(1024), C.c_float(0.25),
                                                   C.c_float(0.1), C.c_float(0.0), algo, s), "ar sgd")
for ep in (1, 2, 3):
    _native.check(lib.tadnn_allreduce_ll_emulated(comms, world, C.c_ulonglong(0), C.c_ulonglong(LLB), C.c_ulonglong(0), C.c_ulonglong(777),
                                                  C.c_float(0.25), 0, C.c_float(0), C.c_float(0), C.c_uint(ep), s), "ll")
done("collectives (one-shot, two-shot, fused SGD, low-latency pairs; 4 emulated ranks)")
for r in range(world):
    lib.tadnn_comm_destroy(C.c_void_p(comms[r]))

# ---- tcgen05 GEMM: four operand-major combinations, bias/act epilogues, split-K, fp32 and bf16 outputs, aux, local push
M, N, Kd = 200, 136, 192
a, b = torch.randn(M, Kd, device=dev).bfloat16(), torch.randn(N, Kd, device=dev).bfloat16()
at, bt = a.t().contiguous(), b.t().contiguous()
bias = torch.randn(N, device=dev)
gemm_bf16(a, b, bias=bias, act="tanh")
gemm_bf16(a, bt, b_mn_major=True, out_dtype=torch.float32)
gemm_bf16(at, bt, a_mn_major=True, b_mn_major=True, out_dtype=torch.float32, split_k=3)
gemm_bf16(at, b, a_mn_major=True, act="relu", bias=bias)
y = torch.randn(M, N, device=dev).bfloat16()
gemm_bf16(a, bt, b_mn_major=True, aux_tanh_out=y)
push = torch.zeros(M, 2 * N, device=dev, dtype=torch.bfloat16)
flags = torch.zeros(((M + 31) // 32) * 2, device=dev, dtype=torch.int32)
gemm_bf16(a, b, push=[push.data_ptr()], push_ldc=2 * N, push_col_offset=N, sig_flags=[flags.data_ptr()], sig_src=1, flag_stride=2)
big_a, big_b = torch.randn(512, 256, device=dev).bfloat16(), torch.randn(256, 256, device=dev).bfloat16()
gemm_bf16(big_a, big_b)  # 2-CTA kernel
done("gemm_bf16 (tcgen05 1-CTA all majors/epilogues, split-K, push+signal; 2-CTA)")

# ---- CNN kernels
x = torch.randn(3, 3, 32, 32, device=dev).bfloat16()
w1, b1 = torch.randn(64, 75, device=dev) * 0.2, torch.randn(64, device=dev) * 0.1
w2, b2 = torch.randn(64, 1600, device=dev) * 0.05, torch.randn(64, device=dev) * 0.1
w1p = torch.empty(64, 80, device=dev, dtype=torch.bfloat16)
w2p = torch.empty(64, 1600, device=dev, dtype=torch.bfloat16)
Tt = torch.empty(576, 4032, device=dev, dtype=torch.bfloat16)
K.cnn_pack_shadows(w1, w2, w1p, w2p, Tt)
p1, idx1 = K.conv1_pool_forward(x, w1p, b1)
p2, idx2 = K.conv2_pool_forward(p1, w2p, b2)
done("conv1_fwd, conv2_fwd (implicit GEMM + pooled epilogues), pack_shadows")
B = 3
tgt = torch.randint(0, 10, (B,), device=dev)
fw1, fb1 = torch.randn(100, 64, device=dev) * 0.1, torch.zeros(100, device=dev)
fw2, fb2 = torch.randn(10, 100, device=dev) * 0.1, torch.zeros(10, device=dev)
hidden, logp, loss = torch.empty(B, 100, device=dev), torch.empty(B, 10, device=dev), torch.empty((), device=dev)
h = K.HeadArgs()
h.p2, h.w1, h.b1, h.w2, h.b2, h.target = p2.data_ptr(), fw1.data_ptr(), fb1.data_ptr(), fw2.data_ptr(), fb2.data_ptr(), tgt.data_ptr()
h.B, h.D, h.H, h.NC, h.dropout_p, h.seed = B, 64, 100, 10, 0.5, 7
h.hidden, h.logp, h.loss = hidden.data_ptr(), logp.data_ptr(), loss.data_ptr()
_native.check(lib.tadnn_head_forward(C.byref(h), C.c_void_p(K.head_workspace(torch.device(dev)).data_ptr()), s), "head fwd")
g = [torch.zeros(100, 64, device=dev), torch.zeros(100, device=dev), torch.zeros(10, 100, device=dev), torch.zeros(10, device=dev)]
da2w = torch.empty(B * 9, 64, device=dev, dtype=torch.bfloat16)
gb2 = torch.zeros(64, device=dev)
h.gw1, h.gb1, h.gw2, h.gb2, h.grad_scale = g[0].data_ptr(), g[1].data_ptr(), g[2].data_ptr(), g[3].data_ptr(), 1.0
h.pool_idx, h.dpre, h.gbias_pre = idx2.data_ptr(), da2w.data_ptr(), gb2.data_ptr()
_native.check(lib.tadnn_head_backward(C.byref(h), s), "head bwd")
done("head_fwd, head_bwd")
dwp = torch.zeros(64, 1600, device=dev)
K.conv2_wgrad(p1, da2w, dwp)
gw2 = torch.zeros(64, 1600, device=dev)
K.conv2_wgrad_unpack(dwp, gw2)
dp1 = torch.zeros(B, 9, 9, 64, device=dev, dtype=torch.bfloat16)
gemm_bf16(da2w.view(B, 576), Tt, b_mn_major=True, out=dp1.view(B, 5184)[:, :4032])
gw1, gb1 = torch.zeros(64, 75, device=dev), torch.zeros(64, device=dev)
K.conv1_pool_backward_bias(x, dp1, p1, idx1, gw1, gb1)
done("conv2_wgrad, Toeplitz dgrad GEMM, conv1_bwd")
xs = torch.randn(40, 64, device=dev).bfloat16()
ws = torch.randn(10, 64, device=dev).bfloat16()
o = smalln.linear_smalln_fwd(xs, ws, torch.zeros(10, device=dev))
gs = torch.randn(40, 16, device=dev).bfloat16()[:, :10]
smalln.linear_smalln_dgrad(gs, ws)
smalln.linear_smalln_wgrad(gs, xs)
u8 = torch.randint(0, 256, (4, 3, 32, 32), device=dev, dtype=torch.uint8)
K.u8_to_bf16(u8, 1024, 3, [1 / 255.] * 3, [0.] * 3)
pp, gg = torch.randn(1000, device=dev), torch.randn(1000, device=dev)
K.sgd_step(pp, gg, 0.1, 0.0, True)
done("smalln fwd/dgrad/wgrad, u8_to_bf16, sgd_step")
print("sanitize_small: done", flush=True)
