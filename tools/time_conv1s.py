"""CUDA-event timing of the conv1s kernels at B = 4096 (L2 flushed between launches)."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from tadnn.ops import cnn as K
from tools.dbg_conv1s import timeit
B = 4096
x = torch.randn(B, 3, 32, 32, device="cuda").bfloat16()
w = torch.randn(64, 75, device="cuda") * 0.2
b = torch.randn(64, device="cuda") * 0.1
xq = K.nhwc4_pack(x); wx = K.conv1s_weight_pack(w, b)
tag = os.environ.get("TADNN_CONV1S_DBG", "0")
print(tag, "conv1s fwd us (median, min):", timeit(lambda: K.conv1s_pool_forward(xq, wx)))
p1, idx1 = K.conv1s_pool_forward(xq, wx)
dp1 = torch.zeros(B, 9, 9, 64, device="cuda", dtype=torch.bfloat16)
dp1[:, :7, :7] = torch.randn(B, 7, 7, 64, device="cuda").bfloat16()
gw, gb = torch.zeros(64, 75, device="cuda"), torch.zeros(64, device="cuda")
print(tag, "conv1s bwd us (median, min):", timeit(lambda: K.conv1s_pool_backward(xq, dp1, idx1, gw, gb, rows=7)))
for Bn in (148, 1024):
    print(tag, f"conv1s fwd B={Bn} us:", timeit(lambda: K.conv1s_pool_forward(xq[:Bn], wx)))
    print(tag, f"conv1s bwd B={Bn} us:", timeit(lambda: K.conv1s_pool_backward(xq[:Bn], dp1[:Bn], idx1[:Bn], gw, gb, rows=7)))
