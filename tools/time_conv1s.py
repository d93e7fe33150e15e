"""CUDA-event timing of conv1s forward at B = 4096 (L2 flushed between launches)."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from tadnn.ops import cnn as K
from tools.dbg_conv1s import timeit
B = 4096
x = torch.randn(B, 3, 32, 32, device="cuda").bfloat16()
w = torch.randn(64, 75, device="cuda") * 0.2
b = torch.randn(64, device="cuda") * 0.1
xq = K.nhwc4_pack(x); wx = K.conv1s_weight_pack(w, b)
print(os.environ.get("TADNN_CONV1S_DBG", "0"), "conv1s fwd us (median, min):", timeit(lambda: K.conv1s_pool_forward(xq, wx)))
