"""Launch the sliding-operand conv1 kernels a few times at B = 4096 (target of an ncu capture)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from tadnn.ops import cnn as K  # noqa: E402

B = 4096
x = torch.randn(B, 3, 32, 32, device="cuda").bfloat16()
w = torch.randn(64, 75, device="cuda") * 0.2
b = torch.randn(64, device="cuda") * 0.1
xq = K.nhwc4_pack(x)
wx = K.conv1s_weight_pack(w, b)
for _ in range(3):
    p1, idx1 = K.conv1s_pool_forward(xq, wx)
if hasattr(K, "conv1s_pool_backward"):
    dp1 = torch.zeros(B, 9, 9, 64, device="cuda", dtype=torch.bfloat16)
    dp1[:, :7, :7] = torch.randn(B, 7, 7, 64, device="cuda").bfloat16()
    gw, gb = torch.zeros(64, 75, device="cuda"), torch.zeros(64, device="cuda")
    for _ in range(3):
        K.conv1s_pool_backward(xq, dp1, idx1, gw, gb, rows=7)
torch.cuda.synchronize()
print("ok")
