import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from tadnn.ops import cnn as K
B = 4096
torch.manual_seed(0)
x = torch.randn(B, 3, 32, 32, device="cuda").bfloat16()
w = torch.randn(64, 75, device="cuda") * 0.2
b = torch.randn(64, device="cuda") * 0.1
wp = K.conv1_weight_pack(w)
p1, idx1 = K.conv1_pool_forward(x, wp, b)
gw = torch.zeros(64, 75, device="cuda")
def t(fn, n=20):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n * 1e3
for shape in ((9, 9), (7, 7), (9, 7), (3, 3)):
    dp = torch.randn(B, shape[0], shape[1], 64, device="cuda").bfloat16()
    print("bwd dp1", shape, round(t(lambda: K.conv1_pool_backward(x, dp, p1, idx1, gw)), 1), "us")
print("fwd", round(t(lambda: K.conv1_pool_forward(x, wp, b)), 1), "us")
