"""A few per-sample (B = 1) training steps of the flagship CNN for an ncu launch list."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import tadnn
from tadnn.models.cifar_cnn import CifarCNN
torch.manual_seed(0)
m = CifarCNN().cuda()
tr = tadnn.BatchTrainer(m, 1e-3)
B = int(sys.argv[1]) if len(sys.argv) > 1 else 1
xs = [m.pack_input(torch.randn(B, 3, 32, 32, device="cuda").bfloat16()) for _ in range(4)]
ys = [torch.randint(0, 10, (B,), device="cuda") for _ in range(4)]
for i in range(8):
    loss = tr.step(xs[i % 4], ys[i % 4])
torch.cuda.synchronize()
print("loss", float(loss))
