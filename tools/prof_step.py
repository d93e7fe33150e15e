"""A few training steps of the flagship CNN (B = 4096, one GPU, no communication) for ncu:
    python tools/prof_step.py [steps]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import tadnn  # noqa: E402
from tadnn.models.cifar_cnn import CifarCNN  # noqa: E402

steps = int(sys.argv[1]) if len(sys.argv) > 1 else 4
torch.manual_seed(0)
m = CifarCNN().cuda()
tr = tadnn.BatchTrainer(m, 1e-3)
xs = [torch.randn(4096, 3, 32, 32, device="cuda").bfloat16() for _ in range(4)]
if m.packs_input:   # the data pipeline packs once (BatchTrainer.fit does it in its widening kernel)
    xs = [m.pack_input(x) for x in xs]
ys = [torch.randint(0, 10, (4096,), device="cuda") for _ in range(4)]
for i in range(steps):
    loss = tr.step(xs[i % 4], ys[i % 4])
torch.cuda.synchronize()
print("loss", float(loss))
