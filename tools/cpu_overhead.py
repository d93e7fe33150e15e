"""How much host time one training step costs (enqueue only), against its device time.
    python tools/cpu_overhead.py"""
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import tadnn  # noqa: E402
from tadnn.models.cifar_cnn import CifarCNN  # noqa: E402

import tadnn.mpi as mpi  # noqa: E402
h = mpi.start(True)     # sets this rank's device
B, N = 4096, 200
torch.manual_seed(1)
m = CifarCNN().cuda()
d = torch.randint(0, 256, (8 * B, 3, 32, 32), dtype=torch.uint8)
t = torch.randint(0, 10, (8 * B,))
d, t, n = tadnn.parallelize(d, t, m, 8 * B, h, None, 1, verbose=False)   # flat parameter buffer, as bench.py
tr = tadnn.BatchTrainer(m, learningRate=1e-3)
nbat = max(1, n // B)
xs = [m.pack_input(d[(i % nbat) * B:(i % nbat + 1) * B].cuda(), m.input_affine) for i in range(8)]
ys = [t[(i % nbat) * B:(i % nbat + 1) * B].cuda() for i in range(8)]
for i in range(10):
    tr.step(xs[i % 8], ys[i % 8])
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
w0 = time.perf_counter()
e0.record()
for i in range(N):
    tr.step(xs[i % 8], ys[i % 8])
e1.record()
w1 = time.perf_counter()
torch.cuda.synchronize()
w2 = time.perf_counter()
print(f"step(): enqueue {1e3 * (w1 - w0) / N:.3f} ms/step on the host, device {e0.elapsed_time(e1) / N:.3f} ms/step, wall {1e3 * (w2 - w0) / N:.3f}")
dp, tp = d.pin_memory(), t.pin_memory()
for _ in tr.fit(dp, tp, B, 10):
    pass
torch.cuda.synchronize()
w0 = time.perf_counter()
for _ in tr.fit(dp, tp, B, N):
    pass
torch.cuda.synchronize()
w1 = time.perf_counter()
print(f"fit(): wall {1e3 * (w1 - w0) / N:.3f} ms/step")
buf = torch.empty((B, 3, 32, 32), dtype=torch.uint8, device="cuda")
torch.cuda.synchronize()
w0 = time.perf_counter()
for i in range(N):
    buf.copy_(dp[(i % nbat) * B:(i % nbat + 1) * B], non_blocking=True)
torch.cuda.synchronize()
w1 = time.perf_counter()
print(f"pinned H2D copy of one batch of uint8 images ({buf.numel() / 1e6:.1f} MB): {1e3 * (w1 - w0) / N:.3f} ms = {buf.numel() * N / (w1 - w0) / 1e9:.1f} GB/s")
if h.size() > 1:
    h.stop()
    sys.exit(0)
import cProfile, pstats
pr = cProfile.Profile()
pr.enable()
for _ in tr.fit(dp, tp, B, N):
    pass
torch.cuda.synchronize()
pr.disable()
pstats.Stats(pr).sort_stats("cumulative").print_stats(22)
