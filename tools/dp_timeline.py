"""Where a data-parallel step's time goes: CUDA events after each phase of BatchTrainer's fused step (forward + backward
down to dp1 | conv1 backward with the overlapped tail synchronisation | final synchronised update), averaged over steps.
    python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 tools/dp_timeline.py"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import tadnn  # noqa: E402
import tadnn.mpi as mpi  # noqa: E402
from tadnn.models.cifar_cnn import CifarCNN  # noqa: E402

h = mpi.start(True)
B, N = 4096, 200
torch.manual_seed(1)
m = CifarCNN().cuda()
d = torch.randint(0, 256, (8 * B * h.size(), 3, 32, 32), dtype=torch.uint8)
t = torch.randint(0, 10, (8 * B * h.size(),))
d, t, n = tadnn.parallelize(d, t, m, d.shape[0], h, None, 1, verbose=False)
tr = tadnn.BatchTrainer(m, learningRate=1e-3)
xs = [m.pack_input(d[i * B:(i + 1) * B].cuda(), m.input_affine) for i in range(8)]
ys = [t[i * B:(i + 1) * B].cuda() for i in range(8)]
st = getattr(m, "_tadnn_sync", None)
lr, wd = 1e-3, 0.0
ev = [[torch.cuda.Event(True) for _ in range(4)] for _ in range(N)]
fused = st is not None and st.fused_step_ok()


def step(i, rec):
    x, y = xs[i % 8], ys[i % 8]
    cur = torch.cuda.current_stream()
    if rec:
        ev[i][0].record(cur)
    if fused:
        def tail(p):
            if rec:
                ev[i][1].record(cur)
            st.fused_tail(p, lr, wd)
        m.forward_backward(x, y, tail_ready=tail)
        if rec:
            ev[i][2].record(cur)
        st.fused_finish(m, lr, wd)
    else:
        m.forward_backward(x, y, tail_ready=(lambda p: ev[i][1].record(cur)) if rec else None)
        if rec:
            ev[i][2].record(cur)
        tr._sgd()
    if rec:
        ev[i][3].record(cur)


for i in range(20):
    step(i, False)
torch.cuda.synchronize()
for i in range(N):
    step(i, True)
torch.cuda.synchronize()
seg = [sum(ev[i][k].elapsed_time(ev[i][k + 1]) for i in range(N)) / N * 1e3 for k in range(3)]
gap = sum(ev[i][3].elapsed_time(ev[i + 1][0]) for i in range(N - 1)) / (N - 1) * 1e3
tot = ev[0][0].elapsed_time(ev[N - 1][3]) / (N - 1) * 1e3
if h.rank() == 0:
    print(f"world {h.size()} fused={fused}: forward+backward to dp1 {seg[0]:.1f} us | conv1 backward {seg[1]:.1f} us | "
          f"update {seg[2]:.1f} us | between steps {gap:.1f} us | step {tot:.1f} us")
h.stop()
