"""Where does a forward layer of the fused model-parallel MLP spend its time at P ranks?

    python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 tools/mp_layer_timeline.py [--reduce nvls]

CUDA events are recorded after every launch group of FusedMPMLP's forward (weight cast, GEMM with the
reduce-scatter push, reduction).  Per label the tool prints the mean over steps of the MAX and the MIN
over ranks: a large max-min gap on a "reduce" row is time spent waiting in the barrier for the slowest
rank (skew), a large min is the kernel itself.  The loss arithmetic of examples/mp_mlp_bench.py is timed
the same way (torch fp32 ops vs the fused criterion).  DESIGN.md section 10, item 1.
"""
import argparse
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import tadnn.mpi as mpi  # noqa: E402
from tadnn.parallel.tp import FusedMPMLP  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--width", type=int, default=16384)
    ap.add_argument("--layers", type=int, default=4)
    ap.add_argument("--batch", type=int, default=4096)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--reduce", default="push", choices=["push", "nvls"])
    ap.add_argument("--chunks", type=int, default=1)
    ap.add_argument("--out", default="gpurun_out")
    a = ap.parse_args()
    h = mpi.start(True)
    dev = h.device()
    P = h.size()
    sizes = [a.width] * (a.layers + 1)
    torch.manual_seed(1)
    model = FusedMPMLP(sizes, mpi=h).to(dev)
    model.reduce_mode, model.nvls_chunks = a.reduce, a.chunks
    x = torch.randn(a.batch, sizes[0], device=dev).bfloat16()
    tg = torch.randn(a.batch, sizes[-1], device=dev)
    for _ in range(3):
        model(x)
    torch.cuda.synchronize()
    dist.barrier()
    sums = {}
    order = []
    for _ in range(a.steps):
        model._trace = []
        with torch.no_grad():
            y = model(x)
        e0, e1, e2 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
        e0.record()
        loss = ((y.float() - tg) ** 2).mean()
        e1.record()
        torch.cuda.synchronize()
        tr = model._trace
        for (_, prev), (label, ev) in zip(tr[:-1], tr[1:]):
            if label not in sums:
                sums[label] = 0.0
                order.append(label)
            sums[label] += prev.elapsed_time(ev)
        if "loss.torch_fp32_forward" not in sums:
            sums["loss.torch_fp32_forward"] = 0.0
            order.append("loss.torch_fp32_forward")
        sums["loss.torch_fp32_forward"] += e0.elapsed_time(e1)
        del loss, e2
    model._trace = None
    vals = torch.tensor([sums[k] / a.steps for k in order], device=dev)
    vmax, vmin = vals.clone(), vals.clone()
    dist.all_reduce(vmax, op=dist.ReduceOp.MAX)
    dist.all_reduce(vmin, op=dist.ReduceOp.MIN)
    if h.rank() == 0:
        rows = [{"phase": k, "ms_max_over_ranks": round(vmax[i].item(), 4), "ms_min_over_ranks": round(vmin[i].item(), 4)}
                for i, k in enumerate(order)]
        for r in rows:
            print(f'{r["phase"]:28s} max {r["ms_max_over_ranks"]:8.4f} ms   min {r["ms_min_over_ranks"]:8.4f} ms')
        os.makedirs(a.out, exist_ok=True)
        with open(os.path.join(a.out, f"mp_layer_timeline_{a.reduce}{a.chunks}_P{P}.json"), "w") as f:
            json.dump({"world": P, "sizes": sizes, "batch": a.batch, "reduce": a.reduce, "chunks": a.chunks, "phases": rows}, f, indent=1)
    h.stop()


if __name__ == "__main__":
    main()
