"""Config 5: gradient all-reduce bandwidth sweep 1 KB - 1 GB, tadnn kernels vs NCCL.

    python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 examples/allreduce_sweep.py

Mirrors the call pattern of synchronizeModel (datamodule.lua:211-224) and the dead comm_speed
probe (datamodule.lua:280-303).  Device-timed with CUDA events, max over ranks; bus GB/s =
algBW * 2(P-1)/P.  Prints one JSON line per size from rank 0 and writes
gpurun_out/allreduce_sweep_P<world>.json.
"""
import argparse
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import tadnn.mpi as mpi  # noqa: E402


def timed(fn, iters, warmup):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    dist.barrier()
    e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1) / iters], device="cuda")
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    return ms.item()


def sweep(h, sizes=None, max_mb=1024, dtype="fp32", quick=False):
    """Run the sweep on a started handle; returns the rows on every rank.  ``frac_of_link`` is the
    best tadnn bus bandwidth over the NVLink-5 per-direction rate (900 GB/s): with the bus-bandwidth
    convention algBW*2(P-1)/P, a ring or two-shot all-reduce that saturates every link reads 1.0."""
    ctx = h.symm
    P = h.size()
    tdt = torch.float32 if dtype == "fp32" else torch.bfloat16
    es = 4 if dtype == "fp32" else 2
    if sizes is None:
        sizes, nbytes = [], 1024
        while nbytes <= (max_mb << 20):
            sizes.append(nbytes)
            nbytes *= 4
    top = max(sizes)
    region = ctx.empty((top // es,), tdt)
    region.normal_()
    nccl_buf = torch.randn(top // es, device="cuda").to(tdt)
    rows = []
    for nbytes in sizes:
        n = nbytes // es
        iters = 200 if nbytes <= (1 << 20) else (50 if nbytes <= (64 << 20) else 10)
        if quick:
            iters = max(5, iters // 4)
        res = {"bytes": nbytes, "world": P, "dtype": dtype}
        view = region[:n]
        best = 0.0
        for name, algo in (("oneshot", 1), ("twoshot", 2), ("nvls", 3)):
            if algo == 1 and nbytes > (8 << 20):
                continue
            if algo == 3 and not ctx.has_multicast:
                continue
            if quick and algo == 2 and ctx.has_multicast and nbytes > (1 << 20):
                continue
            ms = timed(lambda: ctx.allreduce_(view, scale=1.0 / P, algo=algo), iters, 3 if quick else 5)
            res[name + "_us"] = round(ms * 1e3, 2)
            res[name + "_busGBs"] = round(nbytes / (ms * 1e-3) * 2 * (P - 1) / P / 1e9, 2)
            best = max(best, res[name + "_busGBs"])
        ms = timed(lambda: ctx.allreduce_(view, scale=1.0 / P), iters, 3 if quick else 5)
        res["auto_us"] = round(ms * 1e3, 2)
        res["auto_busGBs"] = round(nbytes / (ms * 1e-3) * 2 * (P - 1) / P / 1e9, 2)
        res["frac_of_link_900GBs"] = round(max(best, res["auto_busGBs"]) / 900.0, 3)
        nb = nccl_buf[:n]

        def nccl():
            dist.all_reduce(nb)
            nb.div_(P)
        ms = timed(nccl, iters, 3 if quick else 5)
        res["nccl_div_us"] = round(ms * 1e3, 2)
        res["nccl_div_busGBs"] = round(nbytes / (ms * 1e-3) * 2 * (P - 1) / P / 1e9, 2)
        if not quick:
            ms = timed(lambda: dist.all_reduce(nb), iters, 5)
            res["nccl_us"] = round(ms * 1e3, 2)
            res["nccl_busGBs"] = round(nbytes / (ms * 1e-3) * 2 * (P - 1) / P / 1e9, 2)
        rows.append(res)
    return rows


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--max-mb", type=int, default=1024)
    ap.add_argument("--dtype", default="fp32", choices=["fp32", "bf16"])
    ap.add_argument("--out", default="gpurun_out")
    a = ap.parse_args()
    os.environ.setdefault("TADNN_SYMM_HEAP_MB", str(a.max_mb + 256))
    h = mpi.start(True)
    rows = sweep(h, None, a.max_mb, a.dtype)
    if h.rank() == 0:
        for r in rows:
            print(json.dumps(r), flush=True)
        os.makedirs(a.out, exist_ok=True)
        with open(os.path.join(a.out, f"allreduce_sweep_P{h.size()}_{a.dtype}.json"), "w") as f:
            json.dump(rows, f, indent=1)
    h.stop()


if __name__ == "__main__":
    main()
