"""Configs 2 and 4: model-parallel MLP training throughput, fused tadnn kernels vs the NCCL
restatement of the reference (narrow -> F.linear -> all_reduce per layer).

    torchrun --nproc-per-node 8 examples/mp_mlp_bench.py --config wide   [--impl tadnn|nccl]
    torchrun --nproc-per-node 8 examples/mp_mlp_bench.py --config readme

  readme : README.md:97-101  MPInitialReshape(1024) -> MPInitialLinear(1024,2048) -> MPTanh -> MPBaseLinear(2048,10)
  wide   : 4 x Linear 16384x16384 with MPTanh between (BASELINE.json config 4)

One step = forward + MSE/CE loss + backward + SGD update.  Device-timed (CUDA events), max over
ranks.  Prints one JSON line from rank 0 incl. the roofline estimate: the slower of FLOPs at
the measured GEMM peak and NVLink bytes at the measured 770 GB/s per direction.
"""
import argparse
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def run(h, config="wide", impl="tadnn", batch=0, steps=20, warmup=5, reduce="", chunks=0, fused_loss=None, phases=False):
    """One measurement on an already started handle ``h``; returns the result dict on every rank.
    ``impl="nccl"`` builds the model from the per-op MP layers on ``h.nccl_only()``."""
    from tadnn import nn
    from tadnn.parallel.tp import FusedMPMLP
    P, rank = h.size(), h.rank()
    dev = torch.device("cuda", torch.cuda.current_device())
    torch.manual_seed(1)
    if config == "wide":
        sizes, B = [16384] * 5, (batch or 4096)
    else:
        sizes, B = [1024, 2048, 10], (batch or 8192)
    if impl == "tadnn":
        model = FusedMPMLP(sizes, mpi=h).to(dev)
        if reduce:
            model.reduce_mode = reduce
        if chunks:
            model.nvls_chunks = chunks
    else:
        hn = h.nccl_only()
        model = nn.Sequential()
        for l in range(len(sizes) - 1):
            model.add((nn.MPInitialLinear if l == 0 else nn.MPBaseLinear)(sizes[l], sizes[l + 1], mpi=hn))
            if l < len(sizes) - 2:
                model.add(nn.MPTanh(mpi=hn))
        model = model.to(dev)
        model._tadnn_no_autofuse = True
    params = [p for p in model.parameters()]
    xs = [torch.randn(B, sizes[0], device=dev).bfloat16() for _ in range(4)]
    tg = torch.randn(B, sizes[-1], device=dev)
    if fused_loss is None:
        fused_loss = impl == "tadnn" and sizes[-1] % 8 == 0
    fused_loss = bool(fused_loss) and impl == "tadnn"
    if fused_loss:
        from tadnn.ops.loss import mse_fused
        model.out_bf16 = True
        tgb = tg.bfloat16()
    opt = getattr(model, "sgd_step", None)

    def fwd(i):
        y = model(xs[i % 4]) if impl == "tadnn" else torch.nn.Sequential.forward(model, xs[i % 4])
        return mse_fused(y, tgb) if fused_loss else ((y.float() - tg) ** 2).mean()

    def update():
        if opt is not None:
            opt(1e-4)
            return
        with torch.no_grad():
            torch._foreach_add_(params, [p.grad for p in params], alpha=-1e-4)
            for p in params:
                p.grad = None

    def step(i):
        loss = fwd(i)
        loss.backward()
        update()
        return loss

    for i in range(warmup):
        step(i)
    torch.cuda.synchronize()
    dist.barrier()
    e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
    e0.record()
    for i in range(steps):
        loss = step(i)
    e1.record()
    torch.cuda.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1) / steps], device=dev)
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms = ms.item()
    ph = None
    if phases:
        ev = [torch.cuda.Event(True) for _ in range(4)]
        acc = [0.0, 0.0, 0.0]
        for i in range(5):
            ev[0].record()
            loss = fwd(i)
            ev[1].record()
            loss.backward()
            ev[2].record()
            update()
            ev[3].record()
            torch.cuda.synchronize()
            for k in range(3):
                acc[k] += ev[k].elapsed_time(ev[k + 1]) / 5
        ph = {"forward_ms": acc[0], "backward_ms": acc[1], "update_ms": acc[2]}
        if impl == "tadnn":
            from tadnn import _native
            ph["launch_table"] = _native.launch_table()
    L = len(sizes) - 1
    flops = 0.0
    nvl = 0.0
    for l in range(L):
        f = 2.0 * B * sizes[l] / P * sizes[l + 1]
        flops += f * (3 if l > 0 else 2)  # fwd + wgrad (+ dgrad except for the first layer)
        nvl += (P - 1) / P * B * sizes[l + 1] * 2  # fwd reduce-scatter of bf16 partials
        if l > 0:
            nvl += (P - 1) / P * B * sizes[l] * 2  # bwd all-gather of the slice gradients
    peaks = {"bf16_tflops_sustained": 1387.2}
    try:
        peaks = json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")))
    except OSError:
        pass
    t_compute = flops / (peaks["bf16_tflops_sustained"] * 1e12) * 1e3
    t_link = nvl / 770e9 * 1e3
    res = {"config": config, "impl": impl, "fused_loss": fused_loss, "reduce": (model.reduce_mode if impl == "tadnn" else None),
           "chunks": (model.nvls_chunks if impl == "tadnn" else None), "world": P, "batch": B, "sizes": sizes, "ms_per_step": ms,
           "samples_per_s": B / ms * 1e3, "tflops_per_gpu": flops / ms / 1e9,
           "roofline_ms": max(t_compute, t_link), "compute_ms_at_measured_peak": t_compute,
           "nvlink_ms_at_770GBs": t_link, "fraction_of_roofline": max(t_compute, t_link) / ms,
           "loss": float(loss), "phases": ph}
    del model, params, xs, tg
    torch.cuda.empty_cache()
    return res


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", default="wide", choices=["wide", "readme"])
    ap.add_argument("--impl", default="tadnn", choices=["tadnn", "nccl"])
    ap.add_argument("--batch", type=int, default=0)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--out", default="gpurun_out")
    ap.add_argument("--reduce", default="", choices=["", "push", "nvls"], help="forward reduction of FusedMPMLP (default: its own default)")
    ap.add_argument("--chunks", type=int, default=0, help="row chunks of the NVLS pipeline")
    ap.add_argument("--fused-loss", type=int, default=-1, help="1: bf16 model output + the fused MSE kernel (ops/loss.py); 0: fp32 torch ops; -1: auto")
    ap.add_argument("--phases", action="store_true", help="also time forward / backward / update separately")
    a = ap.parse_args()
    os.environ.setdefault("TADNN_SYMM_HEAP_MB", "3072")
    import tadnn.mpi as mpi
    h = mpi.start(True)
    res = run(h, a.config, a.impl, a.batch, a.steps, a.warmup, a.reduce, a.chunks, None if a.fused_loss < 0 else bool(a.fused_loss), a.phases)
    if h.rank() == 0:
        print(json.dumps(res), flush=True)
        os.makedirs(a.out, exist_ok=True)
        with open(os.path.join(a.out, f"mp_{a.config}_{a.impl}{('_' + a.reduce + str(a.chunks)) if a.reduce else ''}_P{h.size()}.json"), "w") as f:
            json.dump(res, f)
    h.stop()


if __name__ == "__main__":
    main()
