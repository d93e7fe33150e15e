#!/usr/bin/env python
"""Flagship benchmark: BASELINE.json config 3 — the reference example's CIFAR-10-shaped CNN
(conv stack + 2 FC; cifar_example/sgd-torchad_nn-cifar.lua:88-118) trained data-parallel through
``tadnn.parallelize()`` on N B200s, bf16 compute, synthetic CIFAR-shaped data, random-init
weights.  Metric: training samples/s, whole job, device-timed, MAX over ranks.

    python bench.py --gpus N --steps K --warmup W
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port P bench.py --gpus N --steps K --warmup W
    python bench.py --impl reference ...     # the unmodified reference (unavailable: Lua/Torch7)
    python bench.py --impl baseline ...      # PyTorch + cuDNN + NCCL restatement only

Every step is a full training step: forward -> loss -> backward -> the hooked model
synchronisation (fused all-reduce-average, every step: batchSize=1) -> SGD update.  Weak
scaling: the per-GPU batch is fixed.

What one invocation measures, in this order:
  1. ``value`` — the device-timed step (inputs already on the GPU, 8 distinct batches > L2);
  2. ``e2e`` — the same steps through ``BatchTrainer.fit``: every step copies its uint8 images
     and labels from pinned host memory, widens them on the device, and reads its loss back;
  3. ``baseline_arm`` — the comparator BASELINE.md names (PyTorch + cuDNN + NCCL restatement of
     datamodule.lua:211-224), same N, same steps, same two loops, in the same process;
  4. ``selfcheck`` (N > 1) — replicas bit-identical after the run; one tadnn all-reduce equal to
     ``dist.all_reduce`` on the same data;
  5. ``extra`` (N > 1, skippable with --no-extra) — BASELINE.json configs 2, 4 and 5: README and
     wide model-parallel MLP (fused kernels vs NCCL restatement, fraction of roofline) and the
     all-reduce sweep (tadnn vs NCCL, bus GB/s and fraction of the 900 GB/s link rate).
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "train_samples_per_sec"
MODEL = "cifar_example CNN (conv5x5x64-relu-pool3, conv5x5x64-relu-pool3, fc64x100-relu, fc100x10, logsoftmax+CE; 114838 params)"
# BASELINE.md row b1 (PyTorch + cuDNN + NCCL restatement on this box), samples/s by GPU count
PUBLISHED_B1 = {1: 0.82e6, 2: 1.57e6, 4: 3.08e6, 8: 5.99e6}
AFFINE = ([4.0 / 255.0] * 3, [-2.0] * 3)  # uint8 pixel -> roughly N(0,1)-ranged bf16


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="tadnn", choices=["tadnn", "reference", "baseline"])
    ap.add_argument("--batch-per-gpu", type=int, default=4096)
    ap.add_argument("--pool-batches", type=int, default=8,
                    help="distinct batches cycled through (8 x 25 MB of bf16 on the device > 126 MB L2)")
    ap.add_argument("--lr", type=float, default=1e-3)
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-baseline", action="store_true", help="skip the in-process comparator arm")
    ap.add_argument("--no-extra", action="store_true", help="skip the MP-MLP / all-reduce-sweep block (N > 1)")
    ap.add_argument("--extra-steps", type=int, default=10)
    ap.add_argument("--conv1", default="sliding", choices=["sliding", "rows"],
                    help="fused conv1 kernels: the staged NHWC4 image as the tensor-core operand, or im2col rows built in shared memory")
    ap.add_argument("--conv2", default="dense", choices=["dense", "pooled", "legacy"],
                    help="conv2 forward: all 5x5 output pixels (reference behaviour) or only the 3x3 that pool2 reads")
    return ap.parse_args()


def reference_arm(a):
    """The reference is 602 lines of Lua on Torch7 + TorchMPI + OpenMPI 2.0.1 (README.md:9-17).
    `pip install --no-index ... --target baseline/_ref /root/reference` fails ("neither setup.py
    nor pyproject.toml"), and no LuaJIT/Torch7/TorchMPI exists on this image."""
    if int(os.environ.get("RANK", "0")) == 0:
        print(json.dumps({"impl": "reference", "unavailable":
                          "reference is Lua/Torch7+TorchMPI (no setup.py/pyproject; pip install fails; no luajit/th/mpirun on the image)"}))
    return 0


class SyntheticCifar:
    """A sliceable synthetic dataset in CIFAR's own storage format — one byte per value, 3 KB per
    image: ``data[start:finish]`` materialises only that stripe (pinned host memory), so
    ``parallelize()`` can shard a dataset that no rank ever holds whole.  Pixels are noise around
    a class-dependent level (learnable)."""

    def __init__(self, total: int, what: str, seed: int = 1):
        self.total, self.what, self.seed = total, what, seed
        self.shape = (total, 3, 32, 32) if what == "data" else (total,)

    def __len__(self):
        return self.total

    def __getitem__(self, sl):
        import torch
        assert isinstance(sl, slice)
        start, stop, _ = sl.indices(self.total)
        n = stop - start
        g = torch.Generator().manual_seed(self.seed * 1000003 + start)
        labels = torch.randint(0, 10, (n,), generator=g)
        if self.what == "labels":
            return labels.pin_memory() if torch.cuda.is_available() else labels
        x = torch.randn((n, 3, 32, 32), generator=g, dtype=torch.float32)
        x += (labels.float().view(n, 1, 1, 1) - 4.5) * 0.1
        x = (x * 64.0 + 128.0).clamp_(0, 255).to(torch.uint8)
        return x.pin_memory() if torch.cuda.is_available() else x


def main():
    a = parse()
    if a.impl == "reference":
        return reference_arm(a)
    import torch
    import torch.distributed as dist

    if not torch.cuda.is_available():
        print(json.dumps({"metric": METRIC, "value": None, "error": "no CUDA device"}))
        return 1
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if world == 1 and a.gpus > 1:
        print(json.dumps({"metric": METRIC, "value": None,
                          "error": "launch with torch.distributed.run --nproc-per-node N for N>1"}))
        return 1
    os.environ.setdefault("TADNN_SYMM_HEAP_MB", "4096" if (world > 1 and not a.no_extra) else "512")
    import tadnn
    import tadnn.mpi as mpi
    from tadnn import _native
    from tadnn.models.cifar_cnn import CifarCNN
    from tadnn.ops.cnn import u8_to_bf16
    from tadnn.utils.timing import ClockSampler

    h = mpi.start(True)
    rank, P = h.rank(), h.size()
    dev = torch.device("cuda", torch.cuda.current_device())
    torch.manual_seed(1)  # same seed on every rank (sgd-torchad_nn-cifar.lua:48)
    B, nb = a.batch_per_gpu, a.pool_batches
    total = B * nb * P
    data, labels = SyntheticCifar(total, "data"), SyntheticCifar(total, "labels")
    W = max(3, a.warmup)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)

    def barrier():
        torch.cuda.synchronize()
        if P > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_ms(ms):
        t = torch.tensor([ms], device=dev, dtype=torch.float64)
        if P > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return t.item()

    def device_timed(step, d_dev, t_dev):
        def run(k0, k):
            for s in range(k0, k0 + k):
                i = (s % nb) * B
                step(d_dev[i:i + B], t_dev[i:i + B])
        run(0, W)
        barrier()
        e0.record()
        run(W, a.steps)
        e1.record()
        barrier()
        return max_ms(e0.elapsed_time(e1))

    def e2e_timed(fit):
        for _ in fit(W, 0):
            pass
        barrier()
        e0.record()
        last = None
        for last in fit(a.steps, W):
            pass
        e1.record()
        barrier()
        return max_ms(e0.elapsed_time(e1)), last

    h2d = B * (3 * 32 * 32 * 1 + 8)
    sampler = None
    out = {}

    # ================================ tadnn arm ================================
    if a.impl == "tadnn":
        model = CifarCNN(conv2_mode=a.conv2, conv1_kernel=a.conv1).to(dev)
        d, t, n = tadnn.parallelize(data, labels, model, total, h, None, 1, verbose=False)
        assert n == B * nb
        trainer = tadnn.BatchTrainer(model, learningRate=a.lr)
        # device-resident pool for the kernel-path number: 8 batches (> L2), widened once into the model's input layout
        # (NHWC4 bf16 for the sliding-operand conv1: 33.5 MB per batch; plain NCHW bf16, 25 MB, otherwise)
        d_dev = model.pack_input(d.to(dev), AFFINE) if model.packs_input else u8_to_bf16(d.to(dev), 1024, 3, *AFFINE)
        t_dev = t.to(dev)
        torch.cuda.synchronize()
        sampler = ClockSampler(torch.cuda.current_device()).start() if rank == 0 else None
        launches0 = _native.launch_count()
        t_wall = time.perf_counter()
        ms = device_timed(trainer.step, d_dev, t_dev)
        launches_all = _native.launch_count() - launches0
        launches = launches_all * a.steps // (a.steps + W)  # the counter also saw the warm-up steps
        table = _native.launch_table()
        e2e = None
        if not a.no_e2e:
            ms2, last = e2e_timed(lambda k, s0: trainer.fit(d, t, B, k, start=s0, input_affine=AFFINE))
            e2e = {"value": a.steps * B * P / (ms2 / 1e3), "unit": "samples/s",
                   "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": 4,
                   "ms_per_step": ms2 / a.steps, "last_loss": last,
                   "input": "uint8 images + int64 labels from pinned host memory, widened to bf16 on the device "
                            + ("(tadnn_nhwc4_pack)" if model.packs_input else "(tadnn_u8_to_bf16)")}
        if sampler:
            sampler.stop()
        value = a.steps * B * P / (ms / 1e3)
        out = {
            "metric": METRIC, "value": value, "unit": "samples/s", "n_gpus": P, "steps": a.steps,
            "warmup": W, "ms_per_step": ms / a.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": (value / PUBLISHED_B1[P]) if P in PUBLISHED_B1 else None,
            "dtype": "bf16", "data": "synthetic", "impl": "tadnn",
            "config": {"model": MODEL, "global_batch": B * P, "per_gpu_batch": B, "seq_len": None,
                       "input": "3x32x32 (uint8 on the host, bf16 on the device)", "parallelism": f"dp{P}",
                       "sync": "params+grads all-reduce-average every step (batchSize=1)",
                       "conv1_kernel": a.conv1, "conv2_forward": a.conv2, "optimizer": "SGD",
                       "l2_policy": f"inputs cycle through {nb} device batches = {nb * d_dev[0].numel() * B * 2 / 1e6:.0f} MB per GPU (> 126 MB L2)",
                       "vs_baseline_denominator": "BASELINE.md row b1 (PyTorch+cuDNN+NCCL restatement, round-1 measurement at this N)"},
            "clocks": sampler.summary() if sampler else None,
            "gpu_launches": launches,
            "gpu_launch_table": table,
        }
        if e2e is not None:
            out["e2e"] = e2e
        # ---- replicas still identical? one all-reduce vs NCCL? -----------------------------------
        if P > 1:
            out["selfcheck"] = selfcheck(torch, dist, h, model, dev, P, rank)
        del d_dev

    # ================================ comparator arm ================================
    if a.impl == "baseline" or not a.no_baseline:
        from baseline.torch_nccl_baseline import BaselineTrainer
        bt = BaselineTrainer(dev, world=P, lr=a.lr, sync="per_tensor")
        db, tb = data[rank * B * nb:(rank + 1) * B * nb], labels[rank * B * nb:(rank + 1) * B * nb]
        db_dev = db.to(dev).to(torch.bfloat16).mul_(AFFINE[0][0]).add_(AFFINE[1][0])
        tb_dev = tb.to(dev)
        msb = device_timed(bt.step, db_dev, tb_dev)
        del db_dev
        vb = a.steps * B * P / (msb / 1e3)
        arm = {"impl": "PyTorch+cuDNN+NCCL restatement of datamodule.lua:211-224 (baseline/torch_nccl_baseline.py, per-tensor all_reduce+div_)",
               "value": vb, "ms_per_step": msb / a.steps}
        if not a.no_e2e:
            msb2, _ = e2e_timed(lambda k, s0: bt.fit(db, tb, B, k, start=s0, affine=(AFFINE[0][0], AFFINE[1][0])))
            arm["e2e"] = a.steps * B * P / (msb2 / 1e3)
            arm["e2e_ms_per_step"] = msb2 / a.steps
        if a.impl == "baseline":
            out = {"metric": METRIC, "value": vb, "unit": "samples/s", "n_gpus": P, "steps": a.steps, "warmup": W,
                   "ms_per_step": msb / a.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                   "dtype": "bf16", "data": "synthetic", "impl": "baseline",
                   "config": {"model": MODEL, "global_batch": B * P, "per_gpu_batch": B, "parallelism": f"dp{P}"},
                   "gpu_launches": 0}
            if "e2e" in arm:
                out["e2e"] = {"value": arm["e2e"], "unit": "samples/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": 4,
                              "ms_per_step": arm["e2e_ms_per_step"]}
        else:
            arm["ratio"] = out["value"] / vb
            if "e2e" in arm and "e2e" in out:
                arm["e2e_ratio"] = out["e2e"]["value"] / arm["e2e"]
            out["baseline_arm"] = arm
        del bt

    # ================================ other BASELINE.json configs ================================
    if a.impl == "tadnn" and P > 1 and not a.no_extra:
        try:
            out["extra"] = extra_block(h, a.extra_steps)
        except Exception as e:  # the headline must survive a failure here
            out["extra"] = {"error": f"{type(e).__name__}: {e}"}
    if a.impl == "tadnn":
        out["bench_wall_s"] = round(time.perf_counter() - t_wall, 2)

    if rank == 0:
        print(json.dumps(out))
    h.stop()
    return 0


def selfcheck(torch, dist, h, model, dev, P, rank):
    """(a) every rank holds bit-identical parameters after the run (all-gather of a 64-bit
    checksum of the flat buffer); (b) one tadnn all-reduce-average equals NCCL's on the same data."""
    res = {}
    flat = getattr(model, "_tadnn_flat", None)
    vec = flat.flat_params() if flat is not None else torch.cat([p.detach().reshape(-1) for p in model.parameters()])
    bits = vec.view(torch.int32).to(torch.int64)
    idx = torch.arange(1, bits.numel() + 1, device=dev, dtype=torch.int64)
    cs = torch.stack([bits.sum(), (bits * idx).sum()])
    allcs = [torch.zeros_like(cs) for _ in range(P)]
    dist.all_gather(allcs, cs)
    res["replicas_bit_identical"] = bool(all(torch.equal(c, allcs[0]) for c in allcs))
    ctx = getattr(h, "symm", None)
    if ctx is not None:
        n = 1 << 18
        g = torch.Generator(device=dev).manual_seed(1234 + rank)
        src = torch.randn(n, device=dev, generator=g)
        reg = ctx.empty((n,), torch.float32)
        reg.copy_(src)
        ref = src.clone()
        dist.all_reduce(ref)
        ref.div_(P)
        torch.cuda.synchronize()
        dist.barrier()
        ctx.allreduce_(reg, scale=1.0 / P)
        torch.cuda.synchronize()
        err = (reg - ref).abs().max().item()
        res["allreduce_vs_nccl_max_abs_err"] = err
        res["allreduce_ok"] = bool(err < 1e-5)
    res["ok"] = bool(res.get("replicas_bit_identical", False) and res.get("allreduce_ok", True))
    return res


def extra_block(h, steps):
    """BASELINE.json configs 2, 4 and 5 at this N (see the module docstring)."""
    sys.path.insert(0, os.path.join(ROOT, "examples"))
    import allreduce_sweep
    import mp_mlp_bench
    ex = {}
    ex["allreduce"] = allreduce_sweep.sweep(h, sizes=[1 << 10, 1 << 16, 1 << 20, 1 << 24, 1 << 28], dtype="fp32", quick=True)
    for cfg in ("readme", "wide"):
        row = {}
        for impl in ("tadnn", "nccl"):
            r = mp_mlp_bench.run(h, cfg, impl, steps=steps, warmup=3)
            row[impl] = {k: r[k] for k in ("ms_per_step", "samples_per_s", "fraction_of_roofline", "roofline_ms", "loss") if k in r}
        row["speedup_vs_nccl"] = row["nccl"]["ms_per_step"] / row["tadnn"]["ms_per_step"]
        ex["mp_" + cfg] = row
    return ex


if __name__ == "__main__":
    sys.exit(main())
