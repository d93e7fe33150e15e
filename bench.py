#!/usr/bin/env python
"""Flagship benchmark: BASELINE.json config 3 — the reference example's CIFAR-10-shaped CNN
(conv stack + 2 FC; cifar_example/sgd-torchad_nn-cifar.lua:88-118) trained data-parallel through
``tadnn.parallelize()`` on N B200s, bf16 compute, synthetic CIFAR-shaped data, random-init
weights.  Metric: training samples/s, whole job, device-timed, MAX over ranks.

    python bench.py --gpus N --steps K --warmup W
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port P bench.py --gpus N --steps K --warmup W
    python bench.py --impl reference ...     # the unmodified reference (unavailable: Lua/Torch7)
    python bench.py --impl baseline ...      # PyTorch + cuDNN + NCCL restatement

Every step is a full training step: H2D-resident batch -> forward -> loss -> backward -> the
hooked model synchronisation (fused all-reduce-average of params and grads, every step:
batchSize=1) -> SGD update.  Weak scaling: the per-GPU batch is fixed.
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


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=500)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", default="tadnn", choices=["tadnn", "reference", "baseline"])
    ap.add_argument("--batch-per-gpu", type=int, default=4096)
    ap.add_argument("--pool-batches", type=int, default=8,
                    help="distinct device-resident batches cycled through (8 x 25 MB > 126 MB L2)")
    ap.add_argument("--lr", type=float, default=1e-3)
    ap.add_argument("--no-e2e", action="store_true")
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
    """A sliceable synthetic dataset: ``data[start:finish]`` materialises only that stripe
    (pinned host memory), so ``parallelize()`` can shard a dataset that no rank ever holds
    whole.  Values are N(0,1) images with a class-dependent offset (learnable)."""

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
        x = x.to(torch.bfloat16)
        return x.pin_memory() if torch.cuda.is_available() else x


def main():
    a = parse()
    if a.impl == "reference":
        return reference_arm(a)
    import torch
    import torch.distributed as dist
    import tadnn
    import tadnn.mpi as mpi
    from tadnn import _native
    from tadnn.models.cifar_cnn import CifarCNN
    from tadnn.utils.timing import ClockSampler

    if not torch.cuda.is_available():
        print(json.dumps({"metric": METRIC, "value": None, "error": "no CUDA device"}))
        return 1
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if world == 1 and a.gpus > 1:
        print(json.dumps({"metric": METRIC, "value": None,
                          "error": "launch with torch.distributed.run --nproc-per-node N for N>1"}))
        return 1
    h = mpi.start(True) if (world > 1 or a.impl == "tadnn") else None
    rank = h.rank() if h else 0
    P = h.size() if h else 1
    dev = torch.device("cuda", torch.cuda.current_device())
    torch.manual_seed(1)  # same seed on every rank (sgd-torchad_nn-cifar.lua:48)
    B, nb = a.batch_per_gpu, a.pool_batches
    total = B * nb * P
    data, labels = SyntheticCifar(total, "data"), SyntheticCifar(total, "labels")

    if a.impl == "tadnn":
        model = CifarCNN(conv2_mode=a.conv2).to(dev)
        d, t, n = tadnn.parallelize(data, labels, model, total, h, None, 1, verbose=False)
        trainer = tadnn.BatchTrainer(model, learningRate=a.lr)
        step = trainer.step
    else:
        from baseline.torch_nccl_baseline import BaselineTrainer
        trainer = BaselineTrainer(dev, world=P, lr=a.lr, sync="per_tensor")
        d, t = data[rank * B * nb:(rank + 1) * B * nb], labels[rank * B * nb:(rank + 1) * B * nb]
        n = B * nb
        step = trainer.step
    assert n == B * nb
    d_dev, t_dev = d.to(dev), t.to(dev)  # 8 x 25 MB device-resident batches: larger than L2

    def run(k0, k):
        for s in range(k0, k0 + k):
            i = (s % nb) * B
            step(d_dev[i:i + B], t_dev[i:i + B])

    def barrier():
        torch.cuda.synchronize()
        if P > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- device-timed kernel-path number --------------------------------------------------
    run(0, max(3, a.warmup))
    barrier()
    sampler = ClockSampler(torch.cuda.current_device()).start() if rank == 0 else None
    launches0 = _native.launch_count() if a.impl == "tadnn" else 0
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    run(a.warmup, a.steps)
    e1.record()
    barrier()
    launches = (_native.launch_count() - launches0) if a.impl == "tadnn" else 0
    table = _native.launch_table() if a.impl == "tadnn" else {}
    if sampler:
        sampler.stop()
    ms = torch.tensor([e0.elapsed_time(e1)], device=dev, dtype=torch.float64)
    if P > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms = ms.item()
    value = a.steps * B * P / (ms / 1e3)

    # ---- end-to-end through the public API: pinned host -> H2D every step, loss D2H every step
    e2e = None
    if not a.no_e2e and a.impl == "tadnn":
        for _ in trainer.fit(d, t, B, max(3, a.warmup)):
            pass
        barrier()
        e0.record()
        last = None
        for last in trainer.fit(d, t, B, a.steps, start=a.warmup):
            pass
        e1.record()
        barrier()
        ms2 = torch.tensor([e0.elapsed_time(e1)], device=dev, dtype=torch.float64)
        if P > 1:
            dist.all_reduce(ms2, op=dist.ReduceOp.MAX)
        e2e = {"value": a.steps * B * P / (ms2.item() / 1e3), "unit": "samples/s",
               "h2d_bytes_per_step": B * (3 * 32 * 32 * 2 + 8), "d2h_bytes_per_step": 4,
               "ms_per_step": ms2.item() / a.steps, "last_loss": last}

    if rank == 0:
        out = {
            "metric": METRIC, "value": value, "unit": "samples/s", "n_gpus": P, "steps": a.steps,
            "warmup": max(3, a.warmup), "ms_per_step": ms / a.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "impl": a.impl,
            "config": {"model": MODEL, "global_batch": B * P, "per_gpu_batch": B, "seq_len": None,
                       "input": "3x32x32 bf16", "parallelism": f"dp{P}",
                       "sync": "params+grads all-reduce-average every step (batchSize=1)",
                       "conv2_forward": a.conv2, "optimizer": "SGD", "l2_policy": f"inputs cycle through {nb} device batches = "
                       f"{nb * B * 3072 * 2 / 1e6:.0f} MB per GPU (> 126 MB L2)"},
            "clocks": sampler.summary() if sampler else None,
            "gpu_launches": launches,
            "gpu_launch_table": table,
        }
        if e2e is not None:
            out["e2e"] = e2e
        print(json.dumps(out))
    if h is not None:
        h.stop()
    return 0


if __name__ == "__main__":
    sys.exit(main())
