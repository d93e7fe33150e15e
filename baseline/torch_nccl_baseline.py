"""The comparator BASELINE.md names: a plain PyTorch (+cuDNN/cuBLAS) + NCCL restatement of the
reference's algorithm on the same box — none of tadnn's kernels on this path.

b1  sync="per_tensor": after backward, ``dist.all_reduce`` + ``div_`` of every parameter tensor,
    then of every gradient tensor, blocking, no overlap (datamodule.lua:211-224).
b2  sync="flat": one flat-bucket NCCL all-reduce of the gradients (DDP-style).
The model is the example CNN in bf16 autocast with fp32 master weights; SGD step via foreach.
"""
from __future__ import annotations

import torch
import torch.distributed as dist
import torch.nn.functional as F


class BaselineCNN(torch.nn.Module):
    def __init__(self):
        super().__init__()
        self.c1 = torch.nn.Conv2d(3, 64, 5)
        self.c2 = torch.nn.Conv2d(64, 64, 5)
        self.f1 = torch.nn.Linear(64, 100)
        self.f2 = torch.nn.Linear(100, 10)
        self.drop = torch.nn.Dropout(0.5)

    def forward(self, x):
        y = F.max_pool2d(F.relu(self.c1(x)), 3, 3)
        y = F.max_pool2d(F.relu(self.c2(y)), 3, 3).flatten(1)
        y = F.relu(self.f1(self.drop(y)))
        return F.log_softmax(self.f2(y).float(), dim=-1)


class BaselineTrainer:
    def __init__(self, device, world: int, lr: float = 1e-3, sync: str = "per_tensor"):
        self.model = BaselineCNN().to(device).to(memory_format=torch.channels_last)
        self.world, self.lr, self.sync = world, lr, sync
        self.params = list(self.model.parameters())
        if world > 1:
            for p in self.params:
                dist.broadcast(p.data, 0)

    def _sync(self):
        P = self.world
        if P == 1:
            return
        if self.sync == "per_tensor":
            for p in self.params:
                dist.all_reduce(p.data)
                p.data.div_(P)
            for p in self.params:
                dist.all_reduce(p.grad)
                p.grad.div_(P)
        else:
            flat = torch.cat([p.grad.reshape(-1) for p in self.params])
            dist.all_reduce(flat)
            flat.div_(P)
            off = 0
            for p in self.params:
                p.grad.copy_(flat[off:off + p.numel()].view_as(p))
                off += p.numel()

    def step(self, x, y):
        for p in self.params:
            p.grad = None
        with torch.autocast("cuda", dtype=torch.bfloat16):
            logp = self.model(x.contiguous(memory_format=torch.channels_last))
        loss = F.nll_loss(logp, y)
        loss.backward()
        self._sync()
        with torch.no_grad():
            torch._foreach_add_(self.params, [p.grad for p in self.params], alpha=-self.lr)
        return loss.detach()

    def fit(self, data, targets, batch, steps, start=0, affine=(4.0 / 255.0, -2.0)):
        """The same end-to-end loop the tadnn trainer is timed on, in plain PyTorch: every step
        copies its (uint8 or bf16) inputs and labels from pinned host memory on a side stream
        (double-buffered), widens them with torch ops, runs the step and reads the loss back."""
        dev = self.params[0].device
        nb = max(1, data.shape[0] // batch)
        cs = torch.cuda.Stream()
        main = torch.cuda.current_stream()
        bufs = [(torch.empty((batch,) + tuple(data.shape[1:]), dtype=data.dtype, device=dev),
                 torch.empty((batch,), dtype=targets.dtype, device=dev)) for _ in range(2)]
        ready = [torch.cuda.Event(), torch.cuda.Event()]
        done = [torch.cuda.Event(), torch.cuda.Event()]
        host_loss = [torch.zeros((), dtype=torch.float32).pin_memory() for _ in range(2)]
        lev = [torch.cuda.Event(), torch.cuda.Event()]

        def issue(s):
            k = s & 1
            i = ((start + s) % nb) * batch
            with torch.cuda.stream(cs):
                if s >= 2:
                    cs.wait_event(done[k])
                bufs[k][0].copy_(data[i:i + batch], non_blocking=True)
                bufs[k][1].copy_(targets[i:i + batch], non_blocking=True)
                ready[k].record(cs)

        issue(0)
        for s in range(steps):
            k = s & 1
            if s + 1 < steps:
                issue(s + 1)
            main.wait_event(ready[k])
            x = bufs[k][0]
            if x.dtype == torch.uint8:
                x = x.to(torch.bfloat16).mul_(affine[0]).add_(affine[1])
            loss = self.step(x, bufs[k][1])
            done[k].record(main)
            host_loss[k].copy_(loss, non_blocking=True)
            lev[k].record(main)
            if s >= 1:
                lev[k ^ 1].synchronize()
                yield float(host_loss[k ^ 1])
        if steps >= 1:
            k = (steps - 1) & 1
            lev[k].synchronize()
            yield float(host_loss[k])
