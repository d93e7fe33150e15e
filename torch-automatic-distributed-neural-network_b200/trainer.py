"""Batched trainer: the reference's per-sample ``StochasticGradient`` loop
(datamodule.lua:117-184 — forward, criterion, backward, in-place SGD step, periodic sync) at
mini-batch granularity, which is what a B200 needs to be busy.

One ``step(x, y)`` = forward + loss + backward (the ``parallelize()`` hook fires inside
``backward`` and runs the fused all-reduce-average) + one fused SGD kernel over the flat
parameter buffer.  ``fit`` streams mini-batches from (pinned) host memory with a copy stream
and double buffering, and reads every step's loss back to the host one step late so the
device never idles on the read.
"""
from __future__ import annotations

import os
from typing import Callable, Iterator, Optional

import torch


class BatchTrainer:
    def __init__(self, model: torch.nn.Module, learningRate: float = 1e-3, weightDecay: float = 0.0,
                 criterion: Optional[Callable] = None):
        self.model = model
        self.learningRate = float(learningRate)
        self.weightDecay = float(weightDecay)
        self.criterion = criterion
        self.steps = 0
        self._params = [p for p in model.parameters() if p.requires_grad]
        self._cuda = any(p.is_cuda for p in self._params)
        self._copy_stream = torch.cuda.Stream() if self._cuda else None
        self._nvtx = self._cuda and os.environ.get("TADNN_NVTX", "0") not in ("0", "false")

    # ------------------------------------------------------------------ one step
    def _loss(self, x, y):
        if self.criterion is None and hasattr(self.model, "loss"):
            return self.model.loss(x, y)
        out = self.model(x)
        crit = self.criterion
        if crit is None:
            from .nn import CrossEntropyCriterion
            crit = self.criterion = CrossEntropyCriterion()
        return crit(out, y)

    def _sgd(self):
        flat = getattr(self.model, "_tadnn_flat", None)
        lr, wd = self.learningRate, self.weightDecay
        if flat is not None and self._cuda and all(dt == torch.float32 for dt in flat.groups):
            from .ops.cnn import sgd_step
            for g in flat.groups.values():
                sgd_step(g.region[:g.half], g.region[g.half:], lr, wd, zero_grad=True)
            return
        with torch.no_grad():
            for p in self._params:
                if p.grad is None:
                    continue
                if wd:
                    p.grad.add_(p, alpha=wd)
                p.add_(p.grad, alpha=-lr)
                p.grad.zero_()

    def step(self, x: torch.Tensor, y: torch.Tensor) -> torch.Tensor:
        """Returns the (device) loss tensor of this step; does not synchronise.  With
        ``TADNN_NVTX=1`` the step is wrapped in NVTX ranges for Nsight timelines."""
        if self._nvtx:
            torch.cuda.nvtx.range_push(f"tadnn.step {self.steps}")
            try:
                return self._step(x, y)
            finally:
                torch.cuda.nvtx.range_pop()
        return self._step(x, y)

    def _step(self, x: torch.Tensor, y: torch.Tensor) -> torch.Tensor:
        if self.criterion is None and self._cuda and hasattr(self.model, "forward_backward"):
            # explicit forward+backward (no autograd), then the hooked-backward bookkeeping
            loss = self.model.forward_backward(x, y)
            st = getattr(self.model, "_tadnn_sync", None)
            if st is not None:
                st.after_backward(self.model)
        else:
            loss = self._loss(x, y)
            loss.backward()  # parallelize()'s hook synchronises the model in here
        self._sgd()
        self.steps += 1
        return loss.detach()

    # ------------------------------------------------------------------ streaming from the host
    def fit(self, data: torch.Tensor, targets: torch.Tensor, batch: int, steps: int,
            start: int = 0) -> Iterator[float]:
        """Run ``steps`` steps over ``data``/``targets`` (host tensors, ideally pinned; wraps
        around), yielding each step's loss as a Python float (read one step late).  Per step:
        one H2D copy of that step's inputs and one D2H read of its loss."""
        n = data.shape[0]
        nb = max(1, n // batch)
        if not self._cuda:
            for s in range(steps):
                i = ((start + s) % nb) * batch
                yield float(self.step(data[i:i + batch], targets[i:i + batch]))
            return
        dev = self._params[0].device
        cs = self._copy_stream
        main = torch.cuda.current_stream()
        bufs = [(torch.empty((batch,) + tuple(data.shape[1:]), dtype=data.dtype, device=dev),
                 torch.empty((batch,) + tuple(targets.shape[1:]), dtype=targets.dtype, device=dev)) for _ in range(2)]
        ready = [torch.cuda.Event(), torch.cuda.Event()]
        consumed = [torch.cuda.Event(), torch.cuda.Event()]
        host_loss = [torch.zeros((), dtype=torch.float32).pin_memory() for _ in range(2)]
        loss_ev = [torch.cuda.Event(), torch.cuda.Event()]

        def issue(s):
            k = s & 1
            i = ((start + s) % nb) * batch
            with torch.cuda.stream(cs):
                if s >= 2:
                    cs.wait_event(consumed[k])
                bufs[k][0].copy_(data[i:i + batch], non_blocking=True)
                bufs[k][1].copy_(targets[i:i + batch], non_blocking=True)
                ready[k].record(cs)

        issue(0)
        for s in range(steps):
            k = s & 1
            if s + 1 < steps:
                issue(s + 1)
            main.wait_event(ready[k])
            loss = self.step(bufs[k][0], bufs[k][1])
            consumed[k].record(main)
            host_loss[k].copy_(loss, non_blocking=True)
            loss_ev[k].record(main)
            if s >= 1:
                loss_ev[k ^ 1].synchronize()
                yield float(host_loss[k ^ 1])
        if steps >= 1:
            k = (steps - 1) & 1
            loss_ev[k].synchronize()
            yield float(host_loss[k])
        self.h2d_bytes_per_step = batch * (data[0].numel() * data.element_size() + targets[0].numel() * targets.element_size())
        self.d2h_bytes_per_step = 4
