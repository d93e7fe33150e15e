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


_DP_TAIL = None if "TADNN_DP_TAIL" not in os.environ else os.environ["TADNN_DP_TAIL"] not in ("0", "false")


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
            st = getattr(self.model, "_tadnn_sync", None)
            if st is not None and st.fused_step_ok():
                # synchronised SGD step as fused all-reduce+update kernels: the range whose gradients are
                # complete early goes on the side stream from inside backward, the rest at the end
                lr, wd = self.learningRate, self.weightDecay
                # Early synchronisation of the finished gradients on a side stream while conv1's backward runs, or ONE
                # all-reduce + update at the end of backward: the second keeps the step's kernels back to back (dependent
                # launches, no event between them) and measured 3 us faster on two GPUs, where the whole gradient is a 12 us
                # pair exchange; with more ranks the message volume grows and the overlapped tail is kept.
                # TADNN_DP_TAIL=0 / 1 forces either.
                use_tail = _DP_TAIL if _DP_TAIL is not None else st.flat.ctx.world > 2
                tail = (lambda p: st.fused_tail(p, lr, wd)) if use_tail else None
                loss = self.model.forward_backward(x, y, tail_ready=tail)
                st.fused_finish(self.model, lr, wd)
                self.steps += 1
                return loss.detach()
            loss = self.model.forward_backward(x, y)
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
            start: int = 0, depth: int = 4, input_affine=None) -> Iterator[float]:
        """Run ``steps`` steps over ``data``/``targets`` (host tensors, ideally pinned; wraps
        around), yielding each step's loss as a Python float (read two steps late, so that the host stays ahead of the device).  Per step:
        one H2D copy of that step's inputs (and one of its labels) and one D2H read of its loss.

        ``data`` may be ``uint8`` (what CIFAR is on disk: 3 KB per image instead of 6 KB as bf16):
        the bytes are what crosses PCIe, and the device widens them to bf16 with the affine map
        ``input_affine = (scale[C], shift[C])`` (default: the model's ``input_affine`` attribute, else
        ``x/255``) in one small kernel (csrc/preprocess.cu: tadnn_u8_to_bf16; a model with its own input layout —
        ``CifarCNN.pack_input``, NHWC4 — has that kernel write the layout directly) before the step.
        ``depth`` staging buffers are kept in flight on the copy stream (the reference instead
        keeps its whole shard on the GPU, cifar_example/sgd-torchad_nn-cifar.lua:239-245)."""
        n = data.shape[0]
        nb = max(1, n // batch)
        if not self._cuda:
            for s in range(steps):
                i = ((start + s) % nb) * batch
                x = data[i:i + batch]
                if x.dtype == torch.uint8:
                    x = _widen_cpu(x, input_affine or getattr(self.model, "input_affine", None))
                yield float(self.step(x, targets[i:i + batch]))
            return
        dev = self._params[0].device
        cs = self._copy_stream
        main = torch.cuda.current_stream()
        depth = max(2, min(int(depth), max(2, steps)))
        u8 = data.dtype == torch.uint8
        if u8:
            from .ops.cnn import u8_to_bf16
            aff = input_affine or getattr(self.model, "input_affine", None)
            C = int(data.shape[1]) if data.dim() >= 3 else 1
            plane = int(data[0].numel() // C)
            if aff is None:
                aff = ([1.0 / 255.0] * C, [0.0] * C)
        bufs = [(torch.empty((batch,) + tuple(data.shape[1:]), dtype=data.dtype, device=dev),
                 torch.empty((batch,) + tuple(targets.shape[1:]), dtype=targets.dtype, device=dev)) for _ in range(depth)]
        # a model whose fused path has its own input layout (CifarCNN: NHWC4) gets it written by the widening kernel itself
        packer = getattr(self.model, "pack_input", None) if (u8 and getattr(self.model, "packs_input", False)) else None
        if packer is not None:
            wide = [packer(bufs[k][0], aff) for k in range(depth)]
        elif u8:  # one widened buffer per ring slot
            wide = [torch.empty((batch,) + tuple(data.shape[1:]), dtype=torch.bfloat16, device=dev) for _ in range(depth)]
        ready = [torch.cuda.Event() for _ in range(depth)]
        consumed = [torch.cuda.Event() for _ in range(depth)]
        LAG = 2   # a loss is read LAG steps after its step was queued: the host stays that far ahead of the device
        host_loss = [torch.zeros((), dtype=torch.float32).pin_memory() for _ in range(LAG + 1)]
        loss_ev = [torch.cuda.Event() for _ in range(LAG + 1)]
        loss_ready = [torch.cuda.Event() for _ in range(LAG + 1)]
        rs = self.__dict__.setdefault("_read_stream", torch.cuda.Stream())  # loss read-back: never blocks the compute stream
        dev_loss = [torch.zeros((), dtype=torch.float32, device=dev) for _ in range(LAG + 1)]

        ws = self.__dict__.setdefault("_widen_stream", torch.cuda.Stream()) if u8 else None
        copied = [torch.cuda.Event() for _ in range(depth)]
        widened = [torch.cuda.Event() for _ in range(depth)]

        set_stream = torch.cuda.set_stream

        def issue(s):
            k = s % depth
            i = ((start + s) % nb) * batch
            try:
                _issue(s, k, i)
            finally:
                set_stream(main)

        def _issue(s, k, i):
            set_stream(cs)
            if s >= depth:
                cs.wait_event(consumed[k])   # step s - depth has consumed this slot (its bytes were widened before that)
            bufs[k][0].copy_(data[i:i + batch], non_blocking=True)
            bufs[k][1].copy_(targets[i:i + batch], non_blocking=True)
            if not u8:
                ready[k].record(cs)
            else:
                copied[k].record(cs)
            if u8:
                # the widening kernel runs on its own stream: it may have to wait for SMs while the persistent compute
                # kernels are resident, and the copy stream must keep the PCIe link busy meanwhile
                set_stream(ws)
                ws.wait_event(copied[k])
                if s >= depth:
                    ws.wait_event(consumed[k])                     # wide[k] of step s - depth has been consumed
                if packer is not None:
                    packer(bufs[k][0], aff, out=wide[k])
                else:
                    u8_to_bf16(bufs[k][0], plane, C, aff[0], aff[1], out=wide[k])
                widened[k].record(ws)
                ready[k].record(ws)

        for s in range(min(depth - 1, steps)):
            issue(s)
        for s in range(steps):
            k = s % depth
            if s + depth - 1 < steps:
                issue(s + depth - 1)
            main.wait_event(ready[k])
            loss = self.step(wide[k] if u8 else bufs[k][0], bufs[k][1])
            consumed[k].record(main)
            j = s % (LAG + 1)
            dev_loss[j].copy_(loss)          # the step's loss tensor may be reused by the next step: keep a copy
            loss_ready[j].record(main)
            set_stream(rs)                   # (torch.cuda.stream() as a context manager costs ~8 us per use)
            try:
                rs.wait_event(loss_ready[j])
                host_loss[j].copy_(dev_loss[j], non_blocking=True)
                loss_ev[j].record(rs)
            finally:
                set_stream(main)
            if s >= LAG:
                jj = (s - LAG) % (LAG + 1)
                loss_ev[jj].synchronize()
                yield float(host_loss[jj])
        for s2 in range(max(0, steps - LAG), steps):
            jj = s2 % (LAG + 1)
            loss_ev[jj].synchronize()
            yield float(host_loss[jj])
        self.h2d_bytes_per_step = batch * (data[0].numel() * data.element_size() + targets[0].numel() * targets.element_size())
        self.d2h_bytes_per_step = 4


def _widen_cpu(x: torch.Tensor, aff):
    C = int(x.shape[1]) if x.dim() >= 3 else 1
    if aff is None:
        aff = ([1.0 / 255.0] * C, [0.0] * C)
    shape = [1, C] + [1] * (x.dim() - 2) if x.dim() >= 3 else [1] * x.dim()
    sc = torch.tensor(list(aff[0]), dtype=torch.float32).view(shape)
    sh = torch.tensor(list(aff[1]), dtype=torch.float32).view(shape)
    return x.float() * sc + sh
