"""Data-parallel synchronisation state: the counter/firing rule of the reference's hooks and
the machinery that makes the sync one fused, overlapped kernel on B200.

Reference behaviour kept (datamodule.lua:96-110,146-158): a counter starts at 1, the model is
synchronised when ``counter % batchSize == 0 or counter == size`` (size = shard length), then
the counter is incremented; it counts *calls*, never resets.  What changes is the
hooking mechanism — the reference redefines ``nn.Sequential:backward`` and
``nn.StochasticGradient:train`` class-wide and permanently; here ``optimize_sync`` attaches a
``SyncState`` to *the model* (``model._tadnn_sync``) with a ``remove()``, and three entry points
consult it: ``tadnn.nn.Sequential.backward`` (Torch7 style), ``StochasticGradient.train`` and
plain autograd ``loss.backward()`` (post-accumulate-grad hooks + an end-of-backward callback).

Overlap: on a step that will synchronise, each gradient bucket is all-reduced on a side stream
as soon as its last gradient has been accumulated; the parameters are averaged on that stream
once backward has finished (they are still being read by the dgrad kernels until then, and
in "model" mode with batchSize > 1 the replicas differ); the compute stream waits for the side
stream at the end of backward.
"""
from __future__ import annotations

import os
from typing import List, Optional

import torch

from . import symm as _symm


class SyncState:
    def __init__(self, model: torch.nn.Module, mpi, mpinn, batchSize: int, size: int,
                 sync_mode: Optional[str] = None, overlap: Optional[bool] = None,
                 bucket_mb: float = 8.0):
        self.model = model
        self.mpi = mpi
        self.mpinn = mpinn
        self.batchSize = int(batchSize)
        self.size = int(size)
        # syncPrototype=true -> params+grads (datamodule.lua:4,103-107); false -> grads only
        if sync_mode is None:
            sync_mode = "grads" if os.environ.get("TADNN_SYNC_PROTOTYPE", "1") in ("0", "false") else "model"
        assert sync_mode in ("model", "grads")
        self.sync_mode = sync_mode
        self.syncs = 0
        self._handles: List = []
        self._in_backward = False
        self._will_sync = False
        self._pending: List = []
        self._ready = 0
        self.flat = getattr(model, "_tadnn_flat", None)
        use_overlap = self.flat is not None and self.flat.symmetric and self.flat.ctx.world > 1
        if overlap is None:
            env = os.environ.get("TADNN_DP_OVERLAP")
            if env is not None:
                overlap = env not in ("0", "false")
            else:
                # small models: one fused launch at the end of backward is cheaper than the stream
                # fork/join and the SM interference of an overlapped sync (measured: 114K-param CNN,
                # 8 GPUs, 1.203 ms/step without vs 1.227 ms/step with overlap)
                nbytes = sum(g.half * g.region.element_size() for g in self.flat.groups.values()) if self.flat else 0
                overlap = nbytes >= (4 << 20)
        self.overlap = bool(overlap) and use_overlap
        self.comm_stream = torch.cuda.Stream() if self.overlap else None
        self._fused_lo = None
        self.bucket_bytes = int(bucket_mb * (1 << 20))
        self._buckets = self._make_buckets() if self.overlap else []

    # ------------------------------------------------------------------ firing rule
    def should_sync(self, counter: int) -> bool:
        return (counter % self.batchSize == 0) or (counter == self.size)

    def after_backward(self, owner) -> bool:
        """Counter logic of the hooked ``Sequential:backward``; the counter lives on ``owner``
        (whichever module's backward ran) but the captured top-level model is synchronised."""
        c = getattr(owner, "sync_counter", None)
        if c is None:
            c = 1
        fired = self.should_sync(c)
        if fired:
            self.sync()
        owner.sync_counter = c + 1
        return fired

    def sync(self):
        from .. import datamodule
        if self.sync_mode == "model":
            datamodule.synchronizeModel(self.model, self.mpi)
        else:
            self.mpinn.synchronizeGradients(self.model, self.mpi, average=True)
        self.syncs += 1

    # ------------------------------------------------------------------ fused sync + SGD
    def fused_step_ok(self) -> bool:
        """True when a synchronised SGD step can be ONE kernel per gradient range (all-reduce-average +
        update + gradient clear, csrc/comm.cu ``allreduce_sgd_*``): flat fp32 buffer in the symmetric
        heap, more than one rank, a sync on every step (``batchSize == 1``: the replicas then always
        hold identical parameters, so the parameter half of the reference's message,
        datamodule.lua:214-218, is the identity), and a power-of-two world (sum/P of identical floats
        is then exact, i.e. skipping it changes no bit)."""
        f = self.flat
        if f is None or not f.symmetric or os.environ.get("TADNN_DP_FUSED_SGD", "1") in ("0", "false"):
            return False
        P = f.ctx.world
        return (P > 1 and (P & (P - 1)) == 0 and self.batchSize == 1 and len(f.groups) == 1
                and next(iter(f.groups)) == torch.float32)

    def fused_tail(self, first_param, lr: float, wd: float) -> None:
        """Called from inside the backward pass as soon as the gradients of ``first_param`` and of every
        parameter after it (flat order) are complete: their synchronised update runs on the side
        stream while the rest of backward is still computing."""
        g = next(iter(self.flat.groups.values()))
        _g, lo = self.flat._where[id(first_param)]
        lo = lo // 64 * 64  # 256-byte granule (parameters are 256-byte aligned in the flat buffer)
        if self.comm_stream is None:
            self.comm_stream = torch.cuda.Stream()
        ev = torch.cuda.current_stream().record_event()
        self.comm_stream.wait_event(ev)
        P = self.flat.ctx.world
        self.flat.ctx.allreduce_sgd_offset(g.heap_off + (g.half + lo) * 4, g.heap_off + lo * 4, g.half - lo, 1.0 / P, lr, wd,
                                           0, _symm.CH_DP, self.comm_stream)
        self._fused_lo = lo

    def fused_finish(self, owner, lr: float, wd: float) -> None:
        """End of backward: the remaining range on the compute stream, then join the side stream.
        Same counter bookkeeping as ``after_backward``."""
        g = next(iter(self.flat.groups.values()))
        hi = getattr(self, "_fused_lo", None)
        hi = g.half if hi is None else hi
        P = self.flat.ctx.world
        if hi > 0:
            self.flat.ctx.allreduce_sgd_offset(g.heap_off + g.half * 4, g.heap_off, hi, 1.0 / P, lr, wd, 0, _symm.CH_DEFAULT, None)
        if self._fused_lo is not None:
            torch.cuda.current_stream().wait_stream(self.comm_stream)
        self._fused_lo = None
        c = getattr(owner, "sync_counter", None) or 1
        owner.sync_counter = c + 1
        self.syncs += 1

    # ------------------------------------------------------------------ autograd path
    def install_autograd_hooks(self):
        for p in self.model.parameters():
            if p.requires_grad:
                self._handles.append(p.register_post_accumulate_grad_hook(self._on_grad))
        return self

    def remove(self):
        for h in self._handles:
            h.remove()
        self._handles = []
        if getattr(self.model, "_tadnn_sync", None) is self:
            try:
                del self.model._tadnn_sync
            except AttributeError:
                pass

    def _make_buckets(self):
        """Reverse-parameter-order buckets of ~bucket_bytes inside each dtype group's grads half:
        (dtype, heap byte offset, element count, [param ids])."""
        buckets = []
        for g in self.flat.groups.values():
            es = g.region.element_size()
            hi = g.half
            ids: List[int] = []
            for i in range(len(g.params) - 1, -1, -1):
                ids.append(id(g.params[i]))
                lo = g.offsets[i]
                if (hi - lo) * es >= self.bucket_bytes or i == 0:
                    lo = 0 if i == 0 else lo
                    lo_al = lo // 4 * 4 if es == 4 else lo // 8 * 8  # keep 16-byte alignment
                    buckets.append({"dtype": g.dtype, "off": g.heap_off + (g.half + lo_al) * es,
                                    "count": hi - lo_al, "ids": set(ids), "left": 0})
                    hi = lo_al
                    ids = []
        return buckets

    def _on_grad(self, p):
        if not self._in_backward:
            self._in_backward = True
            owner = self.model
            c = getattr(owner, "sync_counter", None) or 1
            self._will_sync = self.should_sync(c)
            torch.autograd.Variable._execution_engine.queue_callback(self._backward_done)
            if self.overlap and self._will_sync:
                for b in self._buckets:
                    b["left"] = len(b["ids"])
                ev = torch.cuda.current_stream().record_event()
                self.comm_stream.wait_event(ev)
        if self.overlap and self._will_sync:
            if self.flat is not None and p.grad is not None:
                gv = self.flat.grad_view(p)
                if p.grad.data_ptr() != gv.data_ptr():
                    with torch.no_grad():
                        gv.copy_(p.grad)
                        p.grad = gv
            pid = id(p)
            for b in self._buckets:
                if pid in b["ids"]:
                    b["left"] -= 1
                    if b["left"] == 0:
                        ev = torch.cuda.current_stream().record_event()
                        self.comm_stream.wait_event(ev)
                        P = self.flat.ctx.world
                        self.flat.ctx.allreduce_offset(b["off"], b["count"], b["dtype"], 1.0 / P, 0, 0,
                                                       _symm.CH_DP, self.comm_stream)
                    break

    def _backward_done(self):
        self._in_backward = False
        owner = self.model
        c = getattr(owner, "sync_counter", None) or 1
        if self.overlap and self._will_sync:
            # any bucket that never completed (unused params) is reduced now
            P = self.flat.ctx.world
            for b in self._buckets:
                if b["left"] > 0:
                    ev = torch.cuda.current_stream().record_event()
                    self.comm_stream.wait_event(ev)
                    self.flat.ctx.allreduce_offset(b["off"], b["count"], b["dtype"], 1.0 / P, 0, 0,
                                                   _symm.CH_DP, self.comm_stream)
                    b["left"] = 0
            if self.sync_mode == "model":
                # Parameters are averaged only now: the rest of backward was still reading W on the
                # compute stream (dX = dY·W), and the replicas of a "model"-mode run with
                # batchSize > 1 have diverged, so reducing them in place any earlier would feed a
                # mix of local and averaged weights into the gradients.  The reference also
                # averages after backward has finished (datamodule.lua:96-107).
                ev = torch.cuda.current_stream().record_event()
                self.comm_stream.wait_event(ev)
                self.flat.allreduce("params", 1.0 / P, stream=self.comm_stream, channel=_symm.CH_DP)
            torch.cuda.current_stream().wait_stream(self.comm_stream)
            self.syncs += 1
            owner.sync_counter = c + 1
        else:
            if self.flat is not None:
                self.flat.rehome_grads()
            self.after_backward(owner)
        self._will_sync = False
