"""Flat fusion buffer: every parameter and every gradient of a model as views into ONE
contiguous region per dtype, laid out ``[params | grads]``.

The reference synchronises a model with 2·N blocking per-tensor collectives plus 2·N ``div``
kernels (datamodule.lua:211-224; N = number of parameter tensors, messages as small as 10
floats).  With the tensors living in one flat region the same semantics are ONE fused kernel
launch (all-reduce + 1/P scale, csrc/comm.cu) — and when the region sits in the symmetric heap
(``tadnn.parallel.symm``) that kernel reads/writes peer memory directly, no staging copy.

Also provides Torch7's ``getParameters()`` flattening (``flat_params``/``flat_grads``).
"""
from __future__ import annotations

from typing import Dict, List, Optional

import torch

from . import symm as _symm

_ALIGN_BYTES = 256


class _Group:
    def __init__(self, dtype):
        self.dtype = dtype
        self.params: List[torch.nn.Parameter] = []
        self.offsets: List[int] = []   # element offsets inside the params half
        self.half = 0                  # elements in each half (padded)
        self.region: Optional[torch.Tensor] = None  # [2*half]
        self.heap_off: Optional[int] = None


class FlatBuffer:
    """Re-homes ``model``'s parameters/gradients into flat storage.

    ``ctx``: a ``SymmContext`` (GPU, world > 1 or forced) or ``None`` (plain device memory;
    collectives then go through the ``mpi`` handle on the flat tensors).
    """

    def __init__(self, model: torch.nn.Module, ctx: Optional[_symm.SymmContext], mpi=None):
        self.model = model
        self.ctx = ctx
        self.mpi = mpi
        self.groups: Dict[torch.dtype, _Group] = {}
        params = [p for p in model.parameters()]
        if not params:
            raise ValueError("model has no parameters")
        dev = params[0].device
        for p in params:
            if p.device != dev:
                raise ValueError("all parameters must live on one device")
            g = self.groups.setdefault(p.dtype, _Group(p.dtype))
            es = p.element_size()
            align = max(1, _ALIGN_BYTES // es)
            g.half = (g.half + align - 1) // align * align
            g.params.append(p)
            g.offsets.append(g.half)
            g.half += p.numel()
        use_heap = ctx is not None and dev.type == "cuda"
        for g in self.groups.values():
            es = torch.empty((), dtype=g.dtype).element_size()
            align = max(1, _ALIGN_BYTES // es)
            g.half = (g.half + align - 1) // align * align
            if use_heap and g.dtype in _symm._DT:
                g.heap_off = ctx.alloc(2 * g.half * es)
                g.region = ctx.tensor(g.heap_off, (2 * g.half,), g.dtype)
            else:
                g.region = torch.empty(2 * g.half, dtype=g.dtype, device=dev)
            g.region.zero_()
            with torch.no_grad():
                for p, off in zip(g.params, g.offsets):
                    pv = g.region[off:off + p.numel()].view(p.shape)
                    gv = g.region[g.half + off:g.half + off + p.numel()].view(p.shape)
                    pv.copy_(p.data)
                    if p.grad is not None:
                        gv.copy_(p.grad)
                    p.data = pv
                    p.grad = gv
        self._where = {id(p): (g, off) for g in self.groups.values() for p, off in zip(g.params, g.offsets)}
        model._tadnn_flat = self

    # ------------------------------------------------------------------ views
    @property
    def symmetric(self) -> bool:
        return self.ctx is not None and all(g.heap_off is not None for g in self.groups.values())

    def _single(self) -> _Group:
        if len(self.groups) != 1:
            raise ValueError("flat_params/flat_grads need a single-dtype model")
        return next(iter(self.groups.values()))

    def flat_params(self) -> torch.Tensor:
        g = self._single()
        return g.region[:g.half]

    def flat_grads(self) -> torch.Tensor:
        g = self._single()
        return g.region[g.half:]

    def grad_view(self, p: torch.nn.Parameter) -> torch.Tensor:
        g, off = self._where[id(p)]
        return g.region[g.half + off:g.half + off + p.numel()].view(p.shape)

    def rehome_grads(self) -> int:
        """Re-point ``p.grad`` at the flat views if user code replaced them (e.g.
        ``zero_grad(set_to_none=True)`` followed by a backward).  Returns #copies made."""
        n = 0
        views = self.__dict__.get("_grad_views")
        if views is None:   # built once: slicing + view per parameter is most of this function's cost on the per-step path
            views = [(p, g.region[g.half + off:g.half + off + p.numel()].view(p.shape))
                     for g in self.groups.values() for p, off in zip(g.params, g.offsets)]
            self.__dict__["_grad_views"] = views
        stale = [(p, gv) for p, gv in views if p.grad is not gv]
        if not stale:
            return 0
        with torch.no_grad():
            for p, gv in stale:
                if p.grad is None:
                    gv.zero_()
                    p.grad = gv
                elif p.grad.data_ptr() != gv.data_ptr():
                    gv.copy_(p.grad)
                    p.grad = gv
                    n += 1
                else:
                    p.grad = gv
        return n

    def zero_grads(self):
        for g in self.groups.values():
            g.region[g.half:].zero_()

    # ------------------------------------------------------------------ collectives
    def allreduce(self, what: str = "model", scale: float = 1.0, stream=None,
                  channel: int = _symm.CH_DP, algo: int = 0):
        """``what``: "model" (params and grads, one launch per dtype), "params" or "grads"."""
        assert what in ("model", "params", "grads")
        P = self.ctx.world if self.ctx is not None else (self.mpi.size() if self.mpi is not None else 1)
        for g in self.groups.values():
            lo, n = {"model": (0, 2 * g.half), "params": (0, g.half), "grads": (g.half, g.half)}[what]
            es = g.region.element_size()
            if g.heap_off is not None:
                if P == 1:
                    if scale != 1.0:
                        g.region[lo:lo + n].mul_(scale)
                    continue
                self.ctx.allreduce_offset(g.heap_off + lo * es, n, g.dtype, scale, algo, 0, channel, stream)
            else:
                t = g.region[lo:lo + n]
                if P > 1:
                    self.mpi.allreduceTensor(t)
                if scale != 1.0:
                    t.mul_(scale)

    def broadcast_params(self, root: int = 0, stream=None):
        for g in self.groups.values():
            if g.heap_off is not None and self.ctx.world > 1:
                self.ctx.broadcast_offset(g.heap_off, g.half * g.region.element_size(), root,
                                          _symm.CH_DP, stream)
            elif self.mpi is not None and self.mpi.size() > 1:
                self.mpi.broadcastTensor(root, g.region[:g.half])
