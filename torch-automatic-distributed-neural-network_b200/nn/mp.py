"""Drop-in model-parallel layers: ``MPInitialLinear``, ``MPBaseLinear``, ``MPTanh``,
``MPInitialReshape``, ``MPBaseReshape`` (reference: nodemodule.lua:34-285).

Layout (identical to the reference, and what ``.t7`` I/O sees): every rank holds
``weight [o, i/P]`` (a slice of the *input* dimension) and a full ``bias [o]``; activations are
replicated; the layer output is the **sum over ranks** of the partial products
(nodemodule.lua:40-54,93-105), so the effective dense layer is
``W = cat_r(W_r, dim=1)``, ``b = sum_r b_r``.

Backward comes in two flavours:

* ``exact`` (default) — a true gradient: the slice gradient ``g·W_r`` is **all-gathered** along
  the last dim (the reference's own experimental ``syncTanh``/``syncReshape`` path,
  nodemodule.lua:165-167,265-267); no reduce of gradInput.
* ``compat`` (``compat=True`` or ``TADNN_COMPAT_BACKWARD=1``) — reproduces the reference's
  arithmetic: gradInput slices are sum-all-reduced (nodemodule.lua:66,117 — SURVEY Q1) and
  ``MPTanh`` tiles its own slice P times instead of gathering (nodemodule.lua:161-164 — Q2).

The communication handle is the process-wide "global ``mpi``" (nodemodule.lua:18-20), i.e.
``tadnn.mpi.current()``, unless one is passed explicitly.

On B200 the hot path is not these per-op modules but the fused kernels: ``fuse_mp(model)``
rewrites an MP ``Sequential`` into ``tadnn.parallel.tp.FusedMPMLP`` (tcgen05 GEMM with the
collective and the tanh in its epilogue).
"""
from __future__ import annotations

import math
import os
from typing import Optional

import torch
import torch.nn.functional as F

from .. import mpi as _mpi
from .protocol import T7Protocol


def _env_flag(name: str) -> bool:
    return os.environ.get(name, "0").lower() in ("1", "true", "yes", "on")


def _print_dims() -> bool:  # nodemodule.lua:3 `printDims`
    return _env_flag("TADNN_PRINT_DIMS")


def _compat_default() -> bool:
    return _env_flag("TADNN_COMPAT_BACKWARD")


def _handle(mpi=None):
    return mpi if mpi is not None else _mpi.current()


def narrowInput(input: torch.Tensor, mpi=None) -> torch.Tensor:
    """This rank's slice of the LAST dim (nodemodule.lua:16-21); a strided view, no copy."""
    h = _handle(mpi)
    n = input.shape[-1]
    P = h.size()
    assert n % P == 0, f"last dim {n} not divisible by world size {P}"
    w = n // P
    return input.narrow(input.dim() - 1, h.rank() * w, w)


# ----------------------------------------------------------------------------
# autograd plumbing
# ----------------------------------------------------------------------------
class _AllReduceSum(torch.autograd.Function):
    """y = sum over ranks (forward); identity (backward): the replicated loss sees every rank's
    partial product with weight one."""

    @staticmethod
    def forward(ctx, x, h):
        y = x.contiguous().clone()
        h.allreduceTensor(y)
        return y

    @staticmethod
    def backward(ctx, g):
        return g, None


class _Narrow(torch.autograd.Function):
    """Forward: this rank's last-dim slice.  Backward: ``exact`` all-gathers the slice gradients
    into the full-width gradient; ``compat`` sum-all-reduces the slice (Q1) and tiles it P times
    (the tiling is what MPTanh/MPBaseReshape do in the reference, Q2)."""

    @staticmethod
    def forward(ctx, x, h, compat):
        ctx.h, ctx.compat = h, compat
        return narrowInput(x, h)

    @staticmethod
    def backward(ctx, g):
        h, P = ctx.h, ctx.h.size()
        if P == 1:
            return g, None, None
        g = g.contiguous()
        if ctx.compat:
            g = g.clone()
            h.allreduceTensor(g)
            return torch.cat([g] * P, dim=-1), None, None
        out = torch.empty(g.shape[:-1] + (g.shape[-1] * P,), dtype=g.dtype, device=g.device)
        h.allgatherTensor(g, out)
        return out, None, None


class _TanhCompat(torch.autograd.Function):
    """Reference ``MPTanh`` backward (nodemodule.lua:145-174): ``gradInput_slice = gradOutput_slice
    * (1 - y_r^2)`` on this rank's slice of the output, tiled P times."""

    @staticmethod
    def forward(ctx, x, h):
        y = torch.tanh(x)
        ctx.h = h
        ctx.save_for_backward(y)
        return y

    @staticmethod
    def backward(ctx, g):
        (y,) = ctx.saved_tensors
        h, P = ctx.h, ctx.h.size()
        yr = narrowInput(y, h)
        gs = g[..., : yr.shape[-1]]  # every slice of the incoming (tiled) gradient is the same
        gi = gs * (1.0 - yr * yr)
        return torch.cat([gi] * P, dim=-1), None


# ----------------------------------------------------------------------------
# layers
# ----------------------------------------------------------------------------
class _MPLinear(torch.nn.Module, T7Protocol):
    t7_class = "nn.Linear"

    def __init__(self, inputSize: int, outputSize: int, mpi=None, compat: Optional[bool] = None,
                 init: str = "reference", device=None, dtype=None):
        super().__init__()
        h = _handle(mpi)
        P = h.size()
        assert inputSize % P == 0, f"i={inputSize} not divisible by {P}"  # nodemodule.lua:36,88
        self._mpi = mpi
        self.in_features = inputSize
        self.out_features = outputSize
        self.world = P
        self.shard = inputSize // P
        self.compat = _compat_default() if compat is None else bool(compat)
        kw = {"device": device, "dtype": dtype}
        self.weight = torch.nn.Parameter(torch.empty(outputSize, self.shard, **kw))
        self.bias = torch.nn.Parameter(torch.empty(outputSize, **kw))
        self.reset(init)

    def reset(self, init: str = "reference"):
        """``reference``: U(±1/sqrt(i/P)) for weight and bias on every rank — the fan-in the
        reference ends up with because it constructs ``nn.Linear(i/P, o)`` (nodemodule.lua:37,90;
        SURVEY Q5).  ``dense_equivalent``: U(±1/sqrt(i)), bias drawn on rank 0 only so that the
        summed bias matches a dense layer's."""
        fan = self.shard if init == "reference" else self.in_features
        stdv = 1.0 / math.sqrt(fan)
        with torch.no_grad():
            self.weight.uniform_(-stdv, stdv)
            self.bias.uniform_(-stdv, stdv)
            if init == "dense_equivalent" and _handle(self._mpi).rank() != 0:
                self.bias.zero_()
        return self

    def forward(self, input: torch.Tensor) -> torch.Tensor:
        h = _handle(self._mpi)
        x = _Narrow.apply(input, h, self.compat) if (input.requires_grad and h.size() > 1) \
            else narrowInput(input, h)
        from ..ops import native_nn
        y = native_nn.linear(x, self.weight, self.bias)   # CUDA: tcgen05 GEMM (strided slice = TMA pitch), else F.linear
        if _print_dims():
            print(f"#{type(self).__name__} UPDATE OUTPUT# input {tuple(x.shape)} output {tuple(y.shape)}")
        if h.size() > 1:
            y = _AllReduceSum.apply(y, h)
        return y

    def extra_repr(self):
        return f"{self.in_features} -> {self.out_features} (shard {self.shard} of {self.world})"

    # ---- dense <-> sharded conversion -----------------------
    @torch.no_grad()
    def load_dense(self, weight: torch.Tensor, bias: Optional[torch.Tensor], mpi=None):
        """Take this rank's slice of a dense ``[o, i]`` weight; bias goes to rank 0 (zeros
        elsewhere) so that the summed bias equals the dense bias."""
        h = _handle(mpi or self._mpi)
        r = h.rank()
        self.weight.copy_(weight[:, r * self.shard:(r + 1) * self.shard])
        if bias is not None:
            if r == 0:
                self.bias.copy_(bias)
            else:
                self.bias.zero_()
        return self

    @torch.no_grad()
    def dense_weight_bias(self, mpi=None):
        """All-gather the shards: returns dense ``(W [o, i], b [o])`` on every rank."""
        h = _handle(mpi or self._mpi)
        P = h.size()
        w = self.weight.detach()
        full = torch.empty(w.shape[0], w.shape[1] * P, dtype=w.dtype, device=w.device)
        h.allgatherTensor(w.contiguous(), full)
        b = self.bias.detach().clone()
        h.allreduceTensor(b)
        return full, b


class MPInitialLinear(_MPLinear):
    """First FC layer of a model-parallel stack (nodemodule.lua:34-73)."""
    t7_name = "nn.MPInitialLinear"


class MPBaseLinear(_MPLinear):
    """Any later FC layer of a model-parallel stack (nodemodule.lua:86-119)."""
    t7_name = "nn.MPBaseLinear"


class MPTanh(torch.nn.Module, T7Protocol):
    """``tanh`` on the full replicated activation (nodemodule.lua:132-143); backward per mode."""
    t7_name = "nn.MPTanh"

    def __init__(self, mpi=None, compat: Optional[bool] = None):
        super().__init__()
        self._mpi = mpi
        self.compat = _compat_default() if compat is None else bool(compat)

    def forward(self, input):
        h = _handle(self._mpi)
        if self.compat and h.size() > 1 and input.requires_grad:
            return _TanhCompat.apply(input, h)
        y = torch.tanh(input)
        if _print_dims():
            print(f"#TANH UPDATE OUTPUT# input {tuple(input.shape)} output {tuple(y.shape)}")
        return y


class _MPReshape(torch.nn.Module, T7Protocol):
    def __init__(self, *size, batchMode: Optional[bool] = None, mpi=None):
        super().__init__()
        if len(size) == 1 and isinstance(size[0], (tuple, list, torch.Size)):
            size = tuple(size[0])
        self.size = tuple(int(s) for s in size)
        self.nelement = 1
        for s in self.size:
            self.nelement *= s
        self.batchMode = batchMode
        self._mpi = mpi

    def forward(self, x):  # nodemodule.lua:188-206 / 240-258 (copy of nn.Reshape.updateOutput)
        if self.batchMode is False or (self.batchMode is None and x.numel() == self.nelement and x.shape[0] != 1):
            return x.reshape(self.size)
        return x.reshape((x.shape[0],) + self.size)

    def extra_repr(self):
        return "x".join(str(s) for s in self.size)


class MPInitialReshape(_MPReshape):
    """Reshape at the top of an MP stack (nodemodule.lua:186-226).  In the reference its
    backward only re-views the (partial) slice gradient coming up from ``MPInitialLinear``; with
    full-width exact gradients it is a plain reshape."""
    t7_name = "nn.MPInitialReshape"


class MPBaseReshape(_MPReshape):
    """Reshape below other layers feeding an ``MPBaseLinear`` (nodemodule.lua:238-285).  The
    reference tiles/gathers the slice gradient here; that step lives in the linear layer's
    narrow-backward in this implementation, so this is a plain reshape."""
    t7_name = "nn.MPBaseReshape"


def fuse_mp(model, **kw):
    """Rewrite an MP ``Sequential`` (``MPInitialReshape?``, ``MPInitialLinear``, then
    ``MPTanh``/``MPBaseLinear`` pairs) into the fused B200 module
    (``tadnn.parallel.tp.FusedMPMLP``): same parameters, tcgen05 GEMMs with the collective
    and the tanh fused into the epilogue."""
    from ..parallel.tp import FusedMPMLP
    return FusedMPMLP.from_sequential(model, **kw)
