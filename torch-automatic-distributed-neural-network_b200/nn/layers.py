"""Stock Torch7 layers with Torch7 parameter layouts and initialisation (see tadnn.nn)."""
from __future__ import annotations

import math
from typing import Optional

import torch
import torch.nn.functional as F

from .protocol import T7Protocol


class Module(torch.nn.Module, T7Protocol):
    pass


class Sequential(torch.nn.Sequential, T7Protocol):
    """``nn.Sequential`` with ``add`` and the Torch7 protocol.  ``model.forward(x)`` /
    ``model(x)`` record ``output`` so that ``model.backward(x, gradOutput)`` works; the
    returned tensor still carries its autograd graph, so ``loss.backward()`` works too."""

    def add(self, module: torch.nn.Module) -> "Sequential":
        self.append(module)
        return self

    def get(self, i: int) -> torch.nn.Module:
        return self[i]

    def _run(self, x):
        fused = self._mp_fused(x)
        if fused is not None:
            from .. import _native
            try:
                return fused(x)
            except (AssertionError, _native.NativeError):  # a shape the fused kernels do not cover: per-layer path
                self.__dict__["_tadnn_mp_fused"] = None
        return torch.nn.Sequential.forward(self, x)

    def _mp_fused(self, x):
        """A stack of ``MP*`` layers on a GPU runs as the fused model-parallel module (``parallel/tp.py``:
        GEMM epilogues that push their tiles to the owner ranks, tanh in the reduction) — what
        ``fuse_mp(model)`` builds by hand, done once on the first CUDA forward.  Same parameters (shared,
        not copied).  Opt out per model with ``model._tadnn_no_autofuse = True`` or globally with
        ``TADNN_MP_AUTOFUSE=0``."""
        import os
        if not (torch.is_tensor(x) and x.is_cuda) or getattr(self, "_tadnn_no_autofuse", False):
            return None
        cached = self.__dict__.get("_tadnn_mp_fused", False)
        if cached is not False:
            return cached
        fused = None
        if os.environ.get("TADNN_MP_AUTOFUSE", "1") not in ("0", "false"):
            from .mp import MPBaseLinear, MPInitialLinear, MPInitialReshape, MPTanh
            mods = list(self)
            if mods and any(isinstance(m, MPInitialLinear) for m in mods) and \
                    all(isinstance(m, (MPInitialReshape, MPInitialLinear, MPBaseLinear, MPTanh)) for m in mods):
                try:
                    from .mp import _handle, fuse_mp
                    lin = next(m for m in mods if isinstance(m, MPInitialLinear))
                    h = _handle(lin._mpi)
                    if getattr(h, "symm", None) is not None:
                        fused = fuse_mp(self, mpi=h)
                except Exception:  # noqa: BLE001 - an unsupported stack simply stays on the per-layer path
                    fused = None
        self.__dict__["_tadnn_mp_fused"] = fused
        return fused

    def forward(self, input):  # noqa: D102
        return self._t7_forward(input)


# ----------------------------------------------------------------------------
# stock layers (Torch7 names / init)
# ----------------------------------------------------------------------------
class Linear(torch.nn.Linear, T7Protocol):
    """``nn.Linear(i, o)``: weight ``[o, i]``, bias ``[o]``, both U(±1/sqrt(i)) (Torch7 reset)."""

    def __init__(self, inputSize: int, outputSize: int, bias: bool = True, **kw):
        super().__init__(inputSize, outputSize, bias=bias, **kw)
        self.reset()

    def reset(self, stdv: Optional[float] = None):
        stdv = stdv if stdv is not None else 1.0 / math.sqrt(self.in_features)
        with torch.no_grad():
            self.weight.uniform_(-stdv, stdv)
            if self.bias is not None:
                self.bias.uniform_(-stdv, stdv)
        return self

    def forward(self, x):
        """CUDA: tcgen05 GEMM forward / dgrad / wgrad (ops/native_nn.py); elsewhere ``F.linear``."""
        from ..ops import native_nn
        return native_nn.linear(x, self.weight, self.bias)


class Tanh(torch.nn.Tanh, T7Protocol):
    pass


class ReLU(torch.nn.ReLU, T7Protocol):
    pass


class LogSoftMax(torch.nn.Module, T7Protocol):
    def forward(self, x):
        return F.log_softmax(x, dim=-1)


class Dropout(torch.nn.Dropout, T7Protocol):
    pass


class Reshape(torch.nn.Module, T7Protocol):
    """``nn.Reshape(d1, d2, ...[, batchMode])`` with Torch7's batch heuristic
    (nodemodule.lua:188-206): a leading batch dimension is kept when the input has more
    elements than ``prod(size)`` or its first dim is 1."""

    def __init__(self, *size, batchMode: Optional[bool] = None):
        super().__init__()
        if len(size) == 1 and isinstance(size[0], (tuple, list, torch.Size)):
            size = tuple(size[0])
        self.size = tuple(int(s) for s in size)
        self.nelement = 1
        for s in self.size:
            self.nelement *= s
        self.batchMode = batchMode

    def _target_shape(self, x: torch.Tensor):
        if self.batchMode is False or (self.batchMode is None and x.numel() == self.nelement and x.shape[0] != 1):
            return self.size
        return (x.shape[0],) + self.size

    def forward(self, x):
        return x.reshape(self._target_shape(x))

    def extra_repr(self):
        return "x".join(str(s) for s in self.size)


class View(torch.nn.Module, T7Protocol):
    """``nn.View(n)``: flattens to ``[B, n]`` (or ``[n]`` for a single sample)."""

    def __init__(self, *size):
        super().__init__()
        self.size = tuple(int(s) for s in size)
        self.nelement = 1
        for s in self.size:
            self.nelement *= s

    def forward(self, x):
        if x.numel() == self.nelement:
            return x.reshape(self.size)
        return x.reshape((-1,) + self.size)


class SpatialConvolutionMM(torch.nn.Module, T7Protocol):
    """``nn.SpatialConvolutionMM(nIn, nOut, kW, kH, dW, dH, padW, padH)`` with the Torch7
    parameter layout: ``weight [nOut, nIn*kH*kW]``, ``bias [nOut]``, U(±1/sqrt(kW*kH*nIn))."""

    def __init__(self, nInputPlane, nOutputPlane, kW, kH, dW=1, dH=1, padW=0, padH=None):
        super().__init__()
        self.nInputPlane, self.nOutputPlane = nInputPlane, nOutputPlane
        self.kW, self.kH, self.dW, self.dH = kW, kH, dW, dH
        self.padW = padW
        self.padH = padW if padH is None else padH
        self.weight = torch.nn.Parameter(torch.empty(nOutputPlane, nInputPlane * kH * kW))
        self.bias = torch.nn.Parameter(torch.empty(nOutputPlane))
        self.reset()

    def reset(self, stdv: Optional[float] = None):
        stdv = stdv if stdv is not None else 1.0 / math.sqrt(self.kW * self.kH * self.nInputPlane)
        with torch.no_grad():
            self.weight.uniform_(-stdv, stdv)
            self.bias.uniform_(-stdv, stdv)
        return self

    def forward(self, x):
        squeeze = x.dim() == 3
        if squeeze:
            x = x.unsqueeze(0)
        from ..ops import native_nn
        if native_nn.conv_ok(x, self.weight, self.kH, self.kW, self.dH, self.dW, self.padH, self.padW):
            # CUDA: im2col + tcgen05 GEMM; the result is NHWC in memory, NCHW in shape (channels_last strides)
            y = native_nn.Conv2dFn.apply(x, self.weight, self.bias, self.kH, self.kW)
            return y.squeeze(0) if squeeze else y
        w = self.weight.view(self.nOutputPlane, self.nInputPlane, self.kH, self.kW)
        y = F.conv2d(x, w.to(x.dtype), self.bias.to(x.dtype), stride=(self.dH, self.dW),
                     padding=(self.padH, self.padW))
        return y.squeeze(0) if squeeze else y


class SpatialMaxPooling(torch.nn.Module, T7Protocol):
    def __init__(self, kW, kH, dW=None, dH=None, padW=0, padH=0):
        super().__init__()
        self.kW, self.kH = kW, kH
        self.dW, self.dH = dW or kW, dH or kH
        self.padW, self.padH = padW, padH

    def forward(self, x):
        squeeze = x.dim() == 3
        if squeeze:
            x = x.unsqueeze(0)
        from ..ops import native_nn
        if native_nn.pool_ok(x, self.kH, self.kW, self.dH, self.dW, self.padH, self.padW):
            y = native_nn.MaxPoolFn.apply(x, self.kH)   # NHWC bf16 kernel with saved arg-max bytes
        else:
            y = F.max_pool2d(x, (self.kH, self.kW), (self.dH, self.dW), (self.padH, self.padW))
        return y.squeeze(0) if squeeze else y


