"""``nn.SpatialSubtractiveNormalization`` / ``SpatialDivisiveNormalization`` / ``SpatialContrastiveNormalization``
with the arithmetic of Torch7 ``nn`` (the example uses the last one on the Y channel,
cifar_example/sgd-torchad_nn-cifar.lua:181-188,203-209).

Torch7 builds them from small networks; collapsed, for an input ``x [C, H, W]`` (or ``[N, C, H, W]``) and a
neighbourhood kernel ``k`` (2-D, or 1-D meaning the separable ``k kᵀ``):

* estimator ``E(x)  = conv2d(zero-pad(x), k / (sum(k) · C))``  summed over all C planes -> one plane
* border coefficient ``coef = E(ones)``  (1 in the interior, < 1 near the border)
* subtractive:  ``x - E(x) / coef``                      (the same local mean is taken off every plane)
* divisive:     ``x / T(sqrt(E(x²)) / coef)``  with ``T(s) = s if s > threshold else thresval``
  (``nn.Threshold(threshold, thresval)``; defaults ``threshold = 1e-4``, ``thresval = threshold``)
* contrastive = subtractive, then divisive.

Note the divisive border rule — the square root is divided by ``coef``, not taken of the corrected sum —
and that the threshold is element-wise, not a per-image mean.
"""
from __future__ import annotations

from typing import Optional

import torch
import torch.nn.functional as F

from .protocol import T7Protocol


def _kernel2d(kernel: torch.Tensor) -> torch.Tensor:
    k = torch.as_tensor(kernel, dtype=torch.float32)
    if k.dim() == 1:
        k = torch.outer(k, k)
    if k.dim() != 2 or k.shape[0] % 2 == 0 or k.shape[1] % 2 == 0:
        raise ValueError("the averaging kernel must be 1-D or 2-D with odd sizes")  # as Torch7 insists
    return k


def local_estimate(x: torch.Tensor, k2: torch.Tensor) -> torch.Tensor:
    """``E(x)`` for ``x [N, C, H, W]``: ``[N, 1, H, W]``."""
    C = x.shape[1]
    w = (k2 / (k2.sum() * C)).to(x).expand(1, C, *k2.shape)
    return F.conv2d(x, w, padding=(k2.shape[0] // 2, k2.shape[1] // 2))


class _SpatialNorm(torch.nn.Module, T7Protocol):
    def __init__(self, nInputPlane: int = 1, kernel: Optional[torch.Tensor] = None):
        super().__init__()
        self.nInputPlane = int(nInputPlane)
        self.register_buffer("kernel", _kernel2d(torch.ones(9, 9) if kernel is None else kernel))  # Torch7's default kernel

    def _batched(self, x: torch.Tensor):
        if x.dim() == 3:
            return x.unsqueeze(0), True
        if x.dim() != 4:
            raise ValueError("expected a [C, H, W] or [N, C, H, W] input")
        return x, False

    def _coef(self, x: torch.Tensor) -> torch.Tensor:
        return local_estimate(torch.ones_like(x[:1]), self.kernel)

    def extra_repr(self):
        return f"{self.nInputPlane} plane(s), kernel {tuple(self.kernel.shape)}"


class SpatialSubtractiveNormalization(_SpatialNorm):
    t7_name = "nn.SpatialSubtractiveNormalization"

    def forward(self, input):
        x, squeeze = self._batched(input)
        out = x - local_estimate(x, self.kernel) / self._coef(x)
        return out.squeeze(0) if squeeze else out


class SpatialDivisiveNormalization(_SpatialNorm):
    t7_name = "nn.SpatialDivisiveNormalization"

    def __init__(self, nInputPlane: int = 1, kernel: Optional[torch.Tensor] = None, threshold: Optional[float] = None,
                 thresval: Optional[float] = None):
        super().__init__(nInputPlane, kernel)
        self.threshold = 1e-4 if threshold is None else float(threshold)
        self.thresval = self.threshold if thresval is None else float(thresval)

    def forward(self, input):
        x, squeeze = self._batched(input)
        std = torch.sqrt(local_estimate(x * x, self.kernel)) / self._coef(x)
        std = torch.where(std > self.threshold, std, torch.full_like(std, self.thresval))
        out = x / std
        return out.squeeze(0) if squeeze else out


class SpatialContrastiveNormalization(_SpatialNorm):
    t7_name = "nn.SpatialContrastiveNormalization"

    def __init__(self, nInputPlane: int = 1, kernel: Optional[torch.Tensor] = None, threshold: Optional[float] = None,
                 thresval: Optional[float] = None):
        super().__init__(nInputPlane, kernel)
        self.sub = SpatialSubtractiveNormalization(nInputPlane, self.kernel)
        self.div = SpatialDivisiveNormalization(nInputPlane, self.kernel, threshold, thresval)

    @property
    def threshold(self) -> float:
        return self.div.threshold

    @property
    def thresval(self) -> float:
        return self.div.thresval

    def forward(self, input):
        return self.div(self.sub(input))


__all__ = ["SpatialSubtractiveNormalization", "SpatialDivisiveNormalization", "SpatialContrastiveNormalization",
           "local_estimate"]
