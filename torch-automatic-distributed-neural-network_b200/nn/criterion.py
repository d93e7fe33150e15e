"""Torch7 criterions (see tadnn.nn)."""
from __future__ import annotations

from typing import Optional

import torch
import torch.nn.functional as F


# ----------------------------------------------------------------------------
# criterions
# ----------------------------------------------------------------------------
class Criterion(torch.nn.Module):
    """Torch7 criterion protocol: ``forward(input, target) -> float`` stores ``output``;
    ``backward``/``updateGradInput(input, target) -> gradInput``.  Calling the object
    (``crit(input, target)``) returns the differentiable loss tensor (PyTorch idiom)."""

    output: float = 0.0
    gradInput: Optional[torch.Tensor] = None
    sizeAverage = True

    def loss(self, input, target):  # differentiable
        raise NotImplementedError

    def __call__(self, input, target):
        return self.loss(input, target)

    def forward(self, input, target):
        with torch.no_grad():
            self.output = float(self.loss(input, target))
        return self.output

    def updateOutput(self, input, target):
        return self.forward(input, target)

    def updateGradInput(self, input, target):
        with torch.enable_grad():
            x = input.detach().requires_grad_(True)
            l = self.loss(x, target)
            (g,) = torch.autograd.grad(l, x)
        self.gradInput = g
        return g

    def backward(self, input, target):
        return self.updateGradInput(input, target)


def _as_target(target, batch: bool):
    if not torch.is_tensor(target):
        target = torch.as_tensor(target)
    target = target.long()
    return target.reshape(-1) if batch else target.reshape(())


class ClassNLLCriterion(Criterion):
    is_cross_entropy = True  # on log-probabilities; lets a fused native step (model.native_sample_step) stand in

    def loss(self, input, target):
        if input.dim() == 1:
            return F.nll_loss(input.unsqueeze(0).float(), _as_target(target, True).to(input.device))
        return F.nll_loss(input.float(), _as_target(target, True).to(input.device),
                          reduction="mean" if self.sizeAverage else "sum")


class CrossEntropyCriterion(Criterion):
    """``LogSoftMax`` + ``ClassNLLCriterion`` (what the example's criterion is,
    cifar_example/sgd-torchad_nn-cifar.lua:237)."""
    is_cross_entropy = True

    def loss(self, input, target):
        if input.dim() == 1:
            return F.cross_entropy(input.unsqueeze(0).float(), _as_target(target, True).to(input.device))
        return F.cross_entropy(input.float(), _as_target(target, True).to(input.device),
                               reduction="mean" if self.sizeAverage else "sum")


class MSECriterion(Criterion):
    def loss(self, input, target):
        return F.mse_loss(input.float(), target.to(input.device).float().reshape(input.shape),
                          reduction="mean" if self.sizeAverage else "sum")


