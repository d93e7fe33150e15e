"""Torch7 ``nn.Module`` protocol as a mixin for ``torch.nn.Module`` (see tadnn.nn)."""
from __future__ import annotations

from typing import List, Optional, Tuple

import torch

# ----------------------------------------------------------------------------
# Torch7 module protocol
# ----------------------------------------------------------------------------
class T7Protocol:
    """Mixin giving a ``torch.nn.Module`` the Torch7 ``nn.Module`` methods.

    ``backward`` accumulates into ``param.grad`` (Torch7's ``gradWeight``/``gradBias``), exactly
    like ``Module:backward = updateGradInput + accGradParameters``.
    """

    output: Optional[torch.Tensor] = None
    gradInput: Optional[torch.Tensor] = None

    # -- forward ---------------------------------------------------------------
    needs_grad_input = False  # Torch7 always returns gradInput; here only on request / when x requires grad

    def _run(self, x):
        return torch.nn.Module.__call__(self, x)

    def _t7_forward(self, input: torch.Tensor) -> torch.Tensor:
        x = input
        if torch.is_grad_enabled() and self.needs_grad_input and torch.is_tensor(x) \
                and x.is_floating_point() and not x.requires_grad:
            x = x.detach().requires_grad_(True)
        out = self._run(x)
        self._t7_in = x
        self._t7_src = input
        self._t7_out = out
        self._t7_cached = None
        self.output = out
        return out

    def updateOutput(self, input):
        with torch.enable_grad():
            return self._t7_forward(input)

    # -- backward --------------------------------------------------------------
    def _t7_grads(self, input, gradOutput):
        """(gradInput, [param grads]) of the last forward; cached so that the Torch7 split
        ``updateGradInput`` + ``accGradParameters`` costs one autograd pass."""
        cached = getattr(self, "_t7_cached", None)
        if cached is not None and cached[0] is gradOutput and getattr(self, "_t7_src", None) is input:
            return cached[1], cached[2]
        if getattr(self, "_t7_out", None) is None or getattr(self, "_t7_src", None) is not input \
                or not self._t7_out.requires_grad:
            with torch.enable_grad():
                self._t7_forward(input)
        params = [p for p in self.parameters() if p.requires_grad]
        want_gi = torch.is_tensor(self._t7_in) and self._t7_in.requires_grad
        wrt = ([self._t7_in] if want_gi else []) + params
        grads = torch.autograd.grad([self._t7_out], wrt, [gradOutput.to(self._t7_out.dtype)],
                                    allow_unused=True, retain_graph=False)
        gi = grads[0] if want_gi else None
        pg = list(grads[1:] if want_gi else grads)
        self._t7_cached = (gradOutput, gi, pg)
        self._t7_out = None
        return gi, pg

    def updateGradInput(self, input, gradOutput):
        gi, _ = self._t7_grads(input, gradOutput)
        self.gradInput = gi
        return gi

    def accGradParameters(self, input, gradOutput, scale: float = 1.0):
        _, pg = self._t7_grads(input, gradOutput)
        with torch.no_grad():
            for p, g in zip((p for p in self.parameters() if p.requires_grad), pg):
                if g is None:
                    continue
                if p.grad is None:
                    p.grad = torch.zeros_like(p)
                p.grad.add_(g, alpha=scale)

    def accUpdateGradParameters(self, input, gradOutput, lr: float):
        """In-place SGD step straight from the gradient, gradWeight untouched
        (datamodule.lua:142)."""
        _, pg = self._t7_grads(input, gradOutput)
        with torch.no_grad():
            for p, g in zip((p for p in self.parameters() if p.requires_grad), pg):
                if g is not None:
                    p.add_(g, alpha=-lr)

    def backward(self, input, gradOutput, scale: float = 1.0):
        gi = self.updateGradInput(input, gradOutput)
        self.accGradParameters(input, gradOutput, scale)
        st = getattr(self, "_tadnn_sync", None)
        if st is not None:
            st.after_backward(self)  # the hooked Sequential:backward, datamodule.lua:96-110
        return gi

    # -- parameters ------------------------------------------------------------
    def t7_parameters(self) -> Tuple[List[torch.Tensor], List[torch.Tensor]]:
        """Torch7 ``net:parameters()``: (weights, gradWeights) as two parallel lists."""
        ps = [p for p in self.parameters()]
        for p in ps:
            if p.grad is None:
                p.grad = torch.zeros_like(p)
        return [p.data for p in ps], [p.grad for p in ps]

    def getParameters(self) -> Tuple[torch.Tensor, torch.Tensor]:
        """Flatten all parameters / gradients into two contiguous vectors and re-point the
        module's tensors at views of them (Torch7 ``getParameters`` semantics)."""
        from ..parallel.fusion import FlatBuffer
        flat = getattr(self, "_tadnn_flat", None)
        if flat is None:
            flat = FlatBuffer(self, None)
        return flat.flat_params(), flat.flat_grads()

    def zeroGradParameters(self):
        for p in self.parameters():
            if p.grad is not None:
                p.grad.zero_()

    def updateParameters(self, lr: float):
        with torch.no_grad():
            for p in self.parameters():
                if p.grad is not None:
                    p.add_(p.grad, alpha=-lr)

    def training(self):
        self.train(True)
        return self

    def evaluate(self):
        self.train(False)
        return self


