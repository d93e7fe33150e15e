"""Fused mean-squared-error criterion for wide bf16 outputs (csrc/loss.cu).

``mse_fused(y, target)`` is ``nn.MSECriterion`` (sizeAverage) as one kernel: it reads the bf16 model
output and the target once and produces the scalar loss *and* the bf16 gradient in the same pass; the
autograd backward only scales that gradient by the upstream value.  Meant for outputs so wide that the
elementwise loss arithmetic in fp32 costs as much as a layer (the 4096 x 16384 model-parallel MLP).

Status: built and bound, first GPU run pending (tests/test_gpu_loss_wip.py carries the ``wip`` marker, not
``gpu``, until it has passed on a B200); nothing on a default path calls it.
"""
from __future__ import annotations

import ctypes as C

import torch

from .. import _native


class _MSEFused(torch.autograd.Function):
    @staticmethod
    def forward(ctx, y: torch.Tensor, target: torch.Tensor, size_average: bool):
        assert y.is_cuda and y.dtype == torch.bfloat16 and y.is_contiguous()
        t = target.to(y.device)
        if t.dtype not in (torch.float32, torch.bfloat16):
            t = t.float()
        t = t.reshape(y.shape).contiguous()
        n = y.numel()
        grad = torch.empty_like(y)
        loss = torch.empty(1, device=y.device, dtype=torch.float32)
        k = 1.0 / n if size_average else 1.0
        _native.check(_native.lib().tadnn_mse_loss_grad(
            C.c_void_p(y.data_ptr()), C.c_void_p(t.data_ptr()), int(t.dtype == torch.float32), C.c_void_p(grad.data_ptr()),
            C.c_void_p(loss.data_ptr()), C.c_longlong(n), C.c_float(k), C.c_float(2.0 * k),
            C.c_void_p(_native.stream_ptr())), "tadnn_mse_loss_grad")
        ctx.save_for_backward(grad)
        return loss.reshape(())

    @staticmethod
    def backward(ctx, g):
        (grad,) = ctx.saved_tensors
        return grad * g.to(grad.dtype), None, None


def mse_fused(y: torch.Tensor, target: torch.Tensor, size_average: bool = True) -> torch.Tensor:
    return _MSEFused.apply(y, target, size_average)


__all__ = ["mse_fused"]
