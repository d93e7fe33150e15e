"""``torchmpi.nn`` equivalent: parameter broadcast and gradient all-reduce.

Reference call sites: ``mpinn.synchronizeParameters(model)`` (datamodule.lua:32-33) makes
all replicas identical before training; ``mpinn.synchronizeGradients(model)``
(datamodule.lua:106,155,293 — the disabled ``syncPrototype=false`` path) all-reduces every
gradient tensor.  Both take the handle from ``tadnn.mpi.current()`` unless one is given.
"""
from __future__ import annotations

import torch

from . import mpi as _mpi


def _params(model):
    return [p for p in model.parameters()]


def synchronizeParameters(model, mpi=None, root: int = 0) -> None:
    """Broadcast every parameter (and buffer) from ``root`` (K4 on a flat symmetric buffer
    when the model has one, per-tensor broadcast otherwise)."""
    mpi = mpi or _mpi.current()
    if mpi.size() == 1:
        return
    flat = getattr(model, "_tadnn_flat", None)
    if flat is not None and flat.symmetric:
        flat.broadcast_params(root)
        with torch.no_grad():  # buffers (e.g. running statistics) are not part of the flat region
            for b in model.buffers():
                if b.is_floating_point():
                    mpi.broadcastTensor(root, b.data)
        return
    with torch.no_grad():
        for p in _params(model):
            mpi.broadcastTensor(root, p.data)
        for b in model.buffers():
            if b.is_floating_point():
                mpi.broadcastTensor(root, b.data)


def synchronizeGradients(model, mpi=None, average: bool = False) -> None:
    """All-reduce (sum; ``average=True`` divides by world size) every gradient tensor."""
    mpi = mpi or _mpi.current()
    P = mpi.size()
    if P == 1:
        return
    flat = getattr(model, "_tadnn_flat", None)
    if flat is not None and flat.symmetric:
        flat.allreduce(what="grads", scale=(1.0 / P) if average else 1.0)
        return
    with torch.no_grad():
        for p in _params(model):
            if p.grad is None:
                p.grad = torch.zeros_like(p)
            mpi.allreduceTensor(p.grad)
            if average:
                p.grad.div_(P)


# snake_case aliases
synchronize_parameters = synchronizeParameters
synchronize_gradients = synchronizeGradients
