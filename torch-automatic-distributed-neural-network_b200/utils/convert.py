"""Dense <-> model-parallel conversion (SURVEY §2.6 layout resolution, Q4).

On disk and in ``tadnn.nn.MP*Linear`` the layout is the reference's: rank r holds
``W_r = W[:, r*i/P:(r+1)*i/P]`` and a full-length bias whose *sum over ranks* is the dense bias.
"""
from __future__ import annotations

from typing import List, Tuple

import torch


def shard_dense_linear(weight: torch.Tensor, bias: torch.Tensor, world: int) -> List[Tuple[torch.Tensor, torch.Tensor]]:
    """Dense ``(W [o,i], b [o])`` -> per-rank ``[(W_r [o,i/P], b_r [o])]`` with ``b_0 = b``, others 0."""
    o, i = weight.shape
    assert i % world == 0
    w = i // world
    out = []
    for r in range(world):
        br = bias.clone() if r == 0 else torch.zeros_like(bias)
        out.append((weight[:, r * w:(r + 1) * w].clone(), br))
    return out


def merge_mp_linear(shards: List[Tuple[torch.Tensor, torch.Tensor]]) -> Tuple[torch.Tensor, torch.Tensor]:
    """Per-rank shards -> dense ``(cat_r W_r, sum_r b_r)``."""
    W = torch.cat([s[0] for s in shards], dim=1)
    b = torch.stack([s[1] for s in shards]).sum(0)
    return W, b


def dense_model_from_mp(model, mpi=None):
    """All-gather an MP ``Sequential`` into an equivalent dense ``tadnn.nn`` model (on every rank)."""
    from .. import nn as tnn
    from ..nn import mp
    out = tnn.Sequential()
    for m in model.children():
        if isinstance(m, mp._MPLinear):
            W, b = m.dense_weight_bias(mpi)
            lin = tnn.Linear(W.shape[1], W.shape[0])
            with torch.no_grad():
                lin.weight.copy_(W)
                lin.bias.copy_(b)
            out.add(lin)
        elif isinstance(m, mp.MPTanh):
            out.add(tnn.Tanh())
        elif isinstance(m, mp._MPReshape):
            out.add(tnn.Reshape(*m.size, batchMode=m.batchMode))
        else:
            out.add(m)
    return out


def mp_model_from_dense(model, mpi=None):
    """Dense ``Sequential`` of Reshape/Linear/Tanh -> MP layers holding this rank's shards."""
    from .. import nn as tnn
    out = tnn.Sequential()
    first = True
    for m in model.children():
        if isinstance(m, torch.nn.Linear):
            cls = tnn.MPInitialLinear if first else tnn.MPBaseLinear
            mpl = cls(m.in_features, m.out_features, mpi=mpi)
            mpl.load_dense(m.weight.detach(), m.bias.detach() if m.bias is not None else None, mpi)
            out.add(mpl)
            first = False
        elif isinstance(m, torch.nn.Tanh):
            out.add(tnn.MPTanh(mpi=mpi))
        elif isinstance(m, tnn.Reshape):
            cls = tnn.MPInitialReshape if first else tnn.MPBaseReshape
            out.add(cls(*m.size, batchMode=m.batchMode, mpi=mpi))
        else:
            out.add(m)
    return out
