"""Checkpoint / resume (absent from the reference: it never calls ``torch.save`` on a model).

* ``save`` / ``load``: native PyTorch checkpoint of model + trainer counters.  Data-parallel replicas
  are identical after a sync, so rank 0 writes; model-parallel models hold different shards per
  rank, so every rank writes ``<path>.rank<r>`` (``per_rank=True``).
* ``.t7`` export/import in the reference's class layout lives in ``tadnn.utils.t7``; dense <->
  sharded conversion in ``tadnn.utils.convert``.
"""
from __future__ import annotations

import os
from typing import Optional

import torch

from .. import mpi as _mpi


def _path(path: str, rank: int, per_rank: bool) -> str:
    return f"{path}.rank{rank}" if per_rank else path


def save(path: str, model: torch.nn.Module, trainer=None, mpi=None, per_rank: bool = False, extra: Optional[dict] = None) -> Optional[str]:
    h = mpi or _mpi.current()
    rank = h.rank()
    if not per_rank and rank != 0:
        h.barrier()
        return None
    state = {"model": {k: v.detach().cpu() for k, v in model.state_dict().items()},
             "world": h.size(), "rank": rank, "extra": extra or {}}
    st = getattr(model, "_tadnn_sync", None)
    state["sync"] = {"model_counter": getattr(model, "sync_counter", None),
                     "batchSize": getattr(st, "batchSize", None), "size": getattr(st, "size", None)}
    if trainer is not None:
        state["trainer"] = {k: getattr(trainer, k) for k in ("learningRate", "learningRateDecay", "maxIteration",
                                                             "shuffleIndices", "sync_counter", "steps")
                            if hasattr(trainer, k)}
    out = _path(path, rank, per_rank)
    os.makedirs(os.path.dirname(os.path.abspath(out)) or ".", exist_ok=True)
    tmp = out + ".tmp"
    torch.save(state, tmp)
    os.replace(tmp, out)  # atomic: a crash never leaves a half-written checkpoint
    if not per_rank:
        h.barrier()
    return out


def load(path: str, model: torch.nn.Module, trainer=None, mpi=None, per_rank: bool = False) -> dict:
    h = mpi or _mpi.current()
    state = torch.load(_path(path, h.rank(), per_rank), map_location="cpu")
    if per_rank and state["world"] != h.size():
        raise ValueError(f"checkpoint was written for world size {state['world']}, running with {h.size()}; "
                         "convert through tadnn.utils.convert (dense <-> sharded)")
    with torch.no_grad():
        own = model.state_dict()
        for k, v in state["model"].items():
            own[k].copy_(v.to(own[k].device, own[k].dtype))  # in place: keeps the flat-buffer views
    if state["sync"].get("model_counter") is not None:
        model.sync_counter = state["sync"]["model_counter"]
    if trainer is not None:
        for k, v in state.get("trainer", {}).items():
            setattr(trainer, k, v)
    return state
