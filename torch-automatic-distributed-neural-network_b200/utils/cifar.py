"""CIFAR-10 in the Torch7 batch format the reference example reads
(cifar_example/sgd-torchad_nn-cifar.lua:124-169): ``cifar-10-batches-t7/data_batch_{1..5}.t7`` and
``test_batch.t7``, ASCII ``.t7`` tables with ``data`` = ByteTensor ``[3072, 10000]`` (one image per
column) and ``labels`` = ByteTensor ``[1, 10000]`` (or ``[10000]``), classes 0..9.

The example loads all five training batches into a 50000-row tensor, transposes each to one image per
row, keeps the first ``trsize`` / ``tesize`` samples and reshapes to ``[n, 3, 32, 32]``.  Here only the
batches that are needed are read, and the labels stay 0-based (the reference adds 1 for Lua).  There is
no download step: this box has no network, and a library should not shell out to ``wget`` (lines 131-136).
"""
from __future__ import annotations

import os
from typing import Tuple

import torch

from . import t7

BATCH = 10000
DIRNAME = "cifar-10-batches-t7"


def _read_batch(path: str, mode: str):
    tab = t7.load(path, mode)
    data, labels = tab["data"], tab["labels"]
    if data.dim() != 2 or data.shape[0] != 3072:
        raise ValueError(f"{path}: expected data of shape [3072, n], got {tuple(data.shape)}")
    labels = labels.reshape(-1)
    if labels.numel() != data.shape[1]:
        raise ValueError(f"{path}: {labels.numel()} labels for {data.shape[1]} images")
    return data.t(), labels  # one image per row (subset.data:t(), line 147)


def load_cifar10_t7(root: str, trsize: int = 4000, tesize: int = 2000, mode: str = "ascii",
                    dtype: torch.dtype = torch.float32) -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor, torch.Tensor]:
    """Returns ``(train_x [trsize,3,32,32], train_y [trsize], test_x [tesize,3,32,32], test_y [tesize])``;
    pixel values 0..255 as ``dtype``, labels int64 in 0..9.  ``root`` is the directory that holds
    ``cifar-10-batches-t7`` (or that directory itself)."""
    d = os.path.join(root, DIRNAME) if os.path.isdir(os.path.join(root, DIRNAME)) else root
    if not os.path.isfile(os.path.join(d, "test_batch.t7")):
        raise FileNotFoundError(f"no CIFAR-10 Torch7 batches under {root!r} (expected {DIRNAME}/data_batch_1.t7 … test_batch.t7)")
    xs, ys, have = [], [], 0
    for i in range(1, 6):
        if have >= trsize:
            break
        path = os.path.join(d, f"data_batch_{i}.t7")
        if not os.path.isfile(path):
            break
        x, y = _read_batch(path, mode)
        xs.append(x)
        ys.append(y)
        have += x.shape[0]
    if have < trsize:
        raise ValueError(f"only {have} training samples on disk, {trsize} requested")
    train_x = torch.cat(xs)[:trsize].to(dtype).reshape(trsize, 3, 32, 32)
    train_y = torch.cat(ys)[:trsize].to(torch.int64)
    x, y = _read_batch(os.path.join(d, "test_batch.t7"), mode)
    if x.shape[0] < tesize:
        raise ValueError(f"only {x.shape[0]} test samples on disk, {tesize} requested")
    test_x = x[:tesize].to(dtype).reshape(tesize, 3, 32, 32)
    test_y = y[:tesize].to(torch.int64)
    return train_x, train_y, test_x, test_y


def write_cifar10_t7(root: str, train_x: torch.Tensor, train_y: torch.Tensor, test_x: torch.Tensor, test_y: torch.Tensor,
                     mode: str = "ascii", batch: int = BATCH) -> str:
    """Write ``uint8`` images ``[n,3,32,32]`` and labels in the same on-disk layout (up to five training
    batches of ``batch`` images + the test batch).  Used by the tests and to export synthetic data."""
    d = os.path.join(root, DIRNAME)
    os.makedirs(d, exist_ok=True)

    def dump(path, x, y):
        t7.save(path, {"data": x.reshape(x.shape[0], 3072).to(torch.uint8).t().contiguous(),
                       "labels": y.reshape(1, -1).to(torch.uint8)}, mode)

    n = train_x.shape[0]
    if n > 5 * batch:
        raise ValueError("at most five training batches")
    for i in range(5):
        lo, hi = i * batch, min((i + 1) * batch, n)
        if lo >= hi:
            break
        dump(os.path.join(d, f"data_batch_{i + 1}.t7"), train_x[lo:hi], train_y[lo:hi])
    dump(os.path.join(d, "test_batch.t7"), test_x, test_y)
    return d


__all__ = ["load_cifar10_t7", "write_cifar10_t7"]
