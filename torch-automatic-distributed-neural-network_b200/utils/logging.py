"""Rank-aware logging and JSONL metrics (the reference only ``print``s: SURVEY §5)."""
from __future__ import annotations

import json
import os
import sys
import time
from typing import Optional


def _rank() -> int:
    try:
        import torch.distributed as dist
        if dist.is_available() and dist.is_initialized():
            return dist.get_rank()
    except Exception:  # noqa: BLE001
        pass
    return int(os.environ.get("RANK", "0"))


def log(msg: str, *, all_ranks: bool = False, file=None) -> None:
    r = _rank()
    if all_ranks or r == 0:
        print(f"[tadnn r{r}] {msg}", file=file or sys.stdout, flush=True)


class MetricsWriter:
    """Append-only JSONL: samples/s, sync ms, bus GB/s, roofline fractions…  Rank 0 writes."""

    def __init__(self, path: Optional[str]):
        self.path = path
        self.enabled = path is not None and _rank() == 0
        if self.enabled:
            os.makedirs(os.path.dirname(os.path.abspath(path)) or ".", exist_ok=True)

    def write(self, **fields) -> None:
        if not self.enabled:
            return
        fields.setdefault("ts", time.time())
        with open(self.path, "a") as f:
            f.write(json.dumps(fields) + "\n")


def write_results_file(directory: str, *, train_size: int, test_size: int, batch_size: int, accuracy: float,
                       learning_rate: float, threads: int, preprocess_s: float, train_s: float) -> str:
    """The example's results file, same name pattern and fields
    (cifar_example/sgd-torchad_nn-cifar.lua:277-299).  Rank 0 writes."""
    path = os.path.join(directory, f"SgdAuto--Size:{train_size}--Batch:{batch_size}.txt")
    if _rank() != 0:
        return path
    with open(path, "w") as f:
        f.write("----------------------------------------\n")
        f.write("----------------DataSet-----------------\n")
        f.write(f"Train set Size:     {train_size}\n")
        f.write(f"Test set Size:      {test_size}\n")
        f.write(f"batchSize:          {batch_size}\n")
        f.write("----------------Network-----------------\n")
        f.write(f"Accuracy:           {accuracy}\n")
        f.write(f"Learning rate:      {learning_rate}\n")
        f.write("----------------System------------------\n")
        f.write(f"Threads:            {threads}\n")
        f.write(f"Pre-process Time:   {preprocess_s}\n")
        f.write(f"Training Time:      {train_s}\n")
        f.write("----------------------------------------\n")
    return path
