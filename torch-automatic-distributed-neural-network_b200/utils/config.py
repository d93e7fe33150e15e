"""One config object for the example/bench entry points.

The reference has three disjoint mechanisms: 8 ``torch.CmdLine`` flags in the
example (cifar_example/sgd-torchad_nn-cifar.lua:29-43), file-local booleans that need source
edits (datamodule.lua:3-4; nodemodule.lua:3-5) and interactive prompts in ``train.sh``.  Here:
one dataclass + argparse mirroring the 8 flags, and ``TADNN_*`` environment switches for the
booleans and for kernel-path forcing.
"""
from __future__ import annotations

import argparse
import dataclasses
import os
from typing import Optional

ENV_SWITCHES = {
    "TADNN_DATA_SHUFFLE": "dataShuffle (datamodule.lua:3): strided instead of contiguous shards",
    "TADNN_SYNC_PROTOTYPE": "syncPrototype (datamodule.lua:4): 1 = params+grads, 0 = grads only",
    "TADNN_PRINT_DIMS": "printDims (nodemodule.lua:3): print shapes in the MP layers",
    "TADNN_COMPAT_BACKWARD": "0 = exact MP backward (syncTanh/syncReshape gather, nodemodule.lua:4-5); 1 = reference arithmetic",
    "TADNN_COMPAT_SYNC_INIT": "1 = broadcast initial weights only when parallelize created mpinn (datamodule.lua:31-34)",
    "TADNN_FORCE_PATH": "nccl = never build the symmetric-memory context (baseline path)",
    "TADNN_FORCE_ALGO": "1 one-shot / 2 two-shot / 3 NVLS for every symmetric all-reduce",
    "TADNN_DP_OVERLAP": "0 = do not overlap the gradient sync with backward",
    "TADNN_MP_REDUCE": "push (default) / nvls: how FusedMPMLP sums the K-split partial products (parallel/tp.py)",
    "TADNN_MP_NVLS_CHUNKS": "row chunks of the NVLS reduction pipeline (default 2)",
    "TADNN_NVLS": "0 = never create the multicast object (VMM heap without NVLS)",
    "TADNN_NVTX": "1 = NVTX ranges around the training-step phases",
    "TADNN_SYMM_HEAP_MB": "size of the symmetric heap per GPU (default 1024)",
    "TADNN_SYMM_STAGING_MB": "staging area inside the heap for tensors that live outside it (default 128)",
    "TADNN_COMM_LL": "0 = never use the low-latency pair all-reduce (fp32 messages up to 512 KB)",
    "TADNN_DP_FUSED_SGD": "0 = separate all-reduce and optimiser launches instead of the fused all-reduce + SGD kernels",
    "TADNN_DP_TAIL": "1 / 0 = force / forbid the early synchronisation of the finished gradients on a side stream during conv1's backward (default: on above two ranks)",
    "TADNN_PDL": "0 = plain stream order instead of programmatic dependent launch between the kernels of a step",
    "TADNN_PDL_GRAPH": "0 = no programmatic-dependency edges inside captured CUDA graphs (per-sample step)",
    "TADNN_MP_AUTOFUSE": "0 = a Sequential of MP* layers keeps its per-layer kernels instead of fusing itself",
    "TADNN_NATIVE_LAYERS": "0 = stock tadnn.nn layers fall back to PyTorch ops on a GPU",
}


@dataclasses.dataclass
class TrainConfig:
    """The example's 8 flags (same names, same defaults) + the B200 additions."""
    data: bool = False            # -data: full dataset (32000 train / 10000 test) instead of 4000 / 2000
    seed: int = 1                 # -seed
    learningRate: float = 1e-3    # -learningRate
    batchSize: int = 1            # -batchSize: sync period handed to parallelize()
    weightDecay: float = 0.0      # -weightDecay (parsed and unused in the reference too)
    iterations: int = 1           # -iterations
    threads: int = 1              # -threads
    usegpu: bool = False          # -usegpu
    # additions
    miniBatch: int = 1            # samples per optimisation step (reference: 1)
    dtype: str = "bf16"
    model: str = "cifar_cnn"
    results_dir: str = "."
    dataDir: Optional[str] = None  # directory with cifar-10-batches-t7 (the reference's dataset files); None = synthetic
    preprocess: str = "yuv"       # "yuv": rgb2yuv + contrast normalisation as in the example (:182-216); "none": skip

    @property
    def trsize(self) -> int:
        return 32000 if self.data else 4000

    @property
    def tesize(self) -> int:
        return 10000 if self.data else 2000


def add_reference_flags(ap: argparse.ArgumentParser) -> argparse.ArgumentParser:
    d = TrainConfig()
    ap.add_argument("-data", "--data", action="store_true", help="use full dataset")
    ap.add_argument("-seed", "--seed", type=int, default=d.seed)
    ap.add_argument("-learningRate", "--learningRate", type=float, default=d.learningRate)
    ap.add_argument("-batchSize", "--batchSize", type=int, default=d.batchSize)
    ap.add_argument("-weightDecay", "--weightDecay", type=float, default=d.weightDecay)
    ap.add_argument("-iterations", "--iterations", type=int, default=d.iterations)
    ap.add_argument("-threads", "--threads", type=int, default=d.threads)
    ap.add_argument("-usegpu", "--usegpu", action="store_true")
    ap.add_argument("--miniBatch", type=int, default=d.miniBatch)
    ap.add_argument("--dtype", default=d.dtype, choices=["bf16", "fp32"])
    ap.add_argument("--results-dir", default=d.results_dir)
    ap.add_argument("--dataDir", default=d.dataDir,
                    help="directory holding cifar-10-batches-t7 (Torch7 ASCII batches, as the reference reads them); "
                         "default: synthetic CIFAR-shaped data")
    ap.add_argument("--preprocess", default=d.preprocess, choices=["yuv", "none"],
                    help="dataset preprocessing of the example (rgb2yuv, U/V standardisation, Y contrast normalisation)")
    return ap


def from_args(ns: argparse.Namespace) -> TrainConfig:
    names = {f.name for f in dataclasses.fields(TrainConfig)}
    kw = {k: v for k, v in vars(ns).items() if k in names}
    if "results_dir" in vars(ns):
        kw["results_dir"] = ns.results_dir
    return TrainConfig(**kw)


def env_summary() -> dict:
    return {k: os.environ.get(k) for k in ENV_SWITCHES if os.environ.get(k) is not None}
