"""Automatic data parallelism — ``tadnn.datamodule`` (reference: datamodule.lua).

``parallelize(data, targets, model, size, mpi, mpinn, batchSize)`` is the one call that turns
a single-device training script into a data-parallel one (datamodule.lua:15-53): it starts the
communication layer if needed, broadcasts the initial weights, gives every rank an even
contiguous stripe of the dataset, installs hooks that synchronise the model every
``batchSize`` backward calls / training samples, and returns ``(data, targets, size)`` for the
local shard.  ``batchSize=-1`` skips the hooks; the user then calls
``synchronizeModel(model, mpi)`` (README.md:66-75).

B200 design: on CUDA the model's parameters and gradients are re-homed into a flat region of
the symmetric heap (``tadnn.parallel.fusion``), so that ``synchronizeModel`` — 2·N blocking
collectives + 2·N ``div`` kernels in the reference (datamodule.lua:211-224) — is one fused
all-reduce-average kernel over NVSwitch peer memory, overlapped with backward
(``tadnn.parallel.dp``).

Module-level switches (file-local constants in the reference, datamodule.lua:3-4) are
environment-driven: ``TADNN_DATA_SHUFFLE`` (``dataShuffle``), ``TADNN_SYNC_PROTOTYPE``
(``syncPrototype``).
"""
from __future__ import annotations

import math
import os
import time
from typing import Optional

import torch

from . import mpi as _mpi_mod
from . import mpinn as _mpinn_mod


def _flag(name: str, default: bool) -> bool:
    v = os.environ.get(name)
    if v is None:
        return default
    return v.lower() in ("1", "true", "yes", "on")


def dataShuffle() -> bool:  # datamodule.lua:3
    return _flag("TADNN_DATA_SHUFFLE", False)


def syncPrototype() -> bool:  # datamodule.lua:4
    return _flag("TADNN_SYNC_PROTOTYPE", True)


_USAGE = "-usage for parallelize(data, targets, model, size)"


# ----------------------------------------------------------------------------
# parallelize (datamodule.lua:15-53)
# ----------------------------------------------------------------------------
def parallelize(data=None, targets=None, model=None, size=None, mpi_obj=None, mpinn_obj=None,
                batchSize=None, *, sync_init: Optional[bool] = None, remainder: str = "drop",
                shuffle: Optional[bool] = None, flatten: Optional[bool] = None,
                sync_mode: Optional[str] = None, overlap: Optional[bool] = None, verbose: bool = True,
                criterion=None):
    """Returns ``(newdata, newtargets, dataSize)`` — three values, as the code does
    (datamodule.lua:52), not the four its header comment promises — or ``-1`` after printing
    the usage line when a required argument is missing (datamodule.lua:18-21).

    Keyword extensions (defaults reproduce the reference): ``sync_init`` — broadcast initial
    weights; ``None`` = always (the reference only does it when it had to create ``mpinn``
    itself, datamodule.lua:31-34; ``TADNN_COMPAT_SYNC_INIT=1`` restores that);
    ``remainder`` "drop" | "last" (Q6); ``shuffle`` strided shard (Q7); ``flatten`` move
    params/grads into the flat symmetric buffer (default: on CUDA); ``sync_mode`` "model" |
    "grads"; ``overlap`` overlap the sync with backward; ``batchSize="auto"`` measures the per-sample
    compute time and the cost of one synchronisation (:func:`comm_speed` with ``criterion``) and picks
    the period that keeps the synchronisation overhead near 10 % (:func:`sync_period_from_speed`).
    """
    if data is None or targets is None or model is None or size is None:
        print(_USAGE)
        return -1
    created_mpinn = False
    if mpi_obj is None:  # datamodule.lua:24-28
        mpi_obj = _mpi_mod.current()
        if not mpi_obj.started:
            mpi_obj.start(True)
    if mpinn_obj is None:  # datamodule.lua:31-34
        mpinn_obj = _mpinn_mod
        created_mpinn = True

    use_cuda = any(p.is_cuda for p in model.parameters())
    if flatten is None:
        flatten = use_cuda
    if flatten and getattr(model, "_tadnn_flat", None) is None:
        from .parallel.fusion import FlatBuffer
        ctx = getattr(mpi_obj, "symm", None) if use_cuda else None
        FlatBuffer(model, ctx, mpi_obj)

    if sync_init is None:
        sync_init = created_mpinn if _flag("TADNN_COMPAT_SYNC_INIT", False) else True
    if sync_init:
        mpinn_obj.synchronizeParameters(model, mpi_obj)

    newdata, dataSize = data_parallel(data, size, mpi_obj, remainder=remainder, shuffle=shuffle, verbose=verbose)
    newtargets, _ = data_parallel(targets, size, mpi_obj, remainder=remainder, shuffle=shuffle, verbose=False)

    if batchSize == "auto":
        # what the reference's commented-out call (datamodule.lua:40-42) was reaching for: measure the
        # compute and the synchronisation cost on this machine and derive the period from them
        speed = comm_speed(newdata, newtargets, model, mpi_obj, mpinn_obj, criterion=criterion,
                           iters=min(10, int(dataSize)), split=True, verbose=verbose)
        optimize_sync(speed, dataSize, model, mpinn_obj, mpi_obj, None, sync_mode=sync_mode, overlap=overlap)
    elif batchSize != -1:  # datamodule.lua:45-47
        optimize_sync(None, dataSize, model, mpinn_obj, mpi_obj, batchSize, sync_mode=sync_mode, overlap=overlap)

    mpi_obj.barrier()  # datamodule.lua:50
    return newdata, newtargets, dataSize


# ----------------------------------------------------------------------------
# optimize_sync (datamodule.lua:65-188)
# ----------------------------------------------------------------------------
def heuristic_batch_size(size: int) -> int:
    """datamodule.lua:68-78."""
    if size < 1000:
        return 1
    if size < 2500:
        return 10
    if size < 5000:
        return 50
    return 100


def sync_period_from_speed(compute_s: float, sync_s: float, size: int, overhead: float = 0.10) -> int:
    """Samples between synchronisations such that ``sync_s / (period * compute_s) <= overhead``, clamped
    to ``[1, size]``.  The measured counterpart of the reference's size table (datamodule.lua:68-78)."""
    if compute_s <= 0.0:
        return heuristic_batch_size(size)
    period = int(math.ceil(sync_s / (overhead * compute_s)))
    return max(1, min(int(size), period))


def optimize_sync(speed, size, model, mpinn, mpi, batchSize=None, *, sync_mode: Optional[str] = None,
                  overlap: Optional[bool] = None):
    """Pick the sync period (if not given) and install the synchronisation hooks on ``model``.
    Returns ``batchSize`` (datamodule.lua:187).  ``speed`` is ignored as in the reference (its call
    site passes an undefined global, datamodule.lua:46) unless it is the ``(compute, sync)`` pair of
    ``comm_speed(split=True)`` and no ``batchSize`` was given: then the period is derived from it."""
    from .parallel.dp import SyncState
    if batchSize is None:
        if isinstance(speed, (tuple, list)) and len(speed) == 2:  # (compute s/sample, s per synchronisation)
            batchSize = sync_period_from_speed(speed[0], speed[1], size)
        else:
            batchSize = heuristic_batch_size(size)
    old = getattr(model, "_tadnn_sync", None)
    if old is not None:
        old.remove()
    if sync_mode is None:
        sync_mode = "model" if syncPrototype() else "grads"
    st = SyncState(model, mpi, mpinn, batchSize, size, sync_mode=sync_mode, overlap=overlap)
    model._tadnn_sync = st
    st.install_autograd_hooks()
    return batchSize


def remove_hooks(model) -> None:
    st = getattr(model, "_tadnn_sync", None)
    if st is not None:
        st.remove()


# ----------------------------------------------------------------------------
# synchronizeModel (+ selectCollective)  (datamodule.lua:199-224)
# ----------------------------------------------------------------------------
def selectCollective(tensor, sync, collective, mpi):
    """cpu/gpu × singlenode/multinode × sync/async lookup (datamodule.lua:199-208)."""
    ttype = "gpu" if tensor.is_cuda else "cpu"
    scope = "multinode" if mpi.needInterNodeCollectives else "singlenode"
    try:
        return mpi.collectiveSelector[ttype][scope][sync][collective]
    except KeyError:
        raise AssertionError(f"Could not find collective {ttype} {scope} {sync} {collective}")


def synchronizeModel(net, mpi=None):
    """All-reduce-average every parameter tensor, then every gradient tensor
    (datamodule.lua:211-224).  Fast path: one fused kernel on the flat symmetric buffer."""
    if not hasattr(net, "parameters"):
        return
    mpi = mpi or _mpi_mod.current()
    P = mpi.size()
    flat = getattr(net, "_tadnn_flat", None)
    if flat is not None:
        flat.rehome_grads()
        flat.allreduce("model", 1.0 / P)
        return
    with torch.no_grad():
        params = list(net.parameters())
        for w in params:
            allreduceTensor = selectCollective(w.data, "sync", "allreduceTensor", mpi)
            allreduceTensor(w.data)
            w.data.div_(P)
        for w in params:
            if w.grad is None:
                w.grad = torch.zeros_like(w)
            allreduceTensor = selectCollective(w.grad, "sync", "allreduceTensor", mpi)
            allreduceTensor(w.grad)
            w.grad.div_(P)


# ----------------------------------------------------------------------------
# data_parallel (datamodule.lua:235-267)
# ----------------------------------------------------------------------------
def shard_bounds(size: int, rank: int, world: int, remainder: str = "drop"):
    """0-based half-open ``[start, finish)`` of rank's stripe and the stripe length.
    ``stripe = floor(size/P)``; the ``size mod P`` trailing samples are dropped, exactly as the
    code does (datamodule.lua:239-246) despite its comment; ``remainder="last"`` implements
    what the comment promises."""
    stripe = (size - (size % world)) // world
    start = rank * stripe
    finish = start + stripe
    rem = 0
    if remainder == "last" and rank == world - 1:
        rem = size % world
        finish += rem
    return start, finish, stripe, rem


def data_parallel(data, size, mpi, *, remainder: str = "drop", shuffle: Optional[bool] = None,
                  verbose: bool = True):
    """Returns ``(shard, shardSize)``.  The shard is a *view* of ``data`` (no copy), like
    ``data[{{start,finish}}]`` (datamodule.lua:251).  ``shuffle`` (``dataShuffle``) gives rank r
    the strided samples ``r, r+P, r+2P, …`` — the intent of datamodule.lua:253-259 without its
    aliasing bug."""
    rank, world = mpi.rank(), mpi.size()
    start, finish, stripe, rem = shard_bounds(int(size), rank, world, remainder)
    if shuffle is None:
        shuffle = dataShuffle()
    if shuffle:
        newdata = data[rank:stripe * world:world]
    else:
        newdata = data[start:finish]
    if verbose:
        # 1-based inclusive bounds, the reference's log line (datamodule.lua:261)
        print(f"Rank: {rank} Start Point: {start + 1} End Point: {finish} Stripe: {stripe} Remainder: {rem}")
    return newdata, stripe + rem


# ----------------------------------------------------------------------------
# comm_speed (datamodule.lua:280-303; dead code there, live here)
# ----------------------------------------------------------------------------
def comm_speed(data, targets, model, mpi=None, mpinn=None, criterion=None, iters: int = 10, split: bool = False,
               verbose: bool = True):
    """Mean seconds of {forward, criterion backward, model backward, synchronizeGradients} over
    ``iters`` samples.  Device-timed with CUDA events on GPU, wall clock on CPU.  ``split=True`` times the
    compute and the synchronisation separately and returns ``(compute_s, sync_s)``, each the MAX over
    ranks (so every rank derives the same period); gradients are left as they were."""
    if split:
        return _comm_speed_split(data, targets, model, mpi or _mpi_mod.current(), mpinn or _mpinn_mod, criterion, iters, verbose)
    mpi = mpi or _mpi_mod.current()
    mpinn = mpinn or _mpinn_mod
    if criterion is None:
        from .nn import CrossEntropyCriterion
        criterion = CrossEntropyCriterion()
    mpi.barrier()
    cuda = any(p.is_cuda for p in model.parameters())
    if cuda:
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
    t0 = time.perf_counter()
    for i in range(iters):
        x, y = data[i], targets[i]
        with torch.enable_grad():
            out = model(x)
            loss = criterion(out, y)
            grads = torch.autograd.grad(loss, [p for p in model.parameters() if p.requires_grad], allow_unused=True)
        with torch.no_grad():
            for p, g in zip([p for p in model.parameters() if p.requires_grad], grads):
                if g is not None:
                    if p.grad is None:
                        p.grad = torch.zeros_like(p)
                    p.grad.add_(g)
        mpinn.synchronizeGradients(model, mpi)
    if cuda:
        e1.record()
        torch.cuda.synchronize()
        speed = e0.elapsed_time(e1) / 1e3 / iters
    else:
        speed = (time.perf_counter() - t0) / iters
    print(f"Rank: {mpi.rank()} comm speed: {speed}")
    return speed


def _comm_speed_split(data, targets, model, mpi, mpinn, criterion, iters, verbose):
    if criterion is None:
        from .nn import CrossEntropyCriterion
        criterion = CrossEntropyCriterion()
    params = [p for p in model.parameters() if p.requires_grad]
    cuda = any(p.is_cuda for p in params)

    def timed(fn, n):
        if cuda:
            torch.cuda.synchronize()
        t0 = time.perf_counter()
        for i in range(n):
            fn(i)
        if cuda:
            torch.cuda.synchronize()
        return (time.perf_counter() - t0) / max(1, n)

    def compute(i):
        x, y = data[i % len(data)], targets[i % len(targets)]
        if cuda:
            x = x.to(params[0].device)
            y = y.to(params[0].device) if torch.is_tensor(y) else y
        with torch.enable_grad():
            torch.autograd.grad(criterion(model(x), y), params, allow_unused=True)

    mpi.barrier()
    compute(0)  # warm-up (allocator, kernels)
    compute_s = timed(compute, iters)
    saved = [None if p.grad is None else p.grad.detach().clone() for p in params]
    for p in params:  # synchronizeModel reduces parameters AND gradients: give it gradients to move
        if p.grad is None:
            p.grad = torch.zeros_like(p)
    before = [p.detach().clone() for p in params]
    synchronizeModel(model, mpi)  # warm-up
    sync_s = timed(lambda i: synchronizeModel(model, mpi), 3)
    with torch.no_grad():  # averaging identical replicas is a no-op, but restore anyway (replicas may differ)
        for p, b, g in zip(params, before, saved):
            p.copy_(b)
            if g is None:
                p.grad = None
            else:
                p.grad.copy_(g)
    t = torch.tensor([compute_s, sync_s], dtype=torch.float64)
    t = t.to(params[0].device) if cuda else t
    for k in range(2):  # MAX over ranks from a sum-only all-reduce: gather one slot per rank
        slots = torch.zeros(mpi.size(), dtype=t.dtype, device=t.device)
        slots[mpi.rank()] = t[k]
        mpi.allreduceTensor(slots)
        t[k] = slots.max()
    compute_s, sync_s = float(t[0]), float(t[1])
    if verbose:
        print(f"Rank: {mpi.rank()} compute {compute_s:.6f} s/sample, sync {sync_s:.6f} s")
    return compute_s, sync_s


# snake_case aliases
synchronize_model = synchronizeModel
select_collective = selectCollective
