"""TorchMPI-shaped communication handle over ``torch.distributed``.

The reference never touches MPI or NCCL itself; it only uses this surface of TorchMPI
(SURVEY §2.4): ``start(useGPU)``, ``rank()``, ``size()``, ``barrier()``,
``allreduceTensor(t)`` (in-place sum), ``allgatherTensor(in, out)``,
``collectiveSelector[cpu|gpu][singlenode|multinode][sync|async][name]`` and
``needInterNodeCollectives`` — call sites datamodule.lua:25-27,50,199-224 and
nodemodule.lua:18-20,52,66,103,117,166,266.

Two access styles are kept: the handle can be passed around (``parallelize(..., mpi)``,
datamodule.lua:15) or picked up as the process-wide "global ``mpi``" that nodemodule.lua
assumes (``tadnn.mpi.current()``).  ``import tadnn.mpi as mpi; mpi.start(True)`` mirrors
``mpi = require('torchmpi'); mpi.start(true)`` (README.md:33-34).

Backends
--------
* ``TorchMPI``  — one process per rank, ``torch.distributed`` (gloo on CPU, NCCL on GPU for
  bootstrap/baseline).  CUDA tensors are reduced by the hand-written sm_100a kernels in
  ``tadnn.parallel.symm`` when a symmetric-memory context is up; NCCL is the baseline path
  (``TADNN_FORCE_PATH=nccl``).
* ``handle.split(color, key)`` / ``mesh(tp_size)`` — sub-communicators with the same API (both
  backends), which is what lets ``parallelize`` (over the DP communicator) and the ``MP*`` layers
  (over the TP communicator) be combined; the reference has one communicator and cannot.
* ``FakeWorld`` / ``FakeMPI`` — P ranks as P threads of one process with a list-of-tensors
  communicator; deterministic; used by the oracle and the unit tests.
"""
from __future__ import annotations

import datetime
import os
import socket
import threading
from typing import Callable, List, Optional

import torch

__all__ = [
    "TorchMPI", "FakeMPI", "FakeWorld", "start", "stop", "rank", "size", "barrier",
    "allreduceTensor", "allgatherTensor", "broadcastTensor", "current", "set_current",
    "started", "mesh",
]

_tls = threading.local()
_default: Optional["TorchMPI"] = None


class _Selector(dict):
    """``collectiveSelector[cpu|gpu][singlenode|multinode][sync|async][name]`` lookup table."""

    def __init__(self, owner):
        super().__init__()
        names = {
            "allreduceTensor": owner.allreduceTensor,
            "allgatherTensor": owner.allgatherTensor,
            "broadcastTensor": owner.broadcastTensor,
        }
        for dev in ("cpu", "gpu"):
            self[dev] = {}
            for scope in ("singlenode", "multinode"):
                self[dev][scope] = {"sync": dict(names), "async": {}}
                for n, f in names.items():
                    self[dev][scope]["async"][n] = owner._make_async(f)


class _Handle:
    """Common TorchMPI-shaped API."""

    needInterNodeCollectives = False

    def _make_async(self, fn):
        def run(*a, **k):
            fn(*a, **k)
            return _DoneHandle()
        return run

    @property
    def collectiveSelector(self):
        sel = self.__dict__.get("_selector")
        if sel is None:
            sel = _Selector(self)
            self.__dict__["_selector"] = sel
        return sel

    # snake_case aliases
    def allreduce_tensor(self, t):
        return self.allreduceTensor(t)

    def allgather_tensor(self, inp, out):
        return self.allgatherTensor(inp, out)

    def broadcast_tensor(self, root, t):
        return self.broadcastTensor(root, t)

    def syncHandle(self, h):
        """TorchMPI's ``mpi.syncHandle(handle)``: block until the asynchronous collective is complete."""
        return h.wait() if h is not None else None


class _DoneHandle:
    """Handle of a collective that needs no further waiting: FakeWorld collectives are synchronous, and
    a CUDA collective is already asynchronous with respect to the host (it is ordered on the stream)."""

    def wait(self):
        return self


class _WorkHandle(_DoneHandle):
    """An in-flight ``torch.distributed`` operation (CPU / gloo)."""

    def __init__(self, work, after=None):
        self._work, self._after = work, after

    def wait(self):
        if self._work is not None:
            self._work.wait()
            self._work = None
            if self._after is not None:
                self._after()
        return self


# ----------------------------------------------------------------------------
# real backend
# ----------------------------------------------------------------------------
class TorchMPI(_Handle):
    """Process-group handle.  ``start(True)`` binds one CUDA device per rank."""

    def __init__(self):
        self._started = False
        self._owns_pg = False
        self._use_gpu = False
        self._group = None   # None = the world; else the torch.distributed group of a split() handle
        self._ranks: Optional[List[int]] = None  # global ranks of the group's members, in group-rank order
        self.symm = None  # tadnn.parallel.symm.SymmContext once the GPU fabric is up

    def _make_async(self, fn):
        """The ``async`` half of the selector table.  CPU tensors: the collective is started with
        ``async_op=True`` and the handle's ``wait()`` / ``mpi.syncHandle`` completes it, so host work can
        overlap the transfer.  CUDA tensors: the collective is enqueued on the current stream like the
        sync variant — the host does not block either way, and stream order gives the dependency."""
        name = fn.__name__

        def run(*a, **k):
            import torch.distributed as dist
            t = a[1] if name == "broadcastTensor" else a[0]
            if (name in ("allreduceTensor", "broadcastTensor") and torch.is_tensor(t) and not t.is_cuda
                    and t.is_contiguous() and dist.is_initialized() and self.size() > 1):
                if name == "allreduceTensor":
                    return _WorkHandle(dist.all_reduce(t, op=dist.ReduceOp.SUM, group=self._group, async_op=True))
                return _WorkHandle(dist.broadcast(t, src=self.global_rank(a[0]), group=self._group, async_op=True))
            fn(*a, **k)
            return _DoneHandle()
        return run

    # -- lifecycle -------------------------------------------------------------
    def start(self, use_gpu: bool = True, backend: Optional[str] = None, *, rank: Optional[int] = None,
              world_size: Optional[int] = None, init_method: Optional[str] = None,
              timeout_s: float = 600.0, symmetric: Optional[bool] = None):
        import torch.distributed as dist
        if self._started:
            return self
        use_gpu = bool(use_gpu) and torch.cuda.is_available()
        self._use_gpu = use_gpu
        if not dist.is_initialized():
            env_rank = int(os.environ.get("RANK", "0")) if rank is None else rank
            env_world = int(os.environ.get("WORLD_SIZE", "1")) if world_size is None else world_size
            if use_gpu:
                local = int(os.environ.get("LOCAL_RANK", env_rank % max(1, torch.cuda.device_count())))
                torch.cuda.set_device(local)
            if backend is None:
                backend = "cpu:gloo,cuda:nccl" if use_gpu else "gloo"
            if init_method is None:
                os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
                os.environ.setdefault("MASTER_PORT", "29511")
                init_method = "env://"
            kw = {}
            if use_gpu:
                kw["device_id"] = torch.device("cuda", torch.cuda.current_device())
            dist.init_process_group(backend=backend, init_method=init_method, rank=env_rank,
                                    world_size=env_world,
                                    timeout=datetime.timedelta(seconds=timeout_s), **kw)
            self._owns_pg = True
        self._started = True
        hosts: List[Optional[str]] = [None] * dist.get_world_size()
        if dist.get_world_size() > 1:
            dist.all_gather_object(hosts, socket.gethostname())
            self.needInterNodeCollectives = len(set(hosts)) > 1
        if symmetric is None:
            symmetric = use_gpu and os.environ.get("TADNN_FORCE_PATH", "") != "nccl"
        if symmetric and use_gpu and not self.needInterNodeCollectives:
            from .parallel import symm
            self.symm = symm.SymmContext.create(self)
        global _default
        if _default is None:
            _default = self
        return self

    def split(self, color: int, key: Optional[int] = None) -> "TorchMPI":
        """``MPI_Comm_split``: ranks that pass the same ``color`` form a new communicator, ordered by
        ``key`` (default: the rank in this one).  Collective over this handle.  The sub-handle has the
        same API (group-relative ``rank()``/``size()``); its collectives go through ``torch.distributed``
        on the sub-group (gloo / NCCL) — the symmetric-memory kernels serve the world communicator.
        Not in the reference (its DP and MP paths share the world communicator); this is
        what makes a DP x TP mesh expressible: see :func:`mesh`."""
        import torch.distributed as dist
        assert self._started, "split() needs a started handle"
        P = self.size()
        me = (int(color), int(self.rank() if key is None else key), dist.get_rank())
        everyone: list = [None] * P
        dist.all_gather_object(everyone, me, group=self._group)
        ranks = [g for _c, _k, g in sorted(e for e in everyone if e[0] == me[0])]
        sub = TorchMPI()
        sub._started, sub._use_gpu = True, self._use_gpu
        sub._ranks = ranks
        # only the members create their group (no world-wide call needed, so sub-handles can be split again)
        sub._group = dist.new_group(ranks, use_local_synchronization=True) if len(ranks) > 1 else None
        if self.needInterNodeCollectives:
            hosts: list = [None] * len(ranks)
            if sub._group is not None:
                dist.all_gather_object(hosts, socket.gethostname(), group=sub._group)
            sub.needInterNodeCollectives = len(set(hosts)) > 1
        return sub

    def stop(self):
        import torch.distributed as dist
        if self._ranks is not None:  # a split() handle: drop the sub-group only
            if self._group is not None and dist.is_initialized():
                dist.destroy_process_group(self._group)
            self._group, self._started = None, False
            return
        if self.symm is not None:
            self.symm.close()
            self.symm = None
        if self._owns_pg and dist.is_initialized():
            dist.destroy_process_group()
        self._started = False
        self._owns_pg = False
        global _default
        if _default is self:
            _default = None

    @property
    def started(self) -> bool:
        return self._started

    @property
    def use_gpu(self) -> bool:
        return self._use_gpu

    # -- queries ---------------------------------------------------------------
    def rank(self) -> int:
        import torch.distributed as dist
        if self._ranks is not None:
            return self._ranks.index(dist.get_rank())
        return dist.get_rank() if dist.is_initialized() else 0

    def size(self) -> int:
        import torch.distributed as dist
        if self._ranks is not None:
            return len(self._ranks)
        return dist.get_world_size() if dist.is_initialized() else 1

    def global_rank(self, group_rank: int) -> int:
        """World rank of member ``group_rank`` of this communicator."""
        return self._ranks[group_rank] if self._ranks is not None else group_rank

    def device(self) -> torch.device:
        return torch.device("cuda", torch.cuda.current_device()) if self._use_gpu else torch.device("cpu")

    # -- collectives -----------------------------------------------------------
    def barrier(self):
        """Host-level barrier (datamodule.lua:50,283).  Always ``torch.distributed``'s: ranks may
        arrive arbitrarily late here (rank 0 writing a checkpoint, uneven data loading, a profiler
        replaying kernels), and a host barrier waits without occupying the GPU.  The device-side
        flag barrier (csrc/comm.cu K5) is used *inside* the collective kernels only, where all
        ranks are already within microseconds of each other; ``symm.barrier()`` exposes it."""
        import torch.distributed as dist
        if dist.is_initialized() and self.size() > 1:
            if self._use_gpu:
                torch.cuda.synchronize()
            dist.barrier(group=self._group)

    def nccl_only(self) -> "TorchMPI":
        """A view of this communicator whose collectives never use the symmetric-memory kernels
        (``torch.distributed``/NCCL only) — the comparator arm of the benchmarks, in-process."""
        v = TorchMPI()
        v.__dict__.update(self.__dict__)
        v.symm = None
        v._owns_pg = False
        return v

    def allreduceTensor(self, t: torch.Tensor) -> torch.Tensor:
        """In-place sum over all ranks (nodemodule.lua:52, datamodule.lua:216)."""
        import torch.distributed as dist
        if self.size() == 1:
            return t
        if t.is_cuda and self.symm is not None and self.symm.supports(t):
            self.symm.allreduce_(t, scale=1.0)
            return t
        if t.is_contiguous():
            dist.all_reduce(t, op=dist.ReduceOp.SUM, group=self._group)
        else:
            c = t.contiguous()
            dist.all_reduce(c, op=dist.ReduceOp.SUM, group=self._group)
            t.copy_(c)
        return t

    def allgatherTensor(self, inp: torch.Tensor, out: torch.Tensor) -> torch.Tensor:
        """``out`` = rank-ordered concatenation of every rank's ``inp`` along the LAST dim
        (the layout the reference's callers assume, nodemodule.lua:161-167,261-267)."""
        import torch.distributed as dist
        P = self.size()
        if P == 1:
            out.copy_(inp.reshape(out.shape))
            return out
        if inp.is_cuda and self.symm is not None and self.symm.supports(inp):
            self.symm.allgather_lastdim(inp, out)
            return out
        parts = [torch.empty_like(inp, memory_format=torch.contiguous_format) for _ in range(P)]
        dist.all_gather(parts, inp.contiguous(), group=self._group)
        out.copy_(torch.cat(parts, dim=-1).reshape(out.shape))
        return out

    def broadcastTensor(self, root: int, t: torch.Tensor) -> torch.Tensor:
        import torch.distributed as dist
        if self.size() == 1:
            return t
        src = self.global_rank(root)  # torch.distributed wants the world rank of the source
        if t.is_contiguous():
            dist.broadcast(t, src=src, group=self._group)
        else:
            c = t.contiguous()
            dist.broadcast(c, src=src, group=self._group)
            t.copy_(c)
        return t


# ----------------------------------------------------------------------------
# fake backend: P ranks = P threads, list-of-tensors communicator
# ----------------------------------------------------------------------------
class FakeWorld:
    """Single-process simulation of ``size`` ranks.  ``run(fn)`` executes ``fn(mpi)`` once per
    rank on its own thread, with ``tadnn.mpi.current()`` bound to that rank's ``FakeMPI``."""

    def __init__(self, size: int):
        assert size >= 1
        self.size = size
        self._slots: List[Optional[torch.Tensor]] = [None] * size
        self._bar = threading.Barrier(size)
        self.calls = {"allreduce": 0, "allgather": 0, "broadcast": 0, "barrier": 0}
        self.bytes = 0
        self._stat_lock = threading.Lock()
        self._subworlds: dict = {}  # (split epoch, color) -> FakeWorld shared by that colour's ranks

    def handle(self, rank: int) -> "FakeMPI":
        return FakeMPI(rank, self)

    def run(self, fn: Callable[["FakeMPI"], object], timeout_s: float = 120.0) -> list:
        results: list = [None] * self.size
        errors: list = [None] * self.size

        def body(r):
            h = self.handle(r)
            set_current(h)
            try:
                results[r] = fn(h)
            except BaseException as e:  # noqa: BLE001 - re-raised on the caller thread
                errors[r] = e
                self._bar.abort()
            finally:
                set_current(None)

        threads = [threading.Thread(target=body, args=(r,), daemon=True) for r in range(self.size)]
        for t in threads:
            t.start()
        for t in threads:
            t.join(timeout_s)
            if t.is_alive():
                self._bar.abort()
                raise TimeoutError("FakeWorld.run: a rank did not finish (collective mismatch?)")
        for e in errors:
            if e is not None and not isinstance(e, threading.BrokenBarrierError):
                raise e
        for e in errors:
            if e is not None:
                raise e
        return results

    def _count(self, what: str, nbytes: int = 0):
        with self._stat_lock:
            self.calls[what] += 1
            self.bytes += nbytes


class FakeMPI(_Handle):
    def __init__(self, rank: int, world: FakeWorld):
        self._rank = rank
        self._world = world
        self.symm = None
        self._nsplit = 0

    def split(self, color: int, key: Optional[int] = None) -> "FakeMPI":
        """``MPI_Comm_split`` among the threads of this FakeWorld (see :meth:`TorchMPI.split`)."""
        w = self._world
        w._slots[self._rank] = (int(color), int(self._rank if key is None else key), self._rank)
        w._bar.wait()
        everyone = list(w._slots)
        w._bar.wait()
        ranks = [r for _c, _k, r in sorted(e for e in everyone if e[0] == int(color))]
        epoch = self._nsplit  # every rank splits in the same order, so the counter agrees
        self._nsplit += 1
        with w._stat_lock:
            sub = w._subworlds.setdefault((epoch, int(color)), FakeWorld(len(ranks)))
        h = FakeMPI(ranks.index(self._rank), sub)
        h._ranks = ranks
        return h

    def global_rank(self, group_rank: int) -> int:
        ranks = getattr(self, "_ranks", None)
        return ranks[group_rank] if ranks is not None else group_rank

    started = True
    use_gpu = False

    def start(self, *a, **k):
        return self

    def stop(self):
        pass

    def rank(self) -> int:
        return self._rank

    def size(self) -> int:
        return self._world.size

    def device(self):
        return torch.device("cpu")

    def barrier(self):
        if self._rank == 0:
            self._world._count("barrier")
        self._world._bar.wait()

    def allreduceTensor(self, t):
        w = self._world
        if self._rank == 0:
            w._count("allreduce", t.numel() * t.element_size())
        w._slots[self._rank] = t.detach().clone()
        w._bar.wait()
        total = w._slots[0].clone()
        for r in range(1, w.size):  # fixed rank order: bit-identical on every rank
            total += w._slots[r]
        w._bar.wait()
        with torch.no_grad():
            t.copy_(total)
        return t

    def allgatherTensor(self, inp, out):
        w = self._world
        if self._rank == 0:
            w._count("allgather", inp.numel() * inp.element_size())
        w._slots[self._rank] = inp.detach().clone()
        w._bar.wait()
        cat = torch.cat([w._slots[r] for r in range(w.size)], dim=-1)
        w._bar.wait()
        with torch.no_grad():
            out.copy_(cat.reshape(out.shape))
        return out

    def broadcastTensor(self, root, t):
        w = self._world
        if self._rank == 0:
            w._count("broadcast", t.numel() * t.element_size())
        if self._rank == root:
            w._slots[root] = t.detach().clone()
        w._bar.wait()
        src = w._slots[root]
        w._bar.wait()
        with torch.no_grad():
            t.copy_(src)
        return t


# ----------------------------------------------------------------------------
# module-level, TorchMPI-style API  (``mpi = require('torchmpi')``)
# ----------------------------------------------------------------------------
def current():
    """The handle the "global ``mpi``" of nodemodule.lua resolves to: a thread-local override
    (FakeWorld ranks), else the process default (created on demand, not started)."""
    h = getattr(_tls, "handle", None)
    if h is not None:
        return h
    global _default
    if _default is None:
        _default = TorchMPI()
    return _default


def set_current(handle) -> None:
    _tls.handle = handle


def start(use_gpu: bool = True, *a, **k):
    return current().start(use_gpu, *a, **k)


def stop():
    return current().stop()


def started() -> bool:
    return bool(current().started)


def rank() -> int:
    return current().rank()


def size() -> int:
    return current().size()


def barrier():
    return current().barrier()


def allreduceTensor(t):
    return current().allreduceTensor(t)


def allgatherTensor(inp, out):
    return current().allgatherTensor(inp, out)


def broadcastTensor(root, t):
    return current().broadcastTensor(root, t)


def mesh(tp_size: int, handle=None):
    """Split ``handle`` (default: the current one) into a DP x TP mesh and return ``(dp, tp)``:
    ``tp`` joins ``tp_size`` consecutive ranks (the model-parallel group: pass it to the ``MP*`` layers),
    ``dp`` joins the ranks that hold the same shard in every such group (pass it to ``parallelize`` /
    ``synchronizeModel``).  The reference cannot combine its two modes because both use the world
    communicator (SURVEY §2.3 "Hybrid DP x TP"); with two communicators they compose."""
    h = handle if handle is not None else current()
    P, r = h.size(), h.rank()
    assert tp_size >= 1 and P % tp_size == 0, f"world size {P} is not a multiple of tp_size {tp_size}"
    tp = h.split(r // tp_size, r % tp_size)
    dp = h.split(r % tp_size, r // tp_size)
    return dp, tp


def __getattr__(name):  # module-level attribute access for the two TorchMPI fields
    if name == "collectiveSelector":
        return current().collectiveSelector
    if name == "needInterNodeCollectives":
        return current().needInterNodeCollectives
    raise AttributeError(name)
