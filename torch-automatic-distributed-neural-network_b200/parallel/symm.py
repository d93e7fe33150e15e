"""Symmetric-memory context: one peer-mapped heap per GPU + the collective kernels on it.

This is the B200 replacement for the TorchMPI library the reference links against
(SURVEY §2.4): every rank allocates an identically sized heap (followed by a signal pad) and
maps its peers' heaps, after which hand-written sm_100a kernels (csrc/comm.cu) do every
collective with plain loads/stores over NVLink 5 / NVSwitch — no NCCL on this path.

Two backends (``TADNN_SYMM_BACKEND=auto|vmm|ipc``).  ``vmm`` (preferred): ``cuMemCreate`` heaps
shared as POSIX file descriptors over an AF_UNIX socket, plus — where the driver exposes NVLS — a
multicast object bound over all heaps, which enables the ``multimem`` all-reduce and the in-switch
reduce-scatter (csrc/comm_vmm.cu).  ``ipc``: ``cudaMalloc`` + CUDA-IPC handles exchanged through
``torch.distributed`` (plumbing only); chosen automatically if any rank fails the VMM bring-up.

Allocation is a symmetric bump allocator: all ranks must call ``alloc`` in the same order
with the same sizes (they do: the callers are SPMD).  ``tensor()`` returns a zero-copy
``torch.Tensor`` view of heap memory, so model parameters and gradients can *live* in the
heap (tadnn.parallel.fusion) and be reduced in place.  ``alloc``/``free`` manage the heap with a
deterministic first-fit free list; size it with ``TADNN_SYMM_HEAP_MB``.

Failure detection: a barrier inside a collective kernel that waits longer than the timeout
(``set_timeout`` / ``TADNN_COMM_TIMEOUT_S``, default 120 s) records which rank it was waiting for
and lets the stream drain; ``check()`` — called before every collective launch — turns that into a
``CommTimeout`` exception that names the missing rank (the reference just hangs).
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional

import torch

from .. import _native

_DT = {torch.float32: 0, torch.bfloat16: 1}

# signal-pad channels: collectives that can be in flight concurrently use different ones
CH_DEFAULT = 0   # mpi.allreduceTensor & friends on the compute stream
CH_DP = 1        # data-parallel gradient/parameter sync on the comm stream
CH_TP = 2        # model-parallel layer collectives
CH_BARRIER = 3
LL_MAX_FLOATS = 131072  # TADNN_LL_MAX_FLOATS: largest message of the low-latency all-reduce


class CommTimeout(_native.NativeError):
    """A device-side collective gave up waiting for a peer (dead or wedged rank)."""


class _RawCuda:
    """Minimal ``__cuda_array_interface__`` carrier for zero-copy torch views of heap memory."""

    def __init__(self, ptr: int, nbytes: int, owner):
        self.owner = owner
        self.__cuda_array_interface__ = {
            "shape": (int(nbytes),), "typestr": "|u1", "data": (int(ptr), False),
            "version": 3, "strides": None,
        }


class SymmContext:
    def __init__(self):
        self.handle = C.c_void_p()
        self.rank = 0
        self.world = 1
        self.heap_bytes = 0
        self.heap_ptrs: list[int] = []
        self.has_multicast = False
        self._bump = 0          # high-water mark (diagnostics)
        self._free: list = []   # sorted [(offset, size)] free blocks
        self._live: dict = {}   # offset -> size of live allocations
        self._heap_u8: Optional[torch.Tensor] = None
        self._staging_off = 0
        self._staging_bytes = 0
        self.mpi = None
        self.device = None

    # ------------------------------------------------------------------ creation
    @classmethod
    def create(cls, mpi, heap_mb: Optional[int] = None, staging_mb: Optional[int] = None) -> "SymmContext":
        import torch.distributed as dist
        self = cls()
        lib = _native.lib()
        self.mpi = mpi
        self.rank, self.world = mpi.rank(), mpi.size()
        if self.world > 8:
            raise _native.NativeError("tadnn symmetric memory supports up to 8 ranks (one NVSwitch domain)")
        heap_mb = int(os.environ.get("TADNN_SYMM_HEAP_MB", heap_mb or 1024))
        staging_mb = int(os.environ.get("TADNN_SYMM_STAGING_MB", staging_mb or 128))
        self.heap_bytes = heap_mb << 20
        self.device = torch.device("cuda", torch.cuda.current_device())
        backend = os.environ.get("TADNN_SYMM_BACKEND", "auto")
        self.backend = "ipc"
        if backend in ("auto", "vmm") and self.world > 1:
            ok, why = self._try_vmm(lib)
            flags: list = [None] * self.world
            dist.all_gather_object(flags, (ok, why))
            if all(f[0] for f in flags):
                self.backend = "vmm"
            else:
                if self.handle:
                    lib.tadnn_comm_destroy(self.handle)
                    self.handle = C.c_void_p()
                self.has_multicast = False
                if backend == "vmm":
                    raise _native.NativeError(f"TADNN_SYMM_BACKEND=vmm failed: {[f[1] for f in flags if not f[0]]}")
                if self.rank == 0:
                    print(f"[tadnn] VMM/multicast heap unavailable ({[f[1] for f in flags if not f[0]][:1]}); using CUDA-IPC", flush=True)
        if self.backend == "ipc":
            _native.check(lib.tadnn_comm_create(C.c_int(self.rank), C.c_int(self.world),
                                                C.c_ulonglong(self.heap_bytes), C.byref(self.handle)),
                          "tadnn_comm_create")
            if self.world > 1:
                mine = C.create_string_buffer(64)
                _native.check(lib.tadnn_comm_ipc_handle(self.handle, mine), "tadnn_comm_ipc_handle")
                allh: list = [None] * self.world
                dist.all_gather_object(allh, bytes(mine.raw))
                blob = b"".join(allh)
                _native.check(lib.tadnn_comm_open_peers(self.handle, C.c_char_p(blob)), "tadnn_comm_open_peers")
        self.heap_ptrs = []
        for r in range(self.world):
            p = C.c_void_p()
            _native.check(lib.tadnn_comm_heap(self.handle, C.c_int(r), C.byref(p)), "tadnn_comm_heap")
            self.heap_ptrs.append(int(p.value))
        raw = _RawCuda(self.heap_ptrs[self.rank], self.heap_bytes, self)
        self._heap_u8 = torch.as_tensor(raw, device=self.device)
        assert self._heap_u8.data_ptr() == self.heap_ptrs[self.rank], "zero-copy heap view failed"
        self._free = [(0, self.heap_bytes)]
        self._live = {}
        self._staging_bytes = staging_mb << 20
        self._staging_off = self.alloc(self._staging_bytes)
        # low-latency all-reduce staging ({value, epoch} pairs), one area per channel that uses it; zeroed before the
        # barrier below, so no peer can write into it earlier
        self._ll = {}
        if self.world > 1:
            ll_bytes = 2 * self.world * LL_MAX_FLOATS * 8
            for ch in (CH_DEFAULT, CH_DP, CH_TP):
                off = self.alloc(ll_bytes)
                self._heap_u8[off:off + ll_bytes].zero_()
                self._ll[ch] = [off, 0]
        if self.world > 1:
            torch.cuda.synchronize()
            dist.barrier()  # pads are zeroed and mapped everywhere before the first kernel
        return self

    # ------------------------------------------------------------------ VMM + multicast bring-up
    def _fd_exchange(self, token: str, send: dict) -> dict:
        """Pass file descriptors between the rank processes over AF_UNIX datagram sockets
        (SCM_RIGHTS).  ``send`` maps destination rank -> fd; returns {source rank: fd} for every
        rank that sent us one (every rank must call this with the same ``token``)."""
        import socket
        import struct
        import torch.distributed as dist
        addr = lambda r: f"\0tadnn-{token}-{r}"  # abstract namespace: no files to clean up
        srv = socket.socket(socket.AF_UNIX, socket.SOCK_DGRAM)
        srv.bind(addr(self.rank))
        srv.settimeout(60.0)
        expect: list = [None] * self.world
        dist.all_gather_object(expect, sorted(send.keys()))
        n_in = sum(1 for src, dsts in enumerate(expect) if self.rank in dsts and src != self.rank)
        out = socket.socket(socket.AF_UNIX, socket.SOCK_DGRAM)
        import array
        for dst, fd in send.items():  # (socket.send_fds drops its address argument before 3.13)
            out.sendmsg([struct.pack("i", self.rank)],
                        [(socket.SOL_SOCKET, socket.SCM_RIGHTS, array.array("i", [fd]))], 0, addr(dst))
        got = {}
        for _ in range(n_in):
            msg, anc, _flags, _a = srv.recvmsg(16, socket.CMSG_LEN(4))
            fds = array.array("i")
            for level, typ, data in anc:
                if level == socket.SOL_SOCKET and typ == socket.SCM_RIGHTS:
                    fds.frombytes(data[:4])
            got[struct.unpack("i", msg[:4])[0]] = fds[0]
        out.close()
        srv.close()
        dist.barrier()
        return got

    def _try_vmm(self, lib):
        import torch.distributed as dist
        try:
            tok = [None]
            if self.rank == 0:
                tok[0] = f"{os.getpid()}-{int.from_bytes(os.urandom(4), 'little')}"
            dist.broadcast_object_list(tok, src=0)
            def agree(ok: bool) -> bool:
                # every phase ends in a vote, so that all ranks leave this function at the same point
                allok: list = [None] * self.world
                dist.all_gather_object(allok, bool(ok))
                return all(allok)

            why = ""
            fd = C.c_int(-1)
            ok = lib.tadnn_comm_create_vmm(C.c_int(self.rank), C.c_int(self.world), C.c_ulonglong(self.heap_bytes),
                                           C.byref(self.handle)) == 0
            ok = ok and lib.tadnn_comm_export_fd(self.handle, C.byref(fd)) == 0
            if not ok:
                why = lib.tadnn_last_error().decode()
            if not agree(ok):
                if fd.value >= 0:
                    os.close(fd.value)
                return False, why or "a peer failed to create its VMM heap"
            got = self._fd_exchange(tok[0] + "-heap", {q: fd.value for q in range(self.world) if q != self.rank})
            os.close(fd.value)
            for q, pfd in got.items():  # import_peer_fd owns (and closes) the descriptor in every case
                if lib.tadnn_comm_import_peer_fd(self.handle, C.c_int(q), C.c_int(pfd)) != 0:
                    ok, why = False, lib.tadnn_last_error().decode()
            if not agree(ok):
                return False, why or "a peer failed to import a heap"
            # multicast (optional): failure here keeps the VMM heap, just without NVLS
            mc = C.c_int(0)
            lib.tadnn_multicast_supported(C.byref(mc))
            want = bool(mc.value) and os.environ.get("TADNN_NVLS", "1") not in ("0", "false")
            votes: list = [None] * self.world
            dist.all_gather_object(votes, want)
            if all(votes):
                mfd = C.c_int(-1)
                ok = True
                if self.rank == 0:
                    ok = lib.tadnn_comm_mc_create(self.handle, C.byref(mfd)) == 0
                oks: list = [None] * self.world
                dist.all_gather_object(oks, ok)
                if all(oks):
                    got = self._fd_exchange(tok[0] + "-mc", {q: mfd.value for q in range(1, self.world)} if self.rank == 0 else {})
                    if self.rank == 0:
                        os.close(mfd.value)
                    else:
                        ok = lib.tadnn_comm_mc_import(self.handle, C.c_int(got[0])) == 0
                    ok = ok and lib.tadnn_comm_mc_add_device(self.handle) == 0
                    dist.all_gather_object(oks, ok)
                    if all(oks):
                        dist.barrier()
                        ok = lib.tadnn_comm_mc_bind(self.handle) == 0
                        dist.all_gather_object(oks, ok)
                        self.has_multicast = all(oks)
                        if not self.has_multicast and self.rank == 0:
                            print(f"[tadnn] multicast bind failed: {lib.tadnn_last_error().decode()}", flush=True)
            return True, ""
        except Exception as e:  # noqa: BLE001 - any failure means "fall back to IPC"
            return False, repr(e)

    def close(self):
        if self.handle:
            torch.cuda.synchronize()
            self._heap_u8 = None
            _native.lib().tadnn_comm_destroy(self.handle)
            self.handle = C.c_void_p()

    # ------------------------------------------------------------------ memory
    def alloc(self, nbytes: int, align: int = 256) -> int:
        """Symmetric allocation: first fit over a free list, 256-byte granules.  Deterministic, so
        ranks that make the same calls in the same order get the same offsets (the callers are
        SPMD).  Returns the heap offset; release it with :meth:`free`."""
        gran = 256
        align = max(gran, (align + gran - 1) // gran * gran)
        size = max(gran, (int(nbytes) + gran - 1) // gran * gran)
        for i, (foff, fsize) in enumerate(self._free):
            off = (foff + align - 1) // align * align
            if off + size <= foff + fsize:
                pieces = []
                if off > foff:
                    pieces.append((foff, off - foff))
                if off + size < foff + fsize:
                    pieces.append((off + size, foff + fsize - off - size))
                self._free[i:i + 1] = pieces
                self._live[off] = size
                self._bump = max(self._bump, off + size)
                return off
        raise MemoryError(
            f"tadnn symmetric heap exhausted: {size} bytes requested, {self.free_bytes()} free "
            f"(largest block {max((f[1] for f in self._free), default=0)}) of {self.heap_bytes}; "
            "raise TADNN_SYMM_HEAP_MB or free() buffers that are no longer used")

    def free(self, off: int) -> None:
        """Return an allocation to the heap.  Like ``alloc`` this is a collective in spirit: every
        rank must free the same offsets in the same order, and no kernel (local or remote) may still
        be using the memory — synchronise first if in doubt."""
        size = self._live.pop(int(off), None)
        if size is None:
            raise ValueError(f"free({off}): not a live symmetric allocation")
        import bisect
        i = bisect.bisect_left(self._free, (off, 0))
        self._free.insert(i, (off, size))
        # merge with the neighbours
        if i + 1 < len(self._free) and self._free[i][0] + self._free[i][1] == self._free[i + 1][0]:
            self._free[i:i + 2] = [(self._free[i][0], self._free[i][1] + self._free[i + 1][1])]
        if i > 0 and self._free[i - 1][0] + self._free[i - 1][1] == self._free[i][0]:
            self._free[i - 1:i + 1] = [(self._free[i - 1][0], self._free[i - 1][1] + self._free[i][1])]

    def free_tensor(self, t: torch.Tensor) -> None:
        off = self.offset_of(t)
        if off is None:
            raise ValueError("free_tensor: not a tensor of this heap")
        self.free(off)

    def free_bytes(self) -> int:
        return sum(f[1] for f in self._free)

    def live_bytes(self) -> int:
        return sum(self._live.values())

    # ------------------------------------------------------------------ failure detection
    def set_timeout(self, seconds: float) -> None:
        """How long a device-side barrier waits for a peer before it gives up (default 120 s or
        ``TADNN_COMM_TIMEOUT_S``)."""
        _native.check(_native.lib().tadnn_comm_set_timeout(self.handle, C.c_double(seconds)), "tadnn_comm_set_timeout")

    def check(self) -> None:
        """Raise :class:`CommTimeout` if a collective on this context has timed out waiting for a
        peer (the message names the ranks).  One read of host memory; call it as often as you like."""
        if self.handle and _native.lib().tadnn_comm_check(self.handle) != 0:
            raise CommTimeout(_native.lib().tadnn_last_error().decode("utf-8", "replace"))

    def tensor(self, off: int, shape, dtype: torch.dtype) -> torch.Tensor:
        n = 1
        for s in shape:
            n *= int(s)
        nbytes = n * torch.empty((), dtype=dtype).element_size()
        return self._heap_u8[off:off + nbytes].view(dtype).view(*shape)

    def empty(self, shape, dtype: torch.dtype) -> torch.Tensor:
        n = 1
        for s in shape:
            n *= int(s)
        off = self.alloc(n * torch.empty((), dtype=dtype).element_size())
        return self.tensor(off, shape, dtype)

    def offset_of(self, t: torch.Tensor) -> Optional[int]:
        """Heap offset of ``t`` if it is a dense view of this rank's heap, else None."""
        base = self.heap_ptrs[self.rank]
        p = t.data_ptr()
        nbytes = t.numel() * t.element_size()
        if t.is_cuda and t.is_contiguous() and base <= p and p + nbytes <= base + self.heap_bytes:
            return p - base
        return None

    def peer_ptr(self, peer: int, off: int) -> int:
        return self.heap_ptrs[peer] + off

    def supports(self, t: torch.Tensor) -> bool:
        if not (t.is_cuda and t.dtype in _DT):
            return False
        return self.offset_of(t) is not None or t.numel() * t.element_size() <= self._staging_bytes

    # ------------------------------------------------------------------ collectives
    def barrier(self, channel: int = CH_BARRIER, stream=None):
        self.check()
        _native.check(_native.lib().tadnn_barrier(self.handle, C.c_int(channel),
                                                  C.c_void_p(_native.stream_ptr(stream))), "tadnn_barrier")

    def _ll_ok(self, count: int, dtype, algo: int, channel: int) -> bool:
        return (algo in (0, 4) and dtype == torch.float32 and 0 < count <= LL_MAX_FLOATS and channel in self._ll
                and os.environ.get("TADNN_COMM_LL", "1") not in ("0", "false"))

    def _ll_launch(self, off: int, param_off: int, count: int, scale: float, sgd: bool, lr: float, wd: float, channel: int, stream):
        ent = self._ll[channel]
        ent[1] = ent[1] % 0xFFFFFFF0 + 1   # non-zero, same sequence on every rank
        _native.check(_native.lib().tadnn_allreduce_ll(
            self.handle, C.c_ulonglong(ent[0]), C.c_ulonglong(off), C.c_ulonglong(param_off), C.c_ulonglong(count),
            C.c_float(scale), C.c_int(int(sgd)), C.c_float(lr), C.c_float(wd), C.c_uint(ent[1]),
            C.c_void_p(_native.stream_ptr(stream))), "tadnn_allreduce_ll")

    def allreduce_offset(self, off: int, count: int, dtype: torch.dtype, scale: float = 1.0,
                         algo: int = 0, max_blocks: int = 0, channel: int = CH_DEFAULT, stream=None):
        """``algo``: 0 auto (low-latency pairs up to 512 KB of fp32, then one-shot / NVLS / two-shot by size),
        1 one-shot, 2 two-shot, 3 NVLS, 4 low-latency."""
        self.check()
        if self._ll_ok(count, dtype, algo, channel):
            return self._ll_launch(off, 0, count, scale, False, 0.0, 0.0, channel, stream)
        if algo == 4:
            raise ValueError("low-latency all-reduce: fp32 messages of at most 512 KB on channels 0-2 only")
        _native.check(_native.lib().tadnn_allreduce(
            self.handle, C.c_ulonglong(off), C.c_ulonglong(count), C.c_int(_DT[dtype]),
            C.c_float(scale), C.c_int(algo), C.c_int(max_blocks), C.c_int(channel),
            C.c_void_p(_native.stream_ptr(stream))), "tadnn_allreduce")

    def allreduce_sgd_offset(self, grad_off: int, param_off: int, count: int, scale: float, lr: float,
                             weight_decay: float = 0.0, algo: int = 0, channel: int = CH_DP, stream=None):
        """Gradient all-reduce-average fused with the SGD update and the gradient clear (fp32; csrc/comm.cu)."""
        self.check()
        if self._ll_ok(count, torch.float32, algo, channel):
            return self._ll_launch(grad_off, param_off, count, scale, True, lr, weight_decay, channel, stream)
        _native.check(_native.lib().tadnn_allreduce_sgd(
            self.handle, C.c_ulonglong(grad_off), C.c_ulonglong(param_off), C.c_ulonglong(count), C.c_float(scale),
            C.c_float(lr), C.c_float(weight_decay), C.c_int(algo), C.c_int(0), C.c_int(channel),
            C.c_void_p(_native.stream_ptr(stream))), "tadnn_allreduce_sgd")

    def allreduce_(self, t: torch.Tensor, scale: float = 1.0, algo: int = 0,
                   channel: int = CH_DEFAULT, stream=None) -> torch.Tensor:
        """In-place ``t = scale * sum_ranks(t)``.  Zero-copy when ``t`` lives in the heap,
        otherwise staged through the heap's staging area."""
        if self.world == 1:
            if scale != 1.0:
                t.mul_(scale)
            return t
        algo = int(os.environ.get("TADNN_FORCE_ALGO", algo)) if algo == 0 else algo
        off = self.offset_of(t)
        if off is not None and off % 16 == 0:
            self.allreduce_offset(off, t.numel(), t.dtype, scale, algo, 0, channel, stream)
            return t
        nbytes = t.numel() * t.element_size()
        if nbytes > self._staging_bytes:
            raise MemoryError("tensor larger than the symmetric staging area; raise TADNN_SYMM_STAGING_MB")
        st = self.tensor(self._staging_off, t.shape, t.dtype)
        st.copy_(t)
        self.allreduce_offset(self._staging_off, t.numel(), t.dtype, scale, algo, 0, channel, stream)
        t.copy_(st)
        return t

    def broadcast_offset(self, off: int, nbytes: int, root: int = 0, channel: int = CH_DEFAULT, stream=None):
        _native.check(_native.lib().tadnn_broadcast(
            self.handle, C.c_ulonglong(off), C.c_ulonglong(nbytes), C.c_int(root), C.c_int(0),
            C.c_int(channel), C.c_void_p(_native.stream_ptr(stream))), "tadnn_broadcast")

    def allgather_cols_offset(self, src: torch.Tensor, off: int, channel: int = CH_DEFAULT, stream=None):
        """Push local ``src[rows, w]`` into heap[off] of every rank viewed as [rows, world*w]."""
        assert src.dim() == 2 and src.stride(1) == 1
        es = src.element_size()
        _native.check(_native.lib().tadnn_allgather_cols(
            self.handle, C.c_void_p(src.data_ptr()), C.c_ulonglong(src.stride(0) * es),
            C.c_ulonglong(off), C.c_int(src.shape[0]), C.c_ulonglong(src.shape[1] * es), C.c_int(0),
            C.c_int(channel), C.c_void_p(_native.stream_ptr(stream))), "tadnn_allgather_cols")

    def allgather_lastdim(self, inp: torch.Tensor, out: torch.Tensor, channel: int = CH_DEFAULT, stream=None):
        """mpi.allgatherTensor semantics: ``out`` = rank-ordered concat of ``inp`` on the last dim."""
        w = inp.shape[-1]
        src = inp.reshape(-1, w)
        if src.stride(1) != 1:
            src = src.contiguous()
        rows = src.shape[0]
        nbytes = rows * w * self.world * inp.element_size()
        off = self.offset_of(out) if out.is_contiguous() else None
        if off is not None:
            self.allgather_cols_offset(src, off, channel, stream)
            return out
        if nbytes > self._staging_bytes:
            raise MemoryError("all-gather larger than the symmetric staging area")
        self.allgather_cols_offset(src, self._staging_off, channel, stream)
        out.copy_(self.tensor(self._staging_off, (rows, w * self.world), inp.dtype).view(out.shape))
        return out

    def reduce_scatter_cols_offset(self, off: int, dst: torch.Tensor, scale: float = 1.0,
                                   bias: Optional[torch.Tensor] = None, act: int = 0,
                                   channel: int = CH_DEFAULT, stream=None):
        assert dst.dim() == 2 and dst.stride(1) == 1 and dst.dtype in _DT
        _native.check(_native.lib().tadnn_reduce_scatter_cols(
            self.handle, C.c_ulonglong(off), C.c_void_p(dst.data_ptr()), C.c_ulonglong(dst.stride(0)),
            C.c_int(dst.shape[0]), C.c_int(dst.shape[1]), C.c_int(_DT[dst.dtype]), C.c_float(scale),
            C.c_void_p(_native.ptr(bias)), C.c_int(act), C.c_int(0), C.c_int(channel),
            C.c_void_p(_native.stream_ptr(stream))), "tadnn_reduce_scatter_cols")
