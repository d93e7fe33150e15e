"""Failure detection of the device-side barrier (csrc/comm.cu) on ONE GPU.

A two-rank context is assembled in this process out of two heaps, and only rank 0 ever launches a
collective: exactly what a survivor sees when its peer has died.  The reference would hang for
ever (every TorchMPI call blocks, datamodule.lua:50,216); here the kernel gives up after the
configured time, the stream drains, and the host gets an error that names the missing rank."""
import ctypes as C
import time

import pytest
import torch

pytestmark = pytest.mark.gpu


def test_dead_peer_becomes_a_python_exception_naming_the_rank():
    from tadnn import _native
    from tadnn.parallel.symm import CommTimeout, SymmContext
    lib = _native.lib()
    lib.tadnn_comm_set_timeout.argtypes = [C.c_void_p, C.c_double]
    heap = 1 << 20
    c0, c1 = C.c_void_p(), C.c_void_p()
    _native.check(lib.tadnn_comm_create(0, 2, C.c_ulonglong(heap), C.byref(c0)), "create 0")
    _native.check(lib.tadnn_comm_create(1, 2, C.c_ulonglong(heap), C.byref(c1)), "create 1")
    try:
        h1 = C.c_void_p()
        _native.check(lib.tadnn_comm_heap(c1, 1, C.byref(h1)), "heap 1")
        heaps = (C.c_void_p * 2)(None, h1)
        pads = (C.c_void_p * 2)(None, C.c_void_p(h1.value + heap))
        _native.check(lib.tadnn_comm_attach(c0, heaps, pads, None), "attach")
        _native.check(lib.tadnn_comm_set_timeout(c0, C.c_double(0.5)), "set_timeout")
        assert lib.tadnn_comm_check(c0) == 0
        ctx = SymmContext()
        ctx.handle, ctx.rank, ctx.world, ctx.heap_bytes = c0, 0, 2, heap
        s = C.c_void_p(_native.stream_ptr())
        t0 = time.perf_counter()
        # two collectives in a row: the first times out on rank 1, the second must fall through quickly
        _native.check(lib.tadnn_allreduce(c0, C.c_ulonglong(0), C.c_ulonglong(1024), 0, C.c_float(0.5), 1, 0, 0, s), "allreduce")
        _native.check(lib.tadnn_allreduce(c0, C.c_ulonglong(0), C.c_ulonglong(1024), 0, C.c_float(0.5), 2, 0, 0, s), "allreduce")
        torch.cuda.synchronize()  # no trap: the context is still usable
        dt = time.perf_counter() - t0
        assert 0.4 < dt < 5.0, dt
        with pytest.raises(CommTimeout, match=r"rank 0 waited .* for rank 1"):
            ctx.check()
        with pytest.raises(CommTimeout):
            ctx.allreduce_offset(0, 1024, torch.float32)  # refused before anything is launched
        assert torch.ones(4, device="cuda").sum().item() == 4.0  # CUDA itself is fine
        ctx.handle = C.c_void_p()
    finally:
        lib.tadnn_comm_destroy(c0)
        lib.tadnn_comm_destroy(c1)
