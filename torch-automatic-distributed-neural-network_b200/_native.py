"""ctypes binding of libtadnn.so (the sm_100a CUDA core, C ABI in include/tadnn.h).

The library is built in-tree by ``build_native.py`` (plain nvcc, ``-gencode
arch=compute_100a,code=sm_100a``).  On a machine with a GPU a missing library is a hard
error: the CUDA path never silently falls back to eager PyTorch.
"""
from __future__ import annotations

import ctypes as C
import os
import threading

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "lib", "libtadnn.so")

_lib = None
_lock = threading.Lock()


class NativeError(RuntimeError):
    pass


class GemmArgs(C.Structure):
    _fields_ = [
        ("M", C.c_int), ("N", C.c_int), ("K", C.c_int),
        ("A", C.c_void_p), ("lda", C.c_int), ("a_mn_major", C.c_int),
        ("B", C.c_void_p), ("ldb", C.c_int), ("b_mn_major", C.c_int),
        ("C", C.c_void_p), ("ldc", C.c_int), ("c_fp32", C.c_int),
        ("bias", C.c_void_p), ("act", C.c_int), ("alpha", C.c_float),
        ("accumulate", C.c_int), ("split_k", C.c_int),
        ("n_push", C.c_int), ("push", C.c_void_p * 8),
        ("push_ldc", C.c_int), ("push_col_offset", C.c_int),
        ("tile_n", C.c_int), ("rs_width", C.c_int), ("aux_tanh_out", C.c_void_p), ("ld_aux", C.c_int), ("n_rotate", C.c_int), ("rs_slab_major", C.c_int),
        ("sig_flags", C.c_void_p * 8), ("n_sig", C.c_int), ("sig_src", C.c_int), ("flag_stride", C.c_int),
        ("wait_flags", C.c_void_p), ("wait_target", C.c_uint), ("wait_cols_per_src", C.c_int),
        ("kblock_mask", C.POINTER(C.c_uint)), ("n_kblock_mask", C.c_int),
    ]


def available() -> bool:
    return os.path.exists(LIB_PATH)


def _declare(lib) -> None:
    lib.tadnn_last_error.restype = C.c_char_p
    lib.tadnn_version.restype = C.c_int
    lib.tadnn_launch_count.restype = C.c_longlong
    lib.tadnn_reset_launch_count.restype = None
    lib.tadnn_launch_table.argtypes = [C.c_char_p, C.c_int]
    lib.tadnn_gemm_bf16.argtypes = [C.POINTER(GemmArgs), C.c_void_p]
    # Everything else takes plain ints / pointers; declare lazily through `fn`.


def lib():
    """Load (once) and return the ctypes handle; raises NativeError if it is not built."""
    global _lib
    if _lib is None:
        with _lock:
            if _lib is None:
                if not os.path.exists(LIB_PATH):
                    raise NativeError(
                        f"{LIB_PATH} is missing: run `python build_native.py` "
                        "(or __graft_entry__.build()) first")
                handle = C.CDLL(LIB_PATH, mode=C.RTLD_GLOBAL)
                _declare(handle)
                _lib = handle
    return _lib


def check(rc: int, what: str = "") -> None:
    if rc != 0:
        msg = lib().tadnn_last_error().decode("utf-8", "replace")
        raise NativeError(f"{what or 'tadnn native call'} failed (rc={rc}): {msg}")


def launch_count() -> int:
    return int(lib().tadnn_launch_count())


def reset_launch_count() -> None:
    lib().tadnn_reset_launch_count()


def launch_table() -> dict:
    buf = C.create_string_buffer(1 << 16)
    lib().tadnn_launch_table(buf, len(buf))
    out = {}
    for line in buf.value.decode().splitlines():
        name, _, cnt = line.rpartition(" ")
        if name:
            out[name] = int(cnt)
    return out


def ptr(t) -> int:
    """Device pointer of a torch tensor (0 for None)."""
    return 0 if t is None else int(t.data_ptr())


def stream_ptr(stream=None) -> int:
    """Raw cudaStream_t of ``stream`` (default: the current stream of the current device).  Called once per kernel launch:
    goes straight to the two C accessors instead of building a ``torch.cuda.Stream`` object (5 us each, ~60 us per step)."""
    if stream is not None:
        return int(stream.cuda_stream)
    import torch
    return int(torch._C._cuda_getCurrentRawStream(torch._C._cuda_getDevice()))
