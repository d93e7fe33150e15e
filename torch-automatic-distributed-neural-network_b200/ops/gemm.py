"""tcgen05 GEMM wrapper (csrc/gemm_tcgen05.cu, ``tadnn_gemm_bf16``).

``gemm_bf16`` computes ``act(alpha * A @ B^T + bias)`` on the 5th-gen tensor cores with
the storage convention of include/tadnn.h.  Replaces the cuBLAS calls behind
``nn.Linear`` in the reference (nodemodule.lua:47,65,72,95,116).
"""
from __future__ import annotations

import ctypes as C

import torch

from .. import _native

_ACT = {None: 0, "none": 0, "tanh": 1, "relu": 2}


def gemm_bf16(a: torch.Tensor, b: torch.Tensor, *, a_mn_major: bool = False,
              b_mn_major: bool = False, out: torch.Tensor | None = None,
              out_dtype: torch.dtype = torch.bfloat16, bias: torch.Tensor | None = None,
              act: str | None = None, alpha: float = 1.0, accumulate: bool = False,
              split_k: int = 0, push=None, push_col_offset: int = 0, push_ldc: int = 0,
              tile_n: int = 0, rs_width: int = 0, aux_tanh_out: torch.Tensor | None = None,
              n_rotate: int = 0, rs_slab_major: bool = False, sig_flags=None, sig_src: int = 0,
              flag_stride: int = 0, wait_flags: int = 0, wait_target: int = 0, wait_cols_per_src: int = 0,
              kblock_mask=None, stream=None) -> torch.Tensor:
    """out[M,N] = act(alpha * A·Bᵀ + bias).

    a: [M,K] (or [K,M] if a_mn_major), b: [N,K] (or [K,N] if b_mn_major); bf16, last dim
    contiguous, row pitch a multiple of 8.  ``push`` is an optional list of raw device
    pointers (peer-mapped buffers) that receive a copy of every output tile at column
    offset ``push_col_offset`` (fused all-gather).  ``sig_flags`` (one raw pointer per target, own rank
    included if it consumes too) makes that push self-announcing and ``wait_flags``/``wait_target``/
    ``wait_cols_per_src`` make this GEMM the consumer of such a gathered A operand (K13, include/tadnn.h).
    ``kblock_mask`` (sequence of ints, one per ``tile_n``-wide column tile): bit kb clear = k-block kb (64 deep) of that
    column tile of B is structurally zero and is skipped.
    """
    assert a.is_cuda and b.is_cuda and a.dtype == torch.bfloat16 and b.dtype == torch.bfloat16
    assert a.dim() == 2 and b.dim() == 2 and a.stride(1) == 1 and b.stride(1) == 1
    M, K = (a.shape[1], a.shape[0]) if a_mn_major else (a.shape[0], a.shape[1])
    N, Kb = (b.shape[1], b.shape[0]) if b_mn_major else (b.shape[0], b.shape[1])
    assert K == Kb, f"inner dims differ: {K} vs {Kb}"
    if rs_width > 0:
        assert out is None and push, "rs_width: partial tiles go to `push` only"
    if out is None and rs_width == 0:
        out = torch.empty((M, N), device=a.device, dtype=out_dtype)
    if out is not None:
        assert out.shape == (M, N) and out.stride(1) == 1
        assert out.dtype in (torch.bfloat16, torch.float32)
    if bias is not None:
        assert bias.dtype == torch.float32 and bias.numel() == N and bias.is_contiguous()
    args = _native.GemmArgs()
    args.M, args.N, args.K = M, N, K
    args.A, args.lda, args.a_mn_major = a.data_ptr(), a.stride(0), int(a_mn_major)
    args.B, args.ldb, args.b_mn_major = b.data_ptr(), b.stride(0), int(b_mn_major)
    if out is not None:
        args.C, args.ldc, args.c_fp32 = out.data_ptr(), out.stride(0), int(out.dtype == torch.float32)
    else:
        args.C, args.ldc, args.c_fp32 = 0, N, 0
    args.bias = _native.ptr(bias)
    args.act = _ACT[act]
    args.alpha = float(alpha)
    args.accumulate = int(accumulate)
    args.split_k = int(split_k)
    push = list(push or [])
    args.n_push = len(push)
    for i, p in enumerate(push):
        args.push[i] = int(p)
    args.push_ldc = int(push_ldc)
    args.push_col_offset = int(push_col_offset)
    args.tile_n = int(tile_n)
    args.rs_width = int(rs_width)
    args.n_rotate = int(n_rotate)
    args.rs_slab_major = int(rs_slab_major)
    sig = list(sig_flags or [])
    args.n_sig = len(sig)
    for i, ptr in enumerate(sig):
        args.sig_flags[i] = int(ptr)
    args.sig_src, args.flag_stride = int(sig_src), int(flag_stride)
    args.wait_flags, args.wait_target, args.wait_cols_per_src = int(wait_flags), int(wait_target), int(wait_cols_per_src)
    if kblock_mask is not None:
        km = (C.c_uint * len(kblock_mask))(*[int(v) for v in kblock_mask])
        args.kblock_mask, args.n_kblock_mask = km, len(kblock_mask)
    if aux_tanh_out is not None:
        assert aux_tanh_out.dtype == torch.bfloat16 and aux_tanh_out.stride(1) == 1 and aux_tanh_out.shape == (M, N)
        args.aux_tanh_out, args.ld_aux = aux_tanh_out.data_ptr(), aux_tanh_out.stride(0)
    rc = _native.lib().tadnn_gemm_bf16(C.byref(args), C.c_void_p(_native.stream_ptr(stream)))
    _native.check(rc, "tadnn_gemm_bf16")
    return out


def gemm_bf16_2cta(a: torch.Tensor, b: torch.Tensor, *, out: torch.Tensor | None = None, bias: torch.Tensor | None = None,
                   act: str | None = None, alpha: float = 1.0, stream=None) -> torch.Tensor:
    """``out[M,N] = act(alpha * a[M,K] @ b[N,K]^T + bias)`` with the 2-CTA kernel (csrc/gemm_2cta.cu:
    clusters of two CTAs, ``tcgen05.mma.cta_group::2``, 256x256 tiles).  K-major bf16 operands, bf16 out."""
    assert a.is_cuda and a.dtype == torch.bfloat16 and b.dtype == torch.bfloat16
    assert a.dim() == 2 and b.dim() == 2 and a.stride(1) == 1 and b.stride(1) == 1 and a.shape[1] == b.shape[1]
    M, K = a.shape
    N = b.shape[0]
    if out is None:
        out = torch.empty((M, N), device=a.device, dtype=torch.bfloat16)
    assert out.dtype == torch.bfloat16 and out.shape == (M, N) and out.stride(1) == 1
    rc = _native.lib().tadnn_gemm_bf16_2cta(
        C.c_void_p(a.data_ptr()), C.c_int(a.stride(0)), C.c_void_p(b.data_ptr()), C.c_int(b.stride(0)),
        C.c_void_p(out.data_ptr()), C.c_int(out.stride(0)), C.c_int(M), C.c_int(N), C.c_int(K),
        C.c_void_p(_native.ptr(bias)), C.c_int(_ACT[act]), C.c_float(alpha), C.c_void_p(_native.stream_ptr(stream)))
    _native.check(rc, "tadnn_gemm_bf16_2cta")
    return out
