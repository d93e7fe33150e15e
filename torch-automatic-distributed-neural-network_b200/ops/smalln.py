"""Linear layers with a tiny output width (csrc/smalln.cu): forward, dgrad, wgrad for N <= 16.

The classifier layer of the README MLP is ``MPBaseLinear(i, 10)`` (nodemodule.lua:86-119).  A 128-wide
tcgen05 tile is the wrong shape for it, so these go through bandwidth-shaped CUDA-core
kernels: x is read once with 16-byte loads over the whole GPU, W stays in L1.
"""
from __future__ import annotations

import ctypes as C

import torch

from .. import _native

MAX_N = 16


def supported(n: int, k: int) -> bool:
    return 1 <= n <= MAX_N and k >= 8 and k % 8 == 0


# Measured on B200 (profiles/smalln_vs_gemm.json, device time per call, N = 10):
#   M x K          fwd  gemm / smalln   dgrad gemm / smalln   wgrad(+bias) gemm / smalln   [us]
#   4096 x 256          6.7 / 5.1           16.7 / 6.2              21.5 / 16.4
#   4096 x 2048        11.1 / 14.7          29.9 / 18.1             23.1 / 24.8
#   32768 x 1024       16.2 / 49.7         115.2 / 86.8             27.3 / 60.2
# The CUDA-core kernels win while the layer is latency-bound; once M*K reaches a few million elements
# the 10 FMAs per loaded element cost more than the tcgen05 tile's wasted columns, except for dgrad,
# whose GEMM form pads K = N up to a 64-deep MMA block and is write-bound either way.
def prefer_fwd(M: int, N: int, K: int) -> bool:
    return supported(N, K) and M * K <= (1 << 21)


def prefer_dgrad(M: int, N: int, K: int) -> bool:
    return supported(N, K)


def prefer_wgrad(M: int, N: int, K: int) -> bool:
    return supported(N, K) and M * K <= (1 << 22)


def _s(stream=None):
    return C.c_void_p(_native.stream_ptr(stream))


def _p(t):
    return C.c_void_p(0 if t is None else t.data_ptr())


def _rows_ok(t: torch.Tensor) -> bool:
    return t.stride(1) == 1 and t.stride(0) % 8 == 0 and t.data_ptr() % 16 == 0


def linear_smalln_fwd(x: torch.Tensor, w: torch.Tensor, bias: torch.Tensor | None = None,
                      out: torch.Tensor | None = None) -> torch.Tensor:
    """out[M,N] (fp32) = x[M,K] · w[N,K]ᵀ (+ bias)."""
    assert x.is_cuda and x.dtype == torch.bfloat16 and w.dtype == torch.bfloat16 and x.dim() == 2 and w.dim() == 2
    M, K = x.shape
    N = w.shape[0]
    assert w.shape[1] == K and supported(N, K) and _rows_ok(x) and _rows_ok(w)
    if out is None:
        out = torch.empty((M, N), device=x.device, dtype=torch.float32)
    assert out.dtype == torch.float32 and out.shape == (M, N) and out.stride(1) == 1
    if bias is not None:
        assert bias.dtype == torch.float32 and bias.numel() == N and bias.is_contiguous()
    _native.check(_native.lib().tadnn_linear_smalln_fwd(_p(x), x.stride(0), _p(w), w.stride(0), _p(bias), _p(out), out.stride(0),
                                                        M, N, K, _s()), "tadnn_linear_smalln_fwd")
    return out


def linear_smalln_dgrad(g: torch.Tensor, w: torch.Tensor, aux_tanh_out: torch.Tensor | None = None) -> torch.Tensor:
    """dx[M,K] (bf16) = g[M,N] · w[N,K], optionally times ``1 - aux²`` (the tanh that produced the layer input)."""
    assert g.is_cuda and g.dtype == torch.bfloat16 and w.dtype == torch.bfloat16 and g.stride(1) == 1
    M, N = g.shape
    K = w.shape[1]
    assert w.shape[0] == N and supported(N, K) and _rows_ok(w)
    dx = torch.empty((M, K), device=g.device, dtype=torch.bfloat16)
    if aux_tanh_out is not None:
        assert aux_tanh_out.dtype == torch.bfloat16 and aux_tanh_out.shape == (M, K) and _rows_ok(aux_tanh_out)
    _native.check(_native.lib().tadnn_linear_smalln_dgrad(
        _p(g), g.stride(0), _p(w), w.stride(0), _p(aux_tanh_out), 0 if aux_tanh_out is None else aux_tanh_out.stride(0),
        _p(dx), dx.stride(0), M, N, K, _s()), "tadnn_linear_smalln_dgrad")
    return dx


def linear_smalln_wgrad(g: torch.Tensor, x: torch.Tensor, dw: torch.Tensor | None = None, db: torch.Tensor | None = None,
                        accumulate: bool = False, want_bias: bool = True):
    """dW[N,K] (fp32) (+)= gᵀ · x and db[N] (+)= column sums of g, in one pass over x."""
    assert g.is_cuda and g.dtype == torch.bfloat16 and x.dtype == torch.bfloat16 and g.stride(1) == 1
    M, N = g.shape
    K = x.shape[1]
    assert x.shape[0] == M and supported(N, K) and _rows_ok(x)
    if dw is None:
        assert not accumulate
        dw = torch.empty((N, K), device=g.device, dtype=torch.float32)
    if db is None and want_bias:
        assert not accumulate
        db = torch.empty((N,), device=g.device, dtype=torch.float32)
    assert dw.dtype == torch.float32 and dw.shape == (N, K) and dw.stride(1) == 1
    _native.check(_native.lib().tadnn_linear_smalln_wgrad(_p(g), g.stride(0), _p(x), x.stride(0), _p(dw), dw.stride(0), _p(db),
                                                          M, N, K, int(accumulate), _s()), "tadnn_linear_smalln_wgrad")
    return dw, db


__all__ = ["MAX_N", "supported", "prefer_fwd", "prefer_dgrad", "prefer_wgrad", "linear_smalln_fwd", "linear_smalln_dgrad", "linear_smalln_wgrad"]
