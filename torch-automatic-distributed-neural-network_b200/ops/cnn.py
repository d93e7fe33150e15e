"""Python wrappers over csrc/cnn.cu (im2col/col2im, max-pool, fused classifier head, SGD).
See include/tadnn.h for the contracts and the reference call sites each kernel replaces."""
from __future__ import annotations

import ctypes as C

import torch

from .. import _native
from .gemm import gemm_bf16


class HeadArgs(C.Structure):
    _fields_ = [
        ("p2", C.c_void_p), ("w1", C.c_void_p), ("b1", C.c_void_p), ("w2", C.c_void_p), ("b2", C.c_void_p),
        ("target", C.c_void_p), ("B", C.c_int), ("D", C.c_int), ("H", C.c_int), ("NC", C.c_int),
        ("dropout_p", C.c_float), ("seed", C.c_uint), ("hidden", C.c_void_p), ("logp", C.c_void_p),
        ("loss", C.c_void_p), ("gw1", C.c_void_p), ("gb1", C.c_void_p), ("gw2", C.c_void_p), ("gb2", C.c_void_p),
        ("dp2", C.c_void_p), ("grad_scale", C.c_float), ("grad_scale_dev", C.c_void_p),
        ("pool_idx", C.c_void_p), ("dpre", C.c_void_p), ("gbias_pre", C.c_void_p), ("seed_dev", C.c_void_p),
    ]


class CnnStepArgs(C.Structure):
    """Mirror of ``tadnn_cnn_step_args`` (include/tadnn.h)."""
    _fields_ = ([("xq", C.c_void_p), ("target", C.c_void_p), ("B", C.c_int)]
                + [(n, C.c_void_p) for n in ("w1", "b1", "w2", "b2", "fw1", "fb1", "fw2", "fb2",
                                             "gw1", "gb1", "gw2", "gb2", "gfw1", "gfb1", "gfw2", "gfb2",
                                             "w1x", "w2p", "Tt", "dwp", "p1", "idx1", "p2", "idx2",
                                             "hidden", "logp", "loss", "da2w", "dp1", "head_ws")]
                + [("dropout_p", C.c_float), ("seed", C.c_uint), ("seed_dev", C.c_void_p), ("grad_scale", C.c_float),
                   ("tt_mask", C.POINTER(C.c_uint)), ("n_tt_mask", C.c_int), ("phase", C.c_int), ("forward_only", C.c_int)])


def cnn_step(args: "CnnStepArgs") -> None:
    """Forward + backward of the fused CNN step as one native call (csrc/cnn_step.cu)."""
    _native.check(_native.lib().tadnn_cnn_step(C.byref(args), _s()), "tadnn_cnn_step")


_F4 = C.c_float * 4
C_ll = C.c_longlong


def _s(stream=None):
    return C.c_void_p(_native.stream_ptr(stream))


def _p(t):
    return C.c_void_p(_native.ptr(t))


def im2col(x: torch.Tensor, KH: int, KW: int, Kpad: int, nchw: bool, out=None) -> torch.Tensor:
    if nchw:
        B, Cc, H, W = x.shape
    else:
        B, H, W, Cc = x.shape
    assert x.dtype == torch.bfloat16 and x.is_contiguous()
    OH, OW = H - KH + 1, W - KW + 1
    if out is None:
        out = torch.empty((B * OH * OW, Kpad), device=x.device, dtype=torch.bfloat16)
    _native.check(_native.lib().tadnn_im2col(_p(x), _p(out), B, Cc, H, W, KH, KW, Kpad, int(nchw), _s()), "tadnn_im2col")
    return out


def im2col_winfirst(x: torch.Tensor, KH: int, KW: int, win: int, only_window: bool) -> torch.Tensor:
    """NHWC im2col with the pooled window's pixels first: rows [0, B*win*win) are (b, oy<win, ox<win)."""
    B, H, W, Cc = x.shape
    OH, OW = H - KH + 1, W - KW + 1
    rows = B * (win * win if only_window else OH * OW)
    out = torch.empty((rows, Cc * KH * KW), device=x.device, dtype=torch.bfloat16)
    _native.check(_native.lib().tadnn_im2col_winfirst(_p(x), _p(out), B, Cc, H, W, KH, KW, win, int(only_window), _s()),
                  "tadnn_im2col_winfirst")
    return out


def conv_toeplitz_pack(wp: torch.Tensor, O: int, Cc: int, KH: int, KW: int, win: int) -> torch.Tensor:
    FH, FW = win + KH - 1, win + KW - 1
    T = torch.empty((FH * FW * Cc, win * win * O), device=wp.device, dtype=torch.bfloat16)
    _native.check(_native.lib().tadnn_conv_toeplitz_pack(_p(wp), _p(T), O, Cc, KH, KW, win, _s()), "tadnn_conv_toeplitz_pack")
    return T


def expand_corner(src: torch.Tensor, H: int, W: int) -> torch.Tensor:
    B, FH, FW, Cc = src.shape
    dst = torch.empty((B, H, W, Cc), device=src.device, dtype=torch.bfloat16)
    _native.check(_native.lib().tadnn_expand_corner(_p(src), _p(dst), B, H, W, FH, FW, Cc, _s()), "tadnn_expand_corner")
    return dst


def col2im(dcol: torch.Tensor, B: int, Cc: int, H: int, W: int, KH: int, KW: int) -> torch.Tensor:
    dx = torch.empty((B, H, W, Cc), device=dcol.device, dtype=torch.bfloat16)
    _native.check(_native.lib().tadnn_col2im(_p(dcol), _p(dx), B, Cc, H, W, KH, KW, _s()), "tadnn_col2im")
    return dx


def maxpool_fwd(x: torch.Tensor, k: int):
    B, H, W, Cc = x.shape
    y = torch.empty((B, H // k, W // k, Cc), device=x.device, dtype=torch.bfloat16)
    idx = torch.empty((B, H // k, W // k, Cc), device=x.device, dtype=torch.uint8)
    _native.check(_native.lib().tadnn_maxpool_fwd(_p(x), _p(y), _p(idx), B, H, W, Cc, k, _s()), "tadnn_maxpool_fwd")
    return y, idx


def maxpool_relu_bwd(dy, y, idx, H: int, W: int, k: int, out=None):
    B, PH, PW, Cc = y.shape
    dx = out if out is not None else torch.empty((B, H, W, Cc), device=y.device, dtype=torch.bfloat16)
    _native.check(_native.lib().tadnn_maxpool_relu_bwd(_p(dy), _p(y), _p(idx), _p(dx), B, H, W, Cc, k, _s()),
                  "tadnn_maxpool_relu_bwd")
    return dx


def colsum_bf16(x: torch.Tensor, out: torch.Tensor):
    R, Cc = x.shape
    _native.check(_native.lib().tadnn_colsum_bf16(_p(x), _p(out), C.c_longlong(R), Cc, x.stride(0), _s()), "tadnn_colsum_bf16")
    return out


def colsum_relu_mask(dy: torch.Tensor, y: torch.Tensor, out: torch.Tensor):
    Cc = y.shape[-1]
    R = y.numel() // Cc
    assert dy.is_contiguous() and y.is_contiguous() and dy.numel() == y.numel()
    _native.check(_native.lib().tadnn_colsum_relu_mask(_p(dy), _p(y), _p(out), C.c_longlong(R), Cc, _s()),
                  "tadnn_colsum_relu_mask")
    return out


def conv_weight_pack(w: torch.Tensor, Cc: int, KH: int, KW: int, Kpad: int) -> torch.Tensor:
    O = w.shape[0]
    wp = torch.empty((O, Kpad), device=w.device, dtype=torch.bfloat16)
    _native.check(_native.lib().tadnn_conv_weight_pack(_p(w), _p(wp), O, Cc, KH, KW, Kpad, _s()), "tadnn_conv_weight_pack")
    return wp


def conv_wgrad_unpack(dwp: torch.Tensor, grad: torch.Tensor, Cc: int, KH: int, KW: int, scale: float = 1.0):
    O, Kpad = dwp.shape
    _native.check(_native.lib().tadnn_conv_wgrad_unpack(_p(dwp), _p(grad), O, Cc, KH, KW, Kpad, C.c_float(scale), _s()),
                  "tadnn_conv_wgrad_unpack")
    return grad


def conv1_weight_pack(w: torch.Tensor) -> torch.Tensor:
    wp = torch.empty((64, 80), device=w.device, dtype=torch.bfloat16)
    _native.check(_native.lib().tadnn_conv1_weight_pack(_p(w), _p(wp), _s()), "tadnn_conv1_weight_pack")
    return wp


def conv1_pool_forward(x: torch.Tensor, wp: torch.Tensor, bias: torch.Tensor):
    """x [B,3,32,32] bf16 -> (p1 [B,9,9,64] bf16, idx1 [B,9,9,64] uint8): conv5x5 + bias + ReLU + maxpool3/3."""
    B = x.shape[0]
    assert x.dtype == torch.bfloat16 and x.is_contiguous() and tuple(x.shape[1:]) == (3, 32, 32)
    p1 = torch.empty((B, 9, 9, 64), device=x.device, dtype=torch.bfloat16)
    idx1 = torch.empty((B, 9, 9, 64), device=x.device, dtype=torch.uint8)
    _native.check(_native.lib().tadnn_conv1_pool_forward(_p(x), _p(wp), _p(bias), _p(p1), _p(idx1), B, _s()),
                  "tadnn_conv1_pool_forward")
    return p1, idx1


def conv1_pool_backward(x, dp1, p1, idx1, grad_weight):
    """dp1: [B, h, w, 64] with h, w <= 9 — the (possibly compact, top-left) pooled gradient."""
    _native.check(_native.lib().tadnn_conv1_pool_backward(_p(x), _p(dp1), int(dp1.shape[1]), int(dp1.shape[2]), _p(p1),
                                                          _p(idx1), _p(grad_weight), x.shape[0], _s()),
                  "tadnn_conv1_pool_backward")
    return grad_weight


def conv2_pool_forward(p1: torch.Tensor, w2p: torch.Tensor, bias: torch.Tensor, store_a2: bool = False):
    """p1 [B,9,9,64] bf16 -> (p2 [B,64] bf16, idx2 [B,64] uint8[, a2 [B,5,5,64]]): conv5x5 + bias + ReLU + maxpool3/3
    as one implicit-GEMM kernel fed by 4-D TMA boxes of p1 (csrc/conv2_fused.cu)."""
    B = p1.shape[0]
    assert p1.dtype == torch.bfloat16 and p1.is_contiguous() and tuple(p1.shape[1:]) == (9, 9, 64)
    p2 = torch.empty((B, 64), device=p1.device, dtype=torch.bfloat16)
    idx2 = torch.empty((B, 64), device=p1.device, dtype=torch.uint8)
    a2 = torch.empty((B, 5, 5, 64), device=p1.device, dtype=torch.bfloat16) if store_a2 else None
    _native.check(_native.lib().tadnn_conv2_pool_forward(_p(p1), _p(w2p), _p(bias), _p(p2), _p(idx2), _p(a2), B, _s()),
                  "tadnn_conv2_pool_forward")
    return (p2, idx2, a2) if store_a2 else (p2, idx2)


def conv2_wgrad(p1: torch.Tensor, da2w: torch.Tensor, dw_packed: torch.Tensor):
    """dw_packed [64,1600] fp32 (k = tap*64 + c) += sum_{n, 3x3 window} da2w[n*9+pos, o] * p1[n, py+kh, px+kw, c]."""
    B = p1.shape[0]
    assert da2w.dtype == torch.bfloat16 and da2w.is_contiguous() and da2w.numel() == B * 9 * 64
    assert dw_packed.dtype == torch.float32 and dw_packed.is_contiguous() and dw_packed.numel() == 64 * 1600
    _native.check(_native.lib().tadnn_conv2_wgrad(_p(p1), _p(da2w), _p(dw_packed), B, _s()), "tadnn_conv2_wgrad")
    return dw_packed


def conv2_wgrad_unpack(dw_packed: torch.Tensor, grad: torch.Tensor):
    """grad[64,1600] (Torch7 order) += dw_packed; dw_packed = 0."""
    _native.check(_native.lib().tadnn_conv2_wgrad_unpack(_p(dw_packed), _p(grad), _s()), "tadnn_conv2_wgrad_unpack")
    return grad


def cnn_pack_shadows(w1, w2, w1p, w2p, Tt):
    _native.check(_native.lib().tadnn_cnn_pack_shadows(_p(w1), _p(w2), _p(w1p), _p(w2p), _p(Tt), _s()), "tadnn_cnn_pack_shadows")


def cnn_pack_shadows_sliding(w1, b1, w2, w1x, w2p, Tt):
    """One launch: conv1 operand of the sliding-operand kernels (bf16 [128,128], bias folded in), packed conv2 filters and
    the Toeplitz matrix of the transposed conv2."""
    _native.check(_native.lib().tadnn_cnn_pack_shadows_sliding(_p(w1), _p(b1), _p(w2), _p(w1x), _p(w2p), _p(Tt), _s()),
                  "tadnn_cnn_pack_shadows_sliding")


def conv1_pool_backward_bias(x, dp1, p1, idx1, grad_weight, grad_bias):
    _native.check(_native.lib().tadnn_conv1_pool_backward_bias(_p(x), _p(dp1), int(dp1.shape[1]), int(dp1.shape[2]), _p(p1), _p(idx1),
                                                               _p(grad_weight), _p(grad_bias), x.shape[0], _s()),
                  "tadnn_conv1_pool_backward_bias")
    return grad_weight


def colsum_relu_mask_corner(dy: torch.Tensor, y: torch.Tensor, out: torch.Tensor):
    """out[c] += sum dy * (y > 0) where dy is [B, FH, FW, C], the top-left corner of y [B, H, W, C]."""
    B, FH, FW, Cc = dy.shape
    assert dy.is_contiguous() and y.is_contiguous() and y.shape[0] == B and y.shape[3] == Cc
    _native.check(_native.lib().tadnn_colsum_relu_mask_corner(_p(dy), _p(y), _p(out), B, int(y.shape[1]), int(y.shape[2]),
                                                              FH, FW, Cc, _s()), "tadnn_colsum_relu_mask_corner")
    return out


_head_ws = {}


def head_workspace(device) -> torch.Tensor:
    """Zero-initialised scratch for tadnn_head_forward, one per device, allocated once."""
    key = str(device)
    ws = _head_ws.get(key)
    if ws is None:
        n = int(_native.lib().tadnn_head_workspace_bytes(0))
        ws = torch.zeros(n, device=device, dtype=torch.uint8)
        _head_ws[key] = ws
    return ws


def sgd_step(p: torch.Tensor, g: torch.Tensor, lr: float, weight_decay: float = 0.0, zero_grad: bool = False):
    assert p.dtype == torch.float32 and g.dtype == torch.float32 and p.is_contiguous() and g.is_contiguous()
    _native.check(_native.lib().tadnn_sgd_step(_p(p), _p(g), C.c_longlong(p.numel()), C.c_float(lr),
                                               C.c_float(weight_decay), int(zero_grad), _s()), "tadnn_sgd_step")


def u8_to_bf16(x: torch.Tensor, plane: int, C: int, scale, shift, out=None) -> torch.Tensor:
    """uint8 [n, C, ...] -> bf16 with y = x * scale[c] + shift[c] (csrc/preprocess.cu)."""
    assert x.dtype == torch.uint8 and x.is_contiguous() and 1 <= C <= 4
    if out is None:
        out = torch.empty(x.shape, device=x.device, dtype=torch.bfloat16)
    fs = (_F4)(*([float(v) for v in scale][:C] + [1.0] * (4 - C)))
    fh = (_F4)(*([float(v) for v in shift][:C] + [0.0] * (4 - C)))
    _native.check(_native.lib().tadnn_u8_to_bf16(_p(x), _p(out), C_ll(x.numel()), int(plane), int(C), fs, fh, _s()), "tadnn_u8_to_bf16")
    return out


def nhwc4_pack(x: torch.Tensor, scale=None, shift=None, out=None) -> torch.Tensor:
    """NCHW [B,3,32,32] uint8 / bf16 / fp32 -> NHWC4 bf16 [B,32,32,4] (channel 3 == 1.0, the bias carrier of
    conv1s_pool_forward); y = x * scale[c] + shift[c] (csrc/conv1_sliding.cu)."""
    assert x.is_contiguous() and tuple(x.shape[1:]) == (3, 32, 32), tuple(x.shape)
    code = {torch.uint8: 0, torch.bfloat16: 1, torch.float32: 2}[x.dtype]
    B = x.shape[0]
    if out is None:
        out = torch.empty((B, 32, 32, 4), device=x.device, dtype=torch.bfloat16)
    fs = (_F4)(*([float(v) for v in scale][:3] + [1.0])) if scale is not None else None
    fh = (_F4)(*([float(v) for v in shift][:3] + [0.0])) if shift is not None else None
    _native.check(_native.lib().tadnn_nhwc4_pack(_p(x), code, _p(out), C_ll(B), fs, fh, _s()), "tadnn_nhwc4_pack")
    return out


def conv1s_weight_pack(w: torch.Tensor, bias: torch.Tensor, out=None) -> torch.Tensor:
    """Torch7 conv1 weight [64,75] fp32 + bias [64] -> bf16 [128,128] operand of conv1s_pool_forward."""
    if out is None:
        out = torch.empty((128, 128), device=w.device, dtype=torch.bfloat16)
    _native.check(_native.lib().tadnn_conv1s_weight_pack(_p(w), _p(bias), _p(out), _s()), "tadnn_conv1s_weight_pack")
    return out


def conv1s_pool_forward(xq: torch.Tensor, wx: torch.Tensor):
    """xq NHWC4 [B,32,32,4] bf16 -> (p1 [B,9,9,64] bf16, idx1 [B,9,9,64] uint8 = 8 - arg-max window position; 15 = dead window)."""
    B = xq.shape[0]
    assert xq.dtype == torch.bfloat16 and xq.is_contiguous() and tuple(xq.shape[1:]) == (32, 32, 4)
    p1 = torch.empty((B, 9, 9, 64), device=xq.device, dtype=torch.bfloat16)
    idx1 = torch.empty((B, 9, 9, 64), device=xq.device, dtype=torch.uint8)
    _native.check(_native.lib().tadnn_conv1s_pool_forward(_p(xq), _p(wx), _p(p1), _p(idx1), B, _s()), "tadnn_conv1s_pool_forward")
    return p1, idx1


def conv1s_pool_backward(xq, dp1, idx1, grad_weight, grad_bias=None, rows: int = 9):
    """grad_weight [64,75] (+ grad_bias [64]) += conv1 gradients from dp1 [B,9,9,64] bf16 (pooled rows < ``rows`` are read)
    and the idx1 of conv1s_pool_forward."""
    assert dp1.dtype == torch.bfloat16 and dp1.is_contiguous() and tuple(dp1.shape[1:]) == (9, 9, 64)
    assert idx1.dtype == torch.uint8 and idx1.is_contiguous() and xq.is_contiguous()
    _native.check(_native.lib().tadnn_conv1s_pool_backward(_p(xq), _p(dp1), _p(idx1), _p(grad_weight),
                                                           _p(grad_bias) if grad_bias is not None else None, int(rows),
                                                           xq.shape[0], _s()), "tadnn_conv1s_pool_backward")
    return grad_weight


def auto_split_k(M: int, N: int, K: int, tile_n: int = 256) -> int:
    tiles = ((M + 127) // 128) * ((N + tile_n - 1) // tile_n)
    kb = (K + 63) // 64
    return max(1, min(kb, 148 // max(1, tiles)))


__all__ = ["CnnStepArgs", "cnn_step", "nhwc4_pack", "conv1s_weight_pack", "conv1s_pool_forward", "conv1s_pool_backward", "cnn_pack_shadows_sliding", "u8_to_bf16", "conv2_pool_forward", "conv2_wgrad", "conv2_wgrad_unpack", "cnn_pack_shadows", "conv1_pool_backward_bias", "im2col_winfirst", "conv_toeplitz_pack", "expand_corner", "conv1_weight_pack", "conv1_pool_forward", "conv1_pool_backward", "im2col", "col2im", "maxpool_fwd", "maxpool_relu_bwd", "colsum_bf16", "colsum_relu_mask", "colsum_relu_mask_corner", "conv_weight_pack",
           "conv_wgrad_unpack", "head_workspace", "sgd_step", "auto_split_k", "HeadArgs", "gemm_bf16"]
