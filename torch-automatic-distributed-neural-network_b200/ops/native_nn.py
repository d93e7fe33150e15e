"""Native (sm_100a) execution of the stock layers: ``Linear``, ``SpatialConvolutionMM`` and
``SpatialMaxPooling`` as autograd Functions over the tadnn kernels.

The reference's value is that an ordinary ``nn.Sequential`` of stock layers gets parallelised
(README.md:92-110, cifar_example/sgd-torchad_nn-cifar.lua:88-118, nodemodule.lua:40-54,93-105);
under it every such layer is a cuBLAS SGEMM / cunn kernel.  Here a CUDA input sends the same
layers to the tcgen05 GEMM (csrc/gemm_tcgen05.cu), the im2col / col2im / pooling kernels
(csrc/cnn.cu) and the small-N kernels (csrc/smalln.cu): no cuBLAS and no cuDNN call on the
path.  bf16 operands, fp32 accumulation, fp32 master weights and fp32 gradients.

Dispatch rule (``native_ok``): CUDA tensor, library built, and a shape the kernels cover; anything
else takes the plain PyTorch statement of the layer (CPU, exotic strides, padded/strided
convolutions).  ``TADNN_NATIVE_LAYERS=0`` forces the PyTorch statement everywhere.
"""
from __future__ import annotations

import ctypes as C
import os

import torch
import torch.nn.functional as F

from .. import _native


def enabled() -> bool:
    return os.environ.get("TADNN_NATIVE_LAYERS", "1") not in ("0", "false") and _native.available()


def _pad8(n: int) -> int:
    return (n + 7) // 8 * 8


def to_bf16_rows(t: torch.Tensor) -> torch.Tensor:
    """2-D tensor -> bf16 with last dim contiguous, row pitch a multiple of 8 elements and a 16-byte
    aligned base: a legal TMA operand whatever the width.  Zero-copy when ``t`` already is one."""
    assert t.dim() == 2 and t.is_cuda
    rows, cols = t.shape
    if t.dtype == torch.bfloat16 and t.stride(1) == 1 and t.stride(0) % 8 == 0 and t.data_ptr() % 16 == 0:
        return t
    ld = _pad8(cols)
    buf = torch.empty((rows, ld), device=t.device, dtype=torch.bfloat16)
    out = buf[:, :cols]
    if t.dtype == torch.float32 and t.stride(1) == 1:
        _native.check(_native.lib().tadnn_cast_f32_bf16_2d(
            C.c_void_p(t.data_ptr()), C.c_longlong(t.stride(0)), C.c_void_p(buf.data_ptr()), C.c_longlong(ld),
            C.c_longlong(rows), C.c_int(cols), C.c_void_p(_native.stream_ptr())), "tadnn_cast_f32_bf16_2d")
    else:
        out.copy_(t)
    return out


# ------------------------------------------------------------------------------------------------
# Linear
# ------------------------------------------------------------------------------------------------
class LinearFn(torch.autograd.Function):
    """y = x W^T + b on the tcgen05 GEMM (small-N kernels for classifier-sized outputs).
    Replaces nn.Linear's addmm / addmv / addr (nodemodule.lua:47,65,72,95,116)."""

    @staticmethod
    def forward(ctx, x, weight, bias):
        from . import smalln
        from .gemm import gemm_bf16
        lead = x.shape[:-1]
        x2 = x.reshape(-1, x.shape[-1])
        xb = to_bf16_rows(x2)
        wb = to_bf16_rows(weight.detach())
        M, K = xb.shape
        N = wb.shape[0]
        out_dtype = torch.float32 if x.dtype == torch.float32 else torch.bfloat16
        b32 = None if bias is None else bias.detach().float().contiguous()
        if smalln.prefer_fwd(M, N, K):
            y = smalln.linear_smalln_fwd(xb, wb, b32)
            y = y if out_dtype == torch.float32 else y.to(out_dtype)
        else:
            y = gemm_bf16(xb, wb, bias=b32, out_dtype=out_dtype)
        ctx.save_for_backward(xb, wb)
        ctx.has_bias = bias is not None
        ctx.x_dtype = x.dtype
        ctx.lead = lead
        return y.view(*lead, N)

    @staticmethod
    def backward(ctx, g):
        from . import cnn as K
        from . import smalln
        from .gemm import gemm_bf16
        xb, wb = ctx.saved_tensors
        M, Kd = xb.shape
        N = wb.shape[0]
        gb = to_bf16_rows(g.reshape(-1, N))
        dx = dw = db = None
        if ctx.needs_input_grad[0]:
            odt = torch.float32 if ctx.x_dtype == torch.float32 else torch.bfloat16
            if smalln.prefer_dgrad(M, N, Kd) and Kd % 8 == 0:
                dx = smalln.linear_smalln_dgrad(gb, wb).to(odt)
            else:
                dx = gemm_bf16(gb, wb, b_mn_major=True, out_dtype=odt)
            dx = dx.view(*ctx.lead, Kd)
        if ctx.needs_input_grad[1]:
            if smalln.prefer_wgrad(M, N, Kd) and ctx.has_bias:
                dw, db = smalln.linear_smalln_wgrad(gb, xb)
            else:
                dw = gemm_bf16(gb, xb, a_mn_major=True, b_mn_major=True, out_dtype=torch.float32,
                               split_k=K.auto_split_k(N, Kd, M))
        if ctx.has_bias and ctx.needs_input_grad[2] and db is None:
            db = torch.zeros(N, device=gb.device, dtype=torch.float32)
            K.colsum_bf16(gb, db)
        return dx, dw, db


def linear_ok(x: torch.Tensor, weight: torch.Tensor) -> bool:
    return (enabled() and x.is_cuda and weight.is_cuda and x.dtype in (torch.float32, torch.bfloat16)
            and weight.dtype == torch.float32 and x.dim() >= 1 and x.numel() > 0)


def linear(x: torch.Tensor, weight: torch.Tensor, bias) -> torch.Tensor:
    """Dispatching ``F.linear``: the native path on CUDA, PyTorch elsewhere."""
    if linear_ok(x, weight):
        if x.dim() == 1:
            return LinearFn.apply(x.unsqueeze(0), weight, bias).squeeze(0)
        return LinearFn.apply(x, weight, bias)
    return F.linear(x, weight.to(x.dtype), None if bias is None else bias.to(x.dtype))


# ------------------------------------------------------------------------------------------------
# SpatialConvolutionMM: im2col + GEMM (stride 1, no padding); the output is NHWC in memory and is
# handed on as an NCHW-shaped view (channels_last strides), which the pooling below consumes as is.
# ------------------------------------------------------------------------------------------------
def _nhwc_view(x: torch.Tensor):
    """``x`` ([B,C,H,W] logical) as a contiguous [B,H,W,C] tensor if its memory already is NHWC, else None."""
    p = x.permute(0, 2, 3, 1)
    return p if p.is_contiguous() else None


class Conv2dFn(torch.autograd.Function):
    """Replaces cunn SpatialConvolutionMM (cifar_example/sgd-torchad_nn-cifar.lua:91,96):
    forward im2col + GEMM(+bias); backward wgrad (split-K GEMM over MN-major operands), bias column
    sum, dgrad GEMM + col2im.  Weight in the Torch7 layout [O, C*KH*KW]."""

    @staticmethod
    def forward(ctx, x, weight, bias, KH, KW):
        from . import cnn as K
        B, Cc, H, W = x.shape
        O = weight.shape[0]
        OH, OW = H - KH + 1, W - KW + 1
        Kpad = _pad8(Cc * KH * KW)
        nhwc = _nhwc_view(x) if (x.dtype == torch.bfloat16 and Cc % 8 == 0) else None
        if nhwc is not None:
            col = K.im2col(nhwc, KH, KW, Kpad, nchw=False)
        elif Cc % 8 == 0:
            xb = x.to(torch.bfloat16).permute(0, 2, 3, 1).contiguous()
            col = K.im2col(xb, KH, KW, Kpad, nchw=False)
        else:
            col = K.im2col(x.to(torch.bfloat16).contiguous(), KH, KW, Kpad, nchw=True)
        wp = K.conv_weight_pack(weight.detach().contiguous(), Cc, KH, KW, Kpad)
        y = K.gemm_bf16(col, wp, bias=None if bias is None else bias.detach().float().contiguous(), tile_n=64 if O <= 64 else 0)
        ctx.save_for_backward(col, wp)
        ctx.geom = (B, Cc, H, W, O, KH, KW, OH, OW, Kpad)
        ctx.has_bias = bias is not None
        ctx.x_dtype = x.dtype
        return y.view(B, OH, OW, O).permute(0, 3, 1, 2)   # NCHW-shaped view of NHWC memory

    @staticmethod
    def backward(ctx, g):
        from . import cnn as K
        col, wp = ctx.saved_tensors
        B, Cc, H, W, O, KH, KW, OH, OW, Kpad = ctx.geom
        gn = _nhwc_view(g)
        if gn is None or gn.dtype != torch.bfloat16:
            gn = g.to(torch.bfloat16).permute(0, 2, 3, 1).contiguous()
        gb = gn.reshape(B * OH * OW, O)
        dx = dw = db = None
        if ctx.needs_input_grad[1]:
            dwp = K.gemm_bf16(gb, col, a_mn_major=True, b_mn_major=True, out_dtype=torch.float32,
                              split_k=K.auto_split_k(O, Kpad, B * OH * OW))
            dw = torch.zeros((O, Cc * KH * KW), device=g.device, dtype=torch.float32)
            K.conv_wgrad_unpack(dwp, dw, Cc, KH, KW)
        if ctx.has_bias and ctx.needs_input_grad[2]:
            db = torch.zeros(O, device=g.device, dtype=torch.float32)
            K.colsum_bf16(gb, db)
        if ctx.needs_input_grad[0]:
            dcol = K.gemm_bf16(gb, wp, b_mn_major=True)                   # [B*OH*OW, Kpad]
            if Cc % 8 == 0:
                dx = K.col2im(dcol, B, Cc, H, W, KH, KW).permute(0, 3, 1, 2)
            else:  # narrow inputs (the image itself): fold on the PyTorch side; rarely asked for
                cols = dcol[:, :Cc * KH * KW].float().view(B, OH * OW, KH * KW, Cc).permute(0, 3, 2, 1).reshape(B, Cc * KH * KW, OH * OW)
                dx = F.fold(cols, (H, W), (KH, KW))
            dx = dx.to(ctx.x_dtype)
        return dx, dw, db, None, None


def conv_ok(x, weight, KH, KW, dH, dW, padH, padW) -> bool:
    if not (enabled() and x.is_cuda and x.dim() == 4 and weight.dtype == torch.float32):
        return False
    if (dH, dW, padH, padW) != (1, 1, 0, 0) or x.dtype not in (torch.float32, torch.bfloat16):
        return False
    O = weight.shape[0]
    return O % 8 == 0 and x.shape[2] >= KH and x.shape[3] >= KW


# ------------------------------------------------------------------------------------------------
# SpatialMaxPooling (k x k, stride k, floor mode) on NHWC bf16 with byte arg-max
# ------------------------------------------------------------------------------------------------
class MaxPoolFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, k):
        from . import cnn as K
        xn = _nhwc_view(x)
        y, idx = K.maxpool_fwd(xn, k)
        ctx.save_for_backward(y, idx)
        ctx.geom = (x.shape[2], x.shape[3], k)
        return y.permute(0, 3, 1, 2)

    @staticmethod
    def backward(ctx, g):
        y, idx = ctx.saved_tensors
        H, W, k = ctx.geom
        gn = _nhwc_view(g)
        if gn is None or gn.dtype != torch.bfloat16:
            gn = g.to(torch.bfloat16).permute(0, 2, 3, 1).contiguous()
        B, PH, PW, Cc = y.shape
        dx = torch.empty((B, H, W, Cc), device=y.device, dtype=torch.bfloat16)
        _native.check(_native.lib().tadnn_maxpool_bwd(
            C.c_void_p(gn.data_ptr()), C.c_void_p(y.data_ptr()), C.c_void_p(idx.data_ptr()), C.c_void_p(dx.data_ptr()),
            B, H, W, Cc, k, C.c_void_p(_native.stream_ptr())), "tadnn_maxpool_bwd")
        return dx.permute(0, 3, 1, 2), None


def pool_ok(x, kH, kW, dH, dW, padH, padW) -> bool:
    if not (enabled() and x.is_cuda and x.dim() == 4 and x.dtype == torch.bfloat16):
        return False
    if not (kH == kW == dH == dW and padH == 0 and padW == 0 and 1 <= kH <= 15):
        return False
    return x.shape[1] % 8 == 0 and x.shape[2] >= kH and x.shape[3] >= kW and _nhwc_view(x) is not None
