"""The example's dataset preprocessing (cifar_example/sgd-torchad_nn-cifar.lua:182-216): RGB -> YUV,
local contrast normalisation of the Y channel with a 7-tap Gaussian neighbourhood
(``nn.SpatialContrastiveNormalization(1, image.gaussian1D(7))``), global standardisation of U and V.

``preprocess`` is pure tensor arithmetic (any device; the CPU path and the test oracle);
``preprocess_cuda`` runs the same transform through the native kernels of csrc/preprocess.cu (two
passes over the shard).  Not on the training hot path.
The reference divides U and V by the *negative* standard deviation (``-std_u``, ``-std_v``, lines
195, 200, 213, 216 — SURVEY Q10); ``reference_sign=True`` reproduces that, the default uses +std.
"""
from __future__ import annotations

import math

import torch
import torch.nn.functional as F


def rgb2yuv(x: torch.Tensor) -> torch.Tensor:
    """``image.rgb2yuv`` on ``[..., 3, H, W]``."""
    r, g, b = x[..., 0, :, :], x[..., 1, :, :], x[..., 2, :, :]
    y = 0.299 * r + 0.587 * g + 0.114 * b
    u = -0.14713 * r - 0.28886 * g + 0.436 * b
    v = 0.615 * r - 0.51499 * g - 0.10001 * b
    return torch.stack([y, u, v], dim=-3)


def gaussian1d(size: int = 7, sigma: float = 0.25, amplitude: float = 1.0) -> torch.Tensor:
    """``image.gaussian1D(size)`` (sigma is relative to the size, Torch ``image`` convention)."""
    center = size / 2.0 + 0.5
    i = torch.arange(1, size + 1, dtype=torch.float64)
    return (amplitude * torch.exp(-(((i - center) / (sigma * size)) ** 2) / 2)).float()


def spatial_contrastive_normalization(y: torch.Tensor, kernel1d: torch.Tensor, threshold: float = 1e-4,
                                      thresval: float | None = None) -> torch.Tensor:
    """``nn.SpatialContrastiveNormalization(1, kernel, threshold, thresval)`` on ``[N, 1, H, W]`` (Torch7 ``nn``
    arithmetic, see :mod:`tadnn.nn.normalization`): subtract the border-corrected local mean, then divide by
    the local standard deviation, element-wise replaced by ``thresval`` where it does not exceed ``threshold``."""
    from ..nn.normalization import _kernel2d, local_estimate
    k2 = _kernel2d(kernel1d).to(y.device)
    thresval = threshold if thresval is None else thresval
    coef = local_estimate(torch.ones_like(y[:1]), k2)  # border correction
    centered = y - local_estimate(y, k2) / coef
    std = torch.sqrt(local_estimate(centered * centered, k2)) / coef
    return centered / torch.where(std > threshold, std, torch.full_like(std, thresval))


def preprocess(train: torch.Tensor, test: torch.Tensor | None = None, reference_sign: bool = False):
    """Returns preprocessed ``(train, test)`` (``test`` may be None).  U/V statistics come from the
    training shard and are applied to the test set, as in the example."""
    out = []
    tr = rgb2yuv(train.float())
    mean_u, std_u = tr[:, 1].mean(), tr[:, 1].std()
    mean_v, std_v = tr[:, 2].mean(), tr[:, 2].std()
    sign = -1.0 if reference_sign else 1.0
    kernel = gaussian1d(7)
    for x in (tr, rgb2yuv(test.float()) if test is not None else None):
        if x is None:
            out.append(None)
            continue
        y = spatial_contrastive_normalization(x[:, 0:1], kernel)
        u = (x[:, 1:2] - mean_u) / (sign * std_u)
        v = (x[:, 2:3] - mean_v) / (sign * std_v)
        out.append(torch.cat([y, u, v], dim=1))
    return out[0], out[1]


def uv_stats_cuda(rgb: torch.Tensor):
    """(mean_u, std_u, mean_v, std_v) of a CUDA fp32 ``[N, 3, H, W]`` shard (unbiased std, as ``torch.std``);
    one pass with fp64 accumulation (csrc/preprocess.cu)."""
    import ctypes as C
    from .. import _native
    assert rgb.is_cuda and rgb.dtype == torch.float32 and rgb.dim() == 4 and rgb.shape[1] == 3 and rgb.is_contiguous()
    n, _, H, W = rgb.shape
    stats = torch.empty(4, device=rgb.device, dtype=torch.float64)
    _native.check(_native.lib().tadnn_preprocess_uv_stats(C.c_void_p(rgb.data_ptr()), C.c_void_p(stats.data_ptr()), C.c_longlong(n), H, W,
                                                          C.c_void_p(_native.stream_ptr())), "tadnn_preprocess_uv_stats")
    su, suu, sv, svv = stats.tolist()
    cnt = n * H * W
    mu, mv = su / cnt, sv / cnt
    var_u = max(suu - cnt * mu * mu, 0.0) / max(cnt - 1, 1)
    var_v = max(svv - cnt * mv * mv, 0.0) / max(cnt - 1, 1)
    return mu, math.sqrt(var_u), mv, math.sqrt(var_v)


def yuv_lcn_cuda(rgb: torch.Tensor, stats, kernel1d: torch.Tensor | None = None, threshold: float = 1e-4,
                 out_dtype: torch.dtype = torch.float32, thresval: float | None = None) -> torch.Tensor:
    """rgb2yuv + U/V standardisation with ``stats`` + contrast normalisation of Y, one CTA per image."""
    import ctypes as C
    from .. import _native
    assert rgb.is_cuda and rgb.dtype == torch.float32 and rgb.dim() == 4 and rgb.shape[1] == 3 and rgb.is_contiguous()
    assert out_dtype in (torch.float32, torch.bfloat16)
    n, _, H, W = rgb.shape
    k = (gaussian1d(7) if kernel1d is None else kernel1d).detach().float().cpu().contiguous()
    taps = (C.c_float * k.numel())(*k.tolist())
    mu, sdu, mv, sdv = stats
    out = torch.empty(rgb.shape, device=rgb.device, dtype=out_dtype)
    _native.check(_native.lib().tadnn_preprocess_yuv_lcn(
        C.c_void_p(rgb.data_ptr()), C.c_void_p(out.data_ptr()), int(out_dtype == torch.bfloat16), C.c_longlong(n), H, W, taps, k.numel(),
        C.c_float(mu), C.c_float(sdu), C.c_float(mv), C.c_float(sdv), C.c_float(threshold),
        C.c_float(threshold if thresval is None else thresval), C.c_void_p(_native.stream_ptr())),
        "tadnn_preprocess_yuv_lcn")
    return out


def preprocess_cuda(train: torch.Tensor, test: torch.Tensor | None = None, reference_sign: bool = False,
                    out_dtype: torch.dtype = torch.float32):
    """:func:`preprocess` on the GPU through the native kernels: two passes over the training shard
    (statistics, then the fused transform) and one over the test set."""
    mu, sdu, mv, sdv = uv_stats_cuda(train)
    sign = -1.0 if reference_sign else 1.0
    stats = (mu, sign * sdu, mv, sign * sdv)
    tr = yuv_lcn_cuda(train, stats, out_dtype=out_dtype)
    te = yuv_lcn_cuda(test, stats, out_dtype=out_dtype) if test is not None else None
    return tr, te


__all__ = ["rgb2yuv", "gaussian1d", "spatial_contrastive_normalization", "preprocess", "uv_stats_cuda", "yuv_lcn_cuda",
           "preprocess_cuda"]
assert math  # noqa: S101 (kept for readers comparing against the Lua formulas)
