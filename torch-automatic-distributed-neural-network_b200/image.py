"""The two functions of Torch's ``image`` package that the reference example calls
(cifar_example/sgd-torchad_nn-cifar.lua:181,185): ``image.rgb2yuv`` and ``image.gaussian1D``."""
from .utils.preprocess import gaussian1d, rgb2yuv


def gaussian1D(size: int = 3, sigma: float = 0.25, amplitude: float = 1.0):  # noqa: N802 - Torch's spelling
    return gaussian1d(size, sigma, amplitude)


__all__ = ["rgb2yuv", "gaussian1D", "gaussian1d"]
