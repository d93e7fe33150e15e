"""``tadnn.nn`` — the Torch7-``nn`` vocabulary the reference is written against, on PyTorch.

The reference builds ``nn.Sequential`` models out of stock Torch7 layers plus its five
model-parallel classes (README.md:92-101; cifar_example/sgd-torchad_nn-cifar.lua:88-118) and
drives them either with ``model:forward`` / ``model:backward`` or with
``nn.StochasticGradient`` (datamodule.lua:83-184).  This module supplies that vocabulary as
``torch.nn.Module`` subclasses, so a model runs under ordinary autograd *and* answers the
Torch7 protocol (``forward/backward/updateOutput/updateGradInput/accGradParameters/
accUpdateGradParameters/zeroGradParameters/updateParameters/getParameters``).

Conventions that differ from Lua and cannot be papered over: indices are 0-based (class
targets are ``0..C-1``; ``dataset[i]`` starts at 0).
"""
from __future__ import annotations

from .protocol import T7Protocol  # noqa: F401
from .layers import (Module, Sequential, Linear, Tanh, ReLU, LogSoftMax, Dropout, Reshape, View,  # noqa: F401
                     SpatialConvolutionMM, SpatialMaxPooling)
from .normalization import (SpatialContrastiveNormalization, SpatialDivisiveNormalization,  # noqa: F401
                            SpatialSubtractiveNormalization)
from .criterion import Criterion, ClassNLLCriterion, CrossEntropyCriterion, MSECriterion  # noqa: F401
from .mp import (MPBaseLinear, MPBaseReshape, MPInitialLinear, MPInitialReshape, MPTanh,  # noqa: F401
                 narrowInput, fuse_mp)
from .trainer import StochasticGradient  # noqa: F401

__all__ = [
    "Module", "Sequential", "Linear", "Tanh", "ReLU", "LogSoftMax", "Dropout", "Reshape", "View",
    "SpatialConvolutionMM", "SpatialMaxPooling", "SpatialSubtractiveNormalization", "SpatialDivisiveNormalization",
    "SpatialContrastiveNormalization", "Criterion", "ClassNLLCriterion",
    "CrossEntropyCriterion", "MSECriterion", "StochasticGradient", "MPInitialLinear",
    "MPBaseLinear", "MPTanh", "MPInitialReshape", "MPBaseReshape", "narrowInput", "fuse_mp",
    "T7Protocol",
]
