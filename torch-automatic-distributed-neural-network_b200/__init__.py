"""tadnn — B200-native automatic distributed neural-network training.

A from-scratch re-design of TorchAD-NN (ngrabaskas/Torch-Automatic-Distributed-Neural-Network)
for one NVSwitch-connected box of B200 GPUs: PyTorch for the front end, hand-written sm_100a
CUDA (tcgen05/TMEM/TMA GEMMs, peer-memory collectives) underneath.

API map (reference → here)::

    automation = require 'torchad_nn.datamodule'      import tadnn as automation   (or tadnn.datamodule)
    automation.parallelize(data, labels, model,        automation.parallelize(data, labels, model,
        size, mpi, mpinn, batchSize)                       size, mpi, mpinn, batchSize)
    automation.synchronizeModel(model, mpi)            automation.synchronizeModel(model, mpi)
    require 'torchad_nn'  (registers nn.MP* classes)   import tadnn  →  tadnn.nn.MPInitialLinear, …
    mpi = require('torchmpi'); mpi.start(true)         import tadnn.mpi as mpi; mpi.start(True)
    nn.StochasticGradient(model, crit):train(ds)       tadnn.nn.StochasticGradient(model, crit).train(ds)

(reference: init.lua:5-6 returns ``{datamodule, nodemodule}``; README.md:33-34,41,56,74,81.)
"""
from __future__ import annotations

__version__ = "0.1.0"

from . import mpi, mpinn  # noqa: F401
from . import datamodule, nodemodule  # noqa: F401
from . import nn  # noqa: F401
from . import image, optim  # noqa: F401  (the two helper packages the example script touches)
from .datamodule import (comm_speed, data_parallel, optimize_sync, parallelize, remove_hooks,  # noqa: F401
                         selectCollective, synchronizeModel, synchronize_model)
from .trainer import BatchTrainer  # noqa: F401
from .nodemodule import model_parallelize  # noqa: F401
from .nn import (MPBaseLinear, MPBaseReshape, MPInitialLinear, MPInitialReshape, MPTanh,  # noqa: F401
                 StochasticGradient)


def native_available() -> bool:
    """True if libtadnn.so (the sm_100a core) is built."""
    from . import _native
    return _native.available()


__all__ = [
    "parallelize", "synchronizeModel", "synchronize_model", "data_parallel", "optimize_sync",
    "selectCollective", "comm_speed", "remove_hooks", "mpi", "mpinn", "nn", "image", "optim", "datamodule", "nodemodule",
    "MPInitialLinear", "MPBaseLinear", "MPTanh", "MPInitialReshape", "MPBaseReshape",
    "StochasticGradient", "BatchTrainer", "native_available", "model_parallelize",
]
