"""The MLPs the reference's README and BASELINE.json name.

* ``readme_mlp``     README.md:92-95 "old components": Reshape(1024) -> Linear(1024,2048) -> Tanh -> Linear(2048,10)
* ``readme_mp_mlp``  README.md:98-101: the same network from MPInitialReshape / MPInitialLinear / MPTanh / MPBaseLinear
* ``wide_mp_mlp``    BASELINE.json config 4: 4 x Linear 16384x16384 with MPTanh between, as the fused B200 module
"""
from __future__ import annotations


def readme_mlp():
    from .. import nn
    return (nn.Sequential().add(nn.Reshape(1024)).add(nn.Linear(1024, 2048)).add(nn.Tanh())
            .add(nn.Linear(2048, 10)))


def readme_mp_mlp(mpi=None, compat=None):
    from .. import nn
    return (nn.Sequential().add(nn.MPInitialReshape(1024, mpi=mpi))
            .add(nn.MPInitialLinear(1024, 2048, mpi=mpi, compat=compat)).add(nn.MPTanh(mpi=mpi, compat=compat))
            .add(nn.MPBaseLinear(2048, 10, mpi=mpi, compat=compat)))


def wide_mp_mlp(width: int = 16384, layers: int = 4, mpi=None):
    from ..parallel.tp import FusedMPMLP
    return FusedMPMLP([width] * (layers + 1), mpi=mpi)
