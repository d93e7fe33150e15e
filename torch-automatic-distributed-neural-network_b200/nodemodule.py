"""``tadnn.nodemodule`` — the model-parallel layer module (reference: nodemodule.lua).

In the reference, loading this file registers the five ``nn.MP*`` classes as a side effect
(``torch.class('nn.MPInitialLinear', 'nn.Linear')`` …, nodemodule.lua:34,86,132,186,238) and
returns an empty table.  Here the classes live in ``tadnn.nn.mp``; this module re-exports them
and the three debug switches (nodemodule.lua:3-5) as environment-driven functions.
"""
from __future__ import annotations

import os

from .nn.mp import (MPBaseLinear, MPBaseReshape, MPInitialLinear, MPInitialReshape, MPTanh,  # noqa: F401
                    fuse_mp, narrowInput)


def _flag(name):
    return os.environ.get(name, "0").lower() in ("1", "true", "yes", "on")


def printDims() -> bool:      # nodemodule.lua:3
    return _flag("TADNN_PRINT_DIMS")


def syncTanh() -> bool:       # nodemodule.lua:4 — exact mode (gather) is the default here
    return not _flag("TADNN_COMPAT_BACKWARD")


def syncReshape() -> bool:    # nodemodule.lua:5
    return not _flag("TADNN_COMPAT_BACKWARD")


def model_parallelize(model, mpi=None, fuse: bool = False):
    """The model-parallel counterpart of ``parallelize``: rewrite a dense ``Sequential`` of
    ``Reshape / Linear / Tanh`` into the ``MP*`` layers, each rank keeping its own slice of every weight
    (the README shows this rewrite done by hand, README.md:77-110).  ``fuse=True`` additionally returns
    the fused B200 module (``FusedMPMLP``) sharing those parameters when the model is on a GPU with the
    symmetric-memory context up.  The dense model must be identical on every rank (same seed)."""
    from .utils.convert import mp_model_from_dense
    out = mp_model_from_dense(model, mpi)
    p = next(model.parameters(), None)
    if p is not None:
        out = out.to(p.device)
    return fuse_mp(out, mpi=mpi) if fuse else out


CLASSES = {
    "nn.MPInitialLinear": MPInitialLinear,
    "nn.MPBaseLinear": MPBaseLinear,
    "nn.MPTanh": MPTanh,
    "nn.MPInitialReshape": MPInitialReshape,
    "nn.MPBaseReshape": MPBaseReshape,
}
