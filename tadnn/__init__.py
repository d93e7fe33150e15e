"""Import alias: ``import tadnn`` loads the package that lives in
``torch-automatic-distributed-neural-network_b200/`` (a directory name Python cannot
import directly).  Sub-modules resolve through ``__path__``, so ``tadnn.nn``,
``tadnn.ops`` … are the files in that directory."""
import os as _os

_real = _os.path.join(_os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))),
                      "torch-automatic-distributed-neural-network_b200")
__path__ = [_real]
__file__ = _os.path.join(_real, "__init__.py")
with open(__file__, "r", encoding="utf-8") as _f:
    exec(compile(_f.read(), __file__, "exec"))
del _f, _os
