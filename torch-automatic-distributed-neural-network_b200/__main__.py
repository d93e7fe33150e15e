"""``python -m tadnn``: one JSON document describing this installation — package and native-library
versions, the exported C entry points, the GPUs and their NVLink/multicast topology, and the
environment switches in effect.  (The reference has no such aid: its only diagnostics are ``print``s,
SURVEY §5.)"""
from __future__ import annotations

import json
import os
import sys


def describe() -> dict:
    import torch

    import tadnn
    from tadnn import _native
    from tadnn.utils import config, topo
    info = {
        "tadnn": tadnn.__version__,
        "torch": torch.__version__,
        "cuda_runtime": torch.version.cuda,
        "native": {"built": _native.available(), "path": _native.LIB_PATH},
        "env": {k: os.environ[k] for k in sorted(config.ENV_SWITCHES) if k in os.environ},
        "env_switches": config.ENV_SWITCHES,
    }
    if _native.available():
        lib = _native.lib()
        info["native"]["version"] = int(lib.tadnn_version())
        try:
            import re
            hdr = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(_native.__file__))), "include", "tadnn.h")
            if os.path.exists(hdr):
                text = re.sub(r"/\*.*?\*/", "", open(hdr).read(), flags=re.S)
                names = sorted(set(re.findall(r"\b(tadnn_[a-z0-9_]+)\s*\(", text)))
                info["native"]["entry_points"] = len(names)
                info["native"]["missing_entry_points"] = [n for n in names if not hasattr(lib, n)]
        except OSError:
            pass
    info["devices"] = topo.describe()
    if torch.cuda.is_available():
        p = torch.cuda.get_device_properties(0)
        info["devices"].update({"sm_count": p.multi_processor_count, "hbm_gb": round(p.total_memory / 2**30, 1),
                                "compute_capability": f"{p.major}.{p.minor}"})
    return info


def main() -> int:
    json.dump(describe(), sys.stdout, indent=1)
    sys.stdout.write("\n")
    return 0


if __name__ == "__main__":
    sys.exit(main())
