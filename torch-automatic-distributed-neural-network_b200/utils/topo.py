"""NVLink / NVSwitch / multicast discovery (SURVEY §5 "distributed communication backend").

The reference runs one MPI rank per node over Ethernet/IB (cifar_example/train.sh:14) and asks
TorchMPI whether collectives cross nodes (``mpi.needInterNodeCollectives``,
datamodule.lua:201).  On a B200 box the questions are: can every GPU reach every peer directly
(NVSwitch), and does the driver expose multicast (NVLS)?"""
from __future__ import annotations

import subprocess
from typing import Dict, List, Optional

import torch


def peer_access_matrix() -> List[List[bool]]:
    n = torch.cuda.device_count()
    return [[i == j or torch.cuda.can_device_access_peer(i, j) for j in range(n)] for i in range(n)]


def all_peers_direct() -> bool:
    m = peer_access_matrix()
    return all(all(row) for row in m)


def nvidia_smi_topo() -> Optional[str]:
    try:
        return subprocess.run(["nvidia-smi", "topo", "-m"], capture_output=True, text=True, timeout=20).stdout
    except (OSError, subprocess.TimeoutExpired):
        return None


def nvlink_link_counts(topo_text: Optional[str] = None) -> Dict[str, int]:
    """{"GPU0-GPU1": 18, ...} parsed from ``nvidia-smi topo -m`` (NV18 = 18 bonded links)."""
    import re
    txt = topo_text if topo_text is not None else (nvidia_smi_topo() or "")
    txt = re.sub(r"\x1b\[[0-9;]*m", "", txt)
    out: Dict[str, int] = {}
    lines = [l for l in txt.splitlines() if l.startswith("GPU")]
    for l in lines:
        cols = l.split()
        src = cols[0]
        for j, c in enumerate(cols[1:]):
            if c.startswith("NV") and c[2:].isdigit():
                out[f"{src}-GPU{j}"] = int(c[2:])
    return out


def multicast_supported(device: int = 0) -> Optional[bool]:
    """CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED via cuda-python if importable, else None."""
    try:
        from cuda import cuda  # type: ignore
        (err,) = cuda.cuInit(0)
        err, dev = cuda.cuDeviceGet(device)
        err, v = cuda.cuDeviceGetAttribute(cuda.CUdevice_attribute.CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, dev)
        return bool(v)
    except Exception:  # noqa: BLE001
        return None


def describe() -> dict:
    n = torch.cuda.device_count() if torch.cuda.is_available() else 0
    d = {"gpus": n}
    if n:
        d["name"] = torch.cuda.get_device_name(0)
        d["all_peers_direct"] = all_peers_direct()
        d["nvlink"] = nvlink_link_counts()
        d["multicast_supported"] = multicast_supported(0)
    return d
