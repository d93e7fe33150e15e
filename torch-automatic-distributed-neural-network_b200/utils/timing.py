"""Timers.  The reference times with ``sys.tic()/sys.toc()`` wall clock
(cifar_example/sgd-torchad_nn-cifar.lua:45,252,259-263) and a dead ``torch.Timer`` probe
(datamodule.lua:285-298).  On B200 every reported number is device-timed with CUDA events on
the launching stream and reduced with MAX over ranks."""
from __future__ import annotations

import subprocess
import threading
import time
from typing import List, Optional

import torch


class Tic:
    """``sys.tic()`` / ``sys.toc()``."""

    def __init__(self):
        self.t0 = time.perf_counter()

    def tic(self):
        self.t0 = time.perf_counter()

    def toc(self) -> float:
        return time.perf_counter() - self.t0


class CudaTimer:
    """``with CudaTimer() as t: ...`` then ``t.ms`` (synchronises on exit)."""

    def __init__(self, stream=None):
        self.stream = stream
        self.e0 = torch.cuda.Event(enable_timing=True)
        self.e1 = torch.cuda.Event(enable_timing=True)
        self.ms = 0.0

    def __enter__(self):
        self.e0.record(self.stream)
        return self

    def __exit__(self, *exc):
        self.e1.record(self.stream)
        self.e1.synchronize()
        self.ms = self.e0.elapsed_time(self.e1)
        return False


def max_over_ranks(value: float) -> float:
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dev = "cuda" if torch.cuda.is_available() and dist.get_backend() != "gloo" else "cpu"
        t = torch.tensor([value], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())
    return float(value)


class ClockSampler:
    """Samples SM clocks / throttle reasons DURING a timed region (the profiling recipe's clocks
    line).  ``summary()`` gives the bench JSON's ``clocks``.

    Primary source: NVML in-process (``pynvml``), polled every ``period_ms`` (default 2 ms) from a
    background thread — a 20-step timed region of the flagship bench lasts ~12 ms, far below the
    granularity of ``nvidia-smi -lms``.  ``start()`` and ``stop()`` each take one synchronous
    sample, so a summary never comes back empty.  Fallback when NVML cannot be loaded: the same
    fields through one ``nvidia-smi`` call per sample."""

    QUERY = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
             "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
    _REASONS = (("hw_slowdown", 0x8), ("hw_thermal_slowdown", 0x40), ("sw_thermal_slowdown", 0x20),
                ("sw_power_cap", 0x4))

    def __init__(self, gpu_index: int = 0, period_ms: float = 2.0):
        self.gpu = gpu_index
        self.period_ms = period_ms
        self.rows: List[tuple] = []      # (sm_mhz, sm_max_mhz, reason bitmask)
        self._thr: Optional[threading.Thread] = None
        self._stop = threading.Event()
        self._nvml = None
        self._h = None
        self._max = None
        self.source = None

    def _open(self):
        try:
            import pynvml
            pynvml.nvmlInit()
            h = None
            try:
                uuid = str(torch.cuda.get_device_properties(self.gpu).uuid)
                if not uuid.startswith("GPU-"):
                    uuid = "GPU-" + uuid
                h = pynvml.nvmlDeviceGetHandleByUUID(uuid)
            except Exception:
                h = pynvml.nvmlDeviceGetHandleByIndex(self.gpu)
            self._nvml, self._h = pynvml, h
            self._max = float(pynvml.nvmlDeviceGetMaxClockInfo(h, pynvml.NVML_CLOCK_SM))
            self.source = "nvml"
        except Exception:
            self._nvml = None
            self.source = "nvidia-smi"

    def sample_once(self):
        if self._nvml is not None:
            try:
                n, h = self._nvml, self._h
                sm = float(n.nvmlDeviceGetClockInfo(h, n.NVML_CLOCK_SM))
                try:
                    bits = int(n.nvmlDeviceGetCurrentClocksEventReasons(h))
                except Exception:
                    bits = int(n.nvmlDeviceGetCurrentClocksThrottleReasons(h))
                self.rows.append((sm, self._max, bits))
                return
            except Exception:
                pass
        try:
            out = subprocess.run(["nvidia-smi", f"--id={self.gpu}", f"--query-gpu={self.QUERY}", "--format=csv,noheader,nounits"],
                                 capture_output=True, text=True, timeout=10).stdout
            for line in out.splitlines():
                parts = [q.strip() for q in line.split(",")]
                if len(parts) >= 9:
                    bits = 0
                    for (name, bit), v in zip(self._REASONS, parts[5:9]):
                        if v.lower().startswith("active"):
                            bits |= bit
                    self.rows.append((float(parts[1]), float(parts[2]), bits))
        except Exception:
            pass

    def start(self):
        self._open()
        self.sample_once()
        if self._nvml is not None:
            def pump():
                while not self._stop.wait(self.period_ms / 1e3):
                    self.sample_once()
            self._thr = threading.Thread(target=pump, daemon=True)
            self._thr.start()
        return self

    def stop(self):
        self._stop.set()
        if self._thr is not None:
            self._thr.join(timeout=2)
        self.sample_once()
        return self

    def summary(self) -> dict:
        if not self.rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0, "source": self.source}
        sm = sorted(r[0] for r in self.rows)
        bits = 0
        for r in self.rows:
            bits |= r[2]
        reasons = [name for name, bit in self._REASONS if bits & bit]
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": max(r[1] for r in self.rows), "reasons": reasons,
                "samples": len(sm), "source": self.source}
