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
    """Samples ``nvidia-smi`` SM clocks / throttle reasons in a background thread DURING a timed
    region (the profiling recipe's clocks line).  ``summary()`` gives the bench JSON's ``clocks``."""

    QUERY = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
             "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int = 0, period_ms: int = 50):
        self.gpu = gpu_index
        self.period_ms = period_ms
        self.rows: List[List[str]] = []
        self._proc: Optional[subprocess.Popen] = None
        self._thr: Optional[threading.Thread] = None

    def start(self):
        try:
            self._proc = subprocess.Popen(
                ["nvidia-smi", f"--id={self.gpu}", f"--query-gpu={self.QUERY}", "--format=csv,noheader,nounits",
                 "-lms", str(self.period_ms)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except OSError:
            self._proc = None
            return self

        def pump():
            for line in self._proc.stdout:
                parts = [p.strip() for p in line.split(",")]
                if len(parts) >= 9:
                    self.rows.append(parts)
        self._thr = threading.Thread(target=pump, daemon=True)
        self._thr.start()
        return self

    def stop(self):
        if self._proc is not None:
            self._proc.terminate()
            try:
                self._proc.wait(timeout=2)
            except subprocess.TimeoutExpired:
                self._proc.kill()
        if self._thr is not None:
            self._thr.join(timeout=2)
        return self

    def summary(self) -> dict:
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[1]))
                mx.append(float(r[2]))
            except ValueError:
                continue
            for name, v in zip(names, r[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": max(mx), "reasons": sorted(reasons),
                "samples": len(sm)}
