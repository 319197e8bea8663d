"""Device-side timing.

The reference times epochs with ``MPI.Wtime()`` (/root/reference/main.py:145,158), a
host wall clock that also swallows data loading and waiting for rank 0.  Here every
number is taken with CUDA events on the launching stream and reduced as the MAX over
ranks; on CPU (tests) it degrades to ``time.perf_counter``.
"""
from __future__ import annotations

import statistics
import subprocess
import threading
import time
from typing import Dict, List, Optional

import torch


class DeviceTimer:
    """``with DeviceTimer(device) as t: ...`` then ``t.ms``."""

    def __init__(self, device: torch.device | str = "cpu", sync: bool = True):
        self.device = torch.device(device)
        self.cuda = self.device.type == "cuda"
        self.sync = sync
        self.ms = 0.0

    def __enter__(self):
        if self.cuda:
            if self.sync:
                torch.cuda.synchronize(self.device)
            self._s = torch.cuda.Event(enable_timing=True)
            self._e = torch.cuda.Event(enable_timing=True)
            self._s.record(torch.cuda.current_stream(self.device))
        else:
            self._t0 = time.perf_counter()
        return self

    def __exit__(self, *exc):
        if self.cuda:
            self._e.record(torch.cuda.current_stream(self.device))
            self._e.synchronize()
            self.ms = float(self._s.elapsed_time(self._e))
        else:
            self.ms = (time.perf_counter() - self._t0) * 1e3
        return False


def wtime() -> float:
    """Drop-in for ``MPI.Wtime()`` (host seconds).  Kept for log-format parity only."""
    return time.perf_counter()


def max_over_ranks(value: float, device: torch.device | str = "cpu") -> float:
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return float(value)
    t = torch.tensor([float(value)], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


_QUERY = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
          "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
          "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")


class ClockSampler:
    """Samples ``nvidia-smi`` SM clocks / throttle reasons while a timed region runs
    (B200_PROFILING.md 'clocks DURING the timed region')."""

    def __init__(self, gpu_index: int = 0, period_ms: int = 200):
        self.gpu_index = gpu_index
        self.period_ms = period_ms
        self.rows: List[List[str]] = []
        self._proc: Optional[subprocess.Popen] = None
        self._thr: Optional[threading.Thread] = None

    def start(self):
        try:
            self._proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.gpu_index), "--query-gpu=" + _QUERY,
                 "--format=csv,noheader,nounits", "-lms", str(self.period_ms)],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except (OSError, ValueError):
            self._proc = None
            return self

        def _pump():
            assert self._proc is not None and self._proc.stdout is not None
            for line in self._proc.stdout:
                parts = [p.strip() for p in line.strip().split(",")]
                if len(parts) >= 9:
                    self.rows.append(parts)

        self._thr = threading.Thread(target=_pump, daemon=True)
        self._thr.start()
        return self

    def stop(self) -> Dict[str, object]:
        if self._proc is not None:
            self._proc.terminate()
            try:
                self._proc.wait(timeout=5)
            except Exception:
                self._proc.kill()
        if self._thr is not None:
            self._thr.join(timeout=2)
        return self.summary()

    def summary(self) -> Dict[str, object]:
        sm: List[float] = []
        mx = 0.0
        reasons = set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[1]))
                mx = max(mx, float(r[2]))
            except ValueError:
                continue
            for name, val in zip(names, r[5:9]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        # median over the upper half = "under load" (idle samples at the edges drag it down)
        sm_sorted = sorted(sm)
        load = sm_sorted[len(sm_sorted) // 2:] if len(sm_sorted) > 3 else sm_sorted
        return {"sm_mhz": statistics.median(load), "sm_max_mhz": mx,
                "reasons": sorted(reasons), "samples": len(sm)}
