"""Device-side timing and clock sampling used by bench.py and the sweeps."""
from __future__ import annotations

import shutil
import statistics
import subprocess
import threading
import time
from typing import List, Optional

import torch


class CudaTimer:
    """CUDA-event stopwatch on the current stream: ``with CudaTimer() as t: ...; t.ms``."""

    def __init__(self, stream: Optional["torch.cuda.Stream"] = None):
        self.stream = stream
        self.start = torch.cuda.Event(enable_timing=True)
        self.end = torch.cuda.Event(enable_timing=True)
        self.ms = float("nan")

    def __enter__(self):
        torch.cuda.synchronize()
        self.start.record(self.stream or torch.cuda.current_stream())
        return self

    def __exit__(self, *exc):
        self.end.record(self.stream or torch.cuda.current_stream())
        torch.cuda.synchronize()
        self.ms = self.start.elapsed_time(self.end)
        return False


class ClockSampler:
    """Samples ``nvidia-smi`` SM clocks and throttle reasons in a background thread while a timed
    region runs (B200_PROFILING.md, "clocks DURING the timed region")."""

    QUERY = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int = 0, period_s: float = 0.2):
        self.gpu_index, self.period_s = gpu_index, period_s
        self.samples: List[List[str]] = []
        self._stop = threading.Event()
        self._thread: Optional[threading.Thread] = None
        self.available = shutil.which("nvidia-smi") is not None

    def _run(self) -> None:
        while not self._stop.is_set():
            try:
                out = subprocess.run(["nvidia-smi", f"--query-gpu={self.QUERY}", "--format=csv,noheader,nounits", "-i", str(self.gpu_index)],
                                     capture_output=True, text=True, timeout=5).stdout.strip()
                if out:
                    self.samples.append([x.strip() for x in out.split(",")])
            except Exception:
                pass
            self._stop.wait(self.period_s)

    def __enter__(self):
        if self.available:
            self._thread = threading.Thread(target=self._run, daemon=True)
            self._thread.start()
        return self

    def __exit__(self, *exc):
        self._stop.set()
        if self._thread is not None:
            self._thread.join(timeout=10)
        return False

    def summary(self) -> dict:
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for s in self.samples:
            try:
                sm.append(float(s[0])); mx.append(float(s[1]))
            except (ValueError, IndexError):
                continue
            for name, val in zip(names, s[3:7]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(self.samples)}
