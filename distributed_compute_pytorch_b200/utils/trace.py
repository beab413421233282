"""Lightweight tracing (the reference has only ``time.time()`` around an epoch, ref:main.py:128-135).

``Tracer`` records named ranges three ways at once:
  * NVTX ranges (``torch.cuda.nvtx``), so a profiler attached from outside sees the same names;
  * CUDA events on the current stream, so device time per range is known without a profiler;
  * host wall-clock, so launch-bound phases show up as the difference of the two.
``export_chrome(path)`` writes a ``chrome://tracing`` / Perfetto JSON with one host row and one device row per rank.
Disabled tracers cost one attribute test per range.  ``DCP_TRACE=/path/prefix`` enables the tracer main.py creates.

    tr = Tracer(enabled=True)
    with tr.range("forward"):
        out = model(x)
    tr.export_chrome("trace_rank0.json")
"""
from __future__ import annotations

import json
import os
import time
from contextlib import contextmanager
from typing import List, Optional

import torch


class Tracer:
    def __init__(self, enabled: Optional[bool] = None, rank: int = 0, device_timing: bool = True):
        self.enabled = bool(os.environ.get("DCP_TRACE")) if enabled is None else enabled
        self.rank = rank
        self.cuda = device_timing and torch.cuda.is_available()
        self._t0 = time.perf_counter()
        self._spans: List[dict] = []          # {"name", "h0", "h1", "e0", "e1"}
        self._origin = None
        if self.enabled and self.cuda:
            self._origin = torch.cuda.Event(enable_timing=True)
            self._origin.record()

    @contextmanager
    def range(self, name: str):
        if not self.enabled:
            yield
            return
        span = {"name": name, "h0": time.perf_counter() - self._t0}
        if self.cuda:
            torch.cuda.nvtx.range_push(name)
            span["e0"] = torch.cuda.Event(enable_timing=True)
            span["e1"] = torch.cuda.Event(enable_timing=True)
            span["e0"].record()
        try:
            yield
        finally:
            if self.cuda:
                span["e1"].record()
                torch.cuda.nvtx.range_pop()
            span["h1"] = time.perf_counter() - self._t0
            self._spans.append(span)

    def mark(self, name: str) -> None:
        """Zero-length host marker (e.g. 'epoch 3')."""
        if self.enabled:
            t = time.perf_counter() - self._t0
            self._spans.append({"name": name, "h0": t, "h1": t})

    def summary(self) -> dict:
        """name -> {"count", "host_ms", "device_ms"} (synchronises once)."""
        if self.cuda and self._spans:
            torch.cuda.synchronize()
        out: dict = {}
        for s in self._spans:
            d = out.setdefault(s["name"], {"count": 0, "host_ms": 0.0, "device_ms": 0.0})
            d["count"] += 1
            d["host_ms"] += (s["h1"] - s["h0"]) * 1e3
            if "e0" in s:
                d["device_ms"] += s["e0"].elapsed_time(s["e1"])
        return out

    def export_chrome(self, path: str) -> str:
        if self.cuda and self._spans:
            torch.cuda.synchronize()
        ev = []
        for s in self._spans:
            ev.append({"name": s["name"], "ph": "X", "pid": self.rank, "tid": 0, "ts": s["h0"] * 1e6, "dur": max(s["h1"] - s["h0"], 0.0) * 1e6})
            if "e0" in s and self._origin is not None:
                ev.append({"name": s["name"], "ph": "X", "pid": self.rank, "tid": 1, "ts": self._origin.elapsed_time(s["e0"]) * 1e3,
                           "dur": s["e0"].elapsed_time(s["e1"]) * 1e3})
        meta = [{"name": "thread_name", "ph": "M", "pid": self.rank, "tid": 0, "args": {"name": "host"}},
                {"name": "thread_name", "ph": "M", "pid": self.rank, "tid": 1, "args": {"name": "device (stream order)"}}]
        with open(path, "w") as f:
            json.dump({"traceEvents": meta + ev, "displayTimeUnit": "ms"}, f)
        return path
