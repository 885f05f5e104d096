"""Lightweight tracing: NVTX ranges + CUDA-event phase timers + a JSONL metrics sink.

The reference has no profiling, timers or structured logs; these are
additions.  ``Tracer.range`` pushes an NVTX range (visible in Nsight) and, when ``timing`` is
on, brackets the region with CUDA events whose elapsed times are summed per name without a
host sync until ``summary()`` is called.
"""
from __future__ import annotations

import contextlib
import json
import time
from collections import defaultdict
from typing import Dict, Optional

import torch


class Tracer:
    def __init__(self, enabled: bool = True, timing: bool = False, jsonl_path: Optional[str] = None):
        self.enabled, self.timing = enabled, timing
        self._cuda = torch.cuda.is_available()
        self._events = defaultdict(list)
        self._cpu = defaultdict(float)
        self._sink = open(jsonl_path, "a") if jsonl_path else None

    @contextlib.contextmanager
    def range(self, name: str):
        if not self.enabled:
            yield
            return
        if self._cuda:
            torch.cuda.nvtx.range_push(name)
            if self.timing:
                a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a.record()
        t0 = time.perf_counter()
        try:
            yield
        finally:
            self._cpu[name] += time.perf_counter() - t0
            if self._cuda:
                if self.timing:
                    b.record()
                    self._events[name].append((a, b))
                torch.cuda.nvtx.range_pop()

    def summary(self) -> Dict[str, Dict[str, float]]:
        out = {}
        if self._cuda and self.timing:
            torch.cuda.synchronize()
        for k, v in self._cpu.items():
            out[k] = {"cpu_ms": v * 1e3}
        for k, evs in self._events.items():
            out.setdefault(k, {})["gpu_ms"] = sum(a.elapsed_time(b) for a, b in evs)
        return out

    def log(self, **record):
        if self._sink:
            record.setdefault("ts", time.time())
            self._sink.write(json.dumps(record) + "\n")
            self._sink.flush()

    def close(self):
        if self._sink:
            self._sink.close()
            self._sink = None
