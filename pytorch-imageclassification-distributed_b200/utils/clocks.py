"""Sample SM clocks / throttle reasons with nvidia-smi while a timed region runs
(B200_PROFILING.md 'clocks DURING the timed region')."""
from __future__ import annotations

import shutil
import statistics
import subprocess
import threading
from typing import Dict, List, Optional

_Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
      "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
      "clocks_event_reasons.sw_power_cap")
_REASONS = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]


class ClockSampler:
    def __init__(self, gpu_index: int = 0, period_ms: int = 200):
        self.gpu_index, self.period_ms = gpu_index, period_ms
        self.rows: List[List[str]] = []
        self._proc: Optional[subprocess.Popen] = None
        self._thread: Optional[threading.Thread] = None

    def start(self):
        if shutil.which("nvidia-smi") is None:
            return self
        self._proc = subprocess.Popen(
            ["nvidia-smi", f"--id={self.gpu_index}", f"--query-gpu={_Q}", "--format=csv,noheader,nounits",
             "-lms", str(self.period_ms)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)

        def pump():
            for line in self._proc.stdout:
                self.rows.append([c.strip() for c in line.split(",")])
        self._thread = threading.Thread(target=pump, daemon=True)
        self._thread.start()
        return self

    def stop(self) -> Dict:
        if self._proc is not None:
            self._proc.terminate()
            try:
                self._proc.wait(timeout=5)
            except Exception:
                self._proc.kill()
            if self._thread:
                self._thread.join(timeout=2)
        sm, mx, reasons = [], [], set()
        for r in self.rows:
            try:
                sm.append(float(r[0])); mx.append(float(r[1]))
            except (ValueError, IndexError):
                continue
            for name, val in zip(_REASONS, r[3:7]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        return {"sm_mhz": statistics.median(sm), "sm_max_mhz": max(mx), "reasons": sorted(reasons),
                "samples": len(sm)}
