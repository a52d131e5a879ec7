"""Device-side step timing (CUDA events on the launching stream, max over ranks) and clock sampling.

The reference reports cumulative wall-clock tokens/sec from before the first batch
(ddp_trainer.py:600-602), which folds warm-up, worker start-up and checkpoint writes into the number;
benchmarks here use this module instead (SURVEY section 5 'Metrics')."""
from __future__ import annotations

import statistics
import subprocess
import threading
import time
from typing import Dict, List, Optional

import torch


class CudaStepTimer:
    def __init__(self):
        self.start = torch.cuda.Event(enable_timing=True)
        self.stop = torch.cuda.Event(enable_timing=True)

    def __enter__(self):
        torch.cuda.synchronize()
        self.start.record()
        return self

    def __exit__(self, *exc):
        self.stop.record()
        torch.cuda.synchronize()
        self.ms = self.start.elapsed_time(self.stop)


class ClockSampler:
    """Samples ``nvidia-smi`` SM clocks / throttle reasons in a background thread during a timed region."""

    QUERY = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
             "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int = 0, period_s: float = 0.2):
        self.gpu_index = gpu_index
        self.period_s = period_s
        self.samples: List[List[str]] = []
        self._stop = threading.Event()
        self._thread: Optional[threading.Thread] = None

    def _run(self):
        while not self._stop.is_set():
            try:
                out = subprocess.run(["nvidia-smi", f"--query-gpu={self.QUERY}", "--format=csv,noheader,nounits",
                                      "-i", str(self.gpu_index)], capture_output=True, text=True, timeout=5).stdout
                row = [c.strip() for c in out.strip().split(",")]
                if len(row) >= 7:
                    self.samples.append(row)
            except Exception:
                pass
            self._stop.wait(self.period_s)

    def __enter__(self):
        self._thread = threading.Thread(target=self._run, daemon=True)
        self._thread.start()
        return self

    def __exit__(self, *exc):
        self._stop.set()
        if self._thread is not None:
            self._thread.join(timeout=10)

    def summary(self) -> Dict[str, object]:
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for row in self.samples:
            try:
                sm.append(float(row[0]))
                mx.append(float(row[1]))
            except ValueError:
                continue
            for name, val in zip(names, row[3:7]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}
