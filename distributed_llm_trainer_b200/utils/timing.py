"""Device-side step timing (CUDA events on the launching stream, max over ranks) and clock sampling.

The reference reports cumulative wall-clock tokens/sec from before the first batch
(ddp_trainer.py:600-602), which folds warm-up, worker start-up and checkpoint writes into the number;
benchmarks here use this module instead (SURVEY section 5 'Metrics')."""
from __future__ import annotations

import os
import statistics
import subprocess
import threading
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


# NVML clocks-event-reason bits (nvml.h)
_REASONS = {
    0x0000000000000004: "sw_power_cap",
    0x0000000000000008: "hw_slowdown",
    0x0000000000000020: "sw_thermal_slowdown",
    0x0000000000000040: "hw_thermal_slowdown",
    0x0000000000000080: "hw_power_brake_slowdown",
}


class ClockSampler:
    """Samples SM clocks and throttle reasons DURING a timed region from a background thread.
    Uses NVML directly (a sample every ~20 ms); falls back to polling ``nvidia-smi`` when NVML is unusable."""

    QUERY = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
             "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int = 0, period_s: float = 0.02):
        self.gpu_index = gpu_index
        self.period_s = period_s
        self.sm: List[float] = []
        self.sm_max: Optional[float] = None
        self.power: List[float] = []
        self.reasons = set()
        self.source = "none"
        self._stop = threading.Event()
        self._thread: Optional[threading.Thread] = None

    # ---- NVML path ------------------------------------------------------------------------------------
    def _run_nvml(self) -> bool:
        try:
            import pynvml as N

            N.nvmlInit()
            idx = self.gpu_index
            vis = [v for v in (os.environ.get("CUDA_VISIBLE_DEVICES") or "").split(",") if v.strip()]
            if vis and idx < len(vis) and vis[idx].strip().isdigit():
                idx = int(vis[idx])
            h = N.nvmlDeviceGetHandleByIndex(idx)
            self.sm_max = float(N.nvmlDeviceGetMaxClockInfo(h, N.NVML_CLOCK_SM))
        except Exception:
            return False
        self.source = "nvml"
        get_reasons = getattr(N, "nvmlDeviceGetCurrentClocksEventReasons", None) or getattr(
            N, "nvmlDeviceGetCurrentClocksThrottleReasons", None)
        while not self._stop.is_set():
            try:
                self.sm.append(float(N.nvmlDeviceGetClockInfo(h, N.NVML_CLOCK_SM)))
                self.power.append(N.nvmlDeviceGetPowerUsage(h) / 1000.0)
                if get_reasons is not None:
                    bits = int(get_reasons(h))
                    for bit, name in _REASONS.items():
                        if bits & bit:
                            self.reasons.add(name)
            except Exception:
                pass
            self._stop.wait(self.period_s)
        return True

    # ---- nvidia-smi fallback ------------------------------------------------------------------------------
    def _run_smi(self) -> None:
        self.source = "nvidia-smi"
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        while not self._stop.is_set():
            try:
                out = subprocess.run(["nvidia-smi", f"--query-gpu={self.QUERY}", "--format=csv,noheader,nounits",
                                      "-i", str(self.gpu_index)], capture_output=True, text=True, timeout=5).stdout
                row = [c.strip() for c in out.strip().split(",")]
                if len(row) >= 7:
                    self.sm.append(float(row[0]))
                    self.sm_max = float(row[1])
                    self.power.append(float(row[2]))
                    for name, val in zip(names, row[3:7]):
                        if val.lower().startswith("active"):
                            self.reasons.add(name)
            except Exception:
                pass
            self._stop.wait(max(self.period_s, 0.1))

    def _run(self):
        if not self._run_nvml():
            self._run_smi()

    def __enter__(self):
        self._thread = threading.Thread(target=self._run, daemon=True)
        self._thread.start()
        return self

    def __exit__(self, *exc):
        self._stop.set()
        if self._thread is not None:
            self._thread.join(timeout=10)

    def summary(self) -> Dict[str, object]:
        return {"sm_mhz": statistics.median(self.sm) if self.sm else None, "sm_max_mhz": self.sm_max,
                "power_w_max": max(self.power) if self.power else None, "samples": len(self.sm),
                "reasons": sorted(self.reasons), "source": self.source}
