"""Tracing / profiling hooks of the training loops.

The reference has none (its guide only says "profile with torch.profiler", docs/LEARNING_GUIDE.md:225;
SURVEY section 5 "Tracing / profiling").  Three independent, opt-in pieces:

* ``--profile_steps N [--profile_start S] [--profile_dir D]``: ``torch.profiler`` (CPU + CUDA activities) over
  N optimizer steps starting at step S; every rank writes ``D/trace_rank{r}.json`` (chrome / perfetto) and
  ``D/kernels_rank{r}.txt`` (kernel table sorted by device time).  The sm_100a engine's kernels appear under
  their C++ names (``b200::g2::gemm2_kernel<...>``, ``b200::attn_fwd_kernel`` ...).
* ``--nvtx``: NVTX ranges ``step N`` / ``train_step`` around every optimizer step, for ``nsys`` / ``ncu --nvtx``.
* device step time: CUDA events around every ``train_step`` on CUDA devices; the loop logs ``step_ms`` (the
  device time of the logged step) next to the wall-clock tokens/sec the reference prints.

Nothing here runs unless asked for, and the event pair costs two stream records per step.
"""
from __future__ import annotations

import os
from typing import Optional

import torch


class StepTracer:
    def __init__(self, train_cfg, rank: int = 0, device: Optional[torch.device] = None):
        self.rank = rank
        self.n = int(getattr(train_cfg, "profile_steps", 0) or 0)
        self.start = int(getattr(train_cfg, "profile_start", 5) or 0)
        self.dir = getattr(train_cfg, "profile_dir", None) or "profiles_run"
        self.nvtx = bool(getattr(train_cfg, "nvtx", False)) and torch.cuda.is_available()
        self.cuda = device is not None and device.type == "cuda"
        self._prof = None
        self._ev = None
        self.step_ms: Optional[float] = None

    # ------------------------------------------------------------------ per step
    def begin(self, step: int) -> None:
        if self.n > 0 and self._prof is None and step == self.start:
            acts = [torch.profiler.ProfilerActivity.CPU]
            if self.cuda:
                acts.append(torch.profiler.ProfilerActivity.CUDA)
            self._prof = torch.profiler.profile(activities=acts, record_shapes=False)
            self._prof.__enter__()
            self._prof_until = step + self.n
        if self.nvtx:
            torch.cuda.nvtx.range_push(f"step {step}")
        if self.cuda:
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            self._ev = (a, b)

    def end(self, step: int) -> None:
        if self._ev is not None:
            self._ev[1].record()
        if self.nvtx:
            torch.cuda.nvtx.range_pop()
        if self._prof is not None and step + 1 >= self._prof_until:
            self._finish_profile()

    def read_step_ms(self) -> Optional[float]:
        """Device time of the last step (synchronises on its end event -- call it where the loop already
        holds the step's loss on the host, i.e. at a log line)."""
        if self._ev is None:
            return None
        self._ev[1].synchronize()
        self.step_ms = self._ev[0].elapsed_time(self._ev[1])
        return self.step_ms

    # ------------------------------------------------------------------ profiler output
    def _finish_profile(self) -> None:
        prof, self._prof, self.n = self._prof, None, 0
        if self.cuda:
            torch.cuda.synchronize()
        prof.__exit__(None, None, None)
        os.makedirs(self.dir, exist_ok=True)
        trace = os.path.join(self.dir, f"trace_rank{self.rank}.json")
        prof.export_chrome_trace(trace)
        sort = "cuda_time_total" if self.cuda else "cpu_time_total"
        try:
            table = prof.key_averages().table(sort_by=sort, row_limit=40)
        except Exception:   # older / newer torch name the column differently
            table = prof.key_averages().table(row_limit=40)
        with open(os.path.join(self.dir, f"kernels_rank{self.rank}.txt"), "w") as f:
            f.write(table)
        if self.rank == 0:
            print(f"Profiler trace written to {trace}")

    def close(self) -> None:
        if self._prof is not None:
            self._finish_profile()
