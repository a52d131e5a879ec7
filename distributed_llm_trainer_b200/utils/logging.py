"""Console + JSONL metrics.

The console line is character-for-character the reference's (ddp_trainer.py:604-609,
fsdp_trainer.py:601-607) so scripts that grep it keep working; the JSONL sink is new (the reference
lists W&B/TensorBoard only in requirements.txt).
"""
from __future__ import annotations

import json
import os
import time
from typing import Any, Dict, Optional


def format_ddp_line(step: int, loss: float, lr: float, tokens_per_sec: float) -> str:
    return (f"Step {step:6d} | "
            f"Loss: {loss:.4f} | "
            f"LR: {lr:.2e} | "
            f"Tokens/sec: {tokens_per_sec:,.0f}")


def format_fsdp_line(step: int, loss: float, lr: float, tokens_per_sec: float, mem_gb: float) -> str:
    return (f"Step {step:6d} | "
            f"Loss: {loss:.4f} | "
            f"LR: {lr:.2e} | "
            f"Tokens/s: {tokens_per_sec:,.0f} | "
            f"Mem: {mem_gb:.1f}GB")


class MetricsLogger:
    """Append-only JSONL writer (rank 0 only); ``None`` path = disabled."""

    def __init__(self, path: Optional[str], enabled: bool = True):
        self.path = path if enabled else None
        self._fh = None
        if self.path:
            os.makedirs(os.path.dirname(os.path.abspath(self.path)), exist_ok=True)
            self._fh = open(self.path, "a", encoding="utf-8")

    def log(self, record: Dict[str, Any]) -> None:
        if self._fh is None:
            return
        rec = {"time": time.time()}
        rec.update(record)
        self._fh.write(json.dumps(rec) + "\n")
        self._fh.flush()

    def close(self) -> None:
        if self._fh is not None:
            self._fh.close()
            self._fh = None
