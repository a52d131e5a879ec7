"""Warm-up + cosine learning-rate schedule.

Both reference trainers use linear warm-up to ``learning_rate`` over ``warmup_steps`` and a cosine
decay to ``0.1 * learning_rate`` at ``max_steps``; they differ in one detail that is kept:

* DDP  (ddp_trainer.py:262-271): the decay ratio is **not clamped**, so past ``max_steps`` the cosine
  keeps oscillating.
* FSDP (fsdp_trainer.py:345-356): ``decay_ratio = min(ratio, 1.0)``.
"""
from __future__ import annotations

import math


def cosine_lr(step: int, base_lr: float, warmup_steps: int, max_steps: int, clamp: bool) -> float:
    if step < warmup_steps:
        return base_lr * (step / warmup_steps)
    span = max_steps - warmup_steps
    ratio = (step - warmup_steps) / span if span != 0 else 1.0
    if clamp:
        ratio = min(ratio, 1.0)
    floor = 0.1 * base_lr
    return floor + 0.5 * (1.0 + math.cos(math.pi * ratio)) * (base_lr - floor)
