"""Process-group bootstrap shared by both trainers (reference ddp_trainer.py:93-113,
fsdp_trainer.py:125-138).  Backend is ``nccl`` as in the reference; ``gloo`` is accepted so the
host-side logic runs under CPU-only multi-process tests."""
from __future__ import annotations

import datetime
import os
from dataclasses import dataclass

import torch
import torch.distributed as dist


@dataclass
class DistInfo:
    distributed: bool
    rank: int
    world_size: int
    local_rank: int

    @property
    def is_main(self) -> bool:
        return self.rank == 0


def init_distributed(backend: str = "nccl", require: bool = False) -> DistInfo:
    want = dist.is_initialized() or "RANK" in os.environ
    if not want:
        if require:
            raise RuntimeError("this trainer must be launched with torchrun (RANK not set)")
        return DistInfo(False, 0, 1, 0)
    local_rank = int(os.environ.get("LOCAL_RANK", 0))
    if not dist.is_initialized():
        if backend == "nccl" and not torch.cuda.is_available():
            backend = "gloo"
        kwargs = {}
        if backend == "nccl":
            torch.cuda.set_device(local_rank)
            kwargs["device_id"] = torch.device(f"cuda:{local_rank}")
        dist.init_process_group(backend=backend, timeout=datetime.timedelta(minutes=10), **kwargs)
    return DistInfo(True, dist.get_rank(), dist.get_world_size(), local_rank)


def all_reduce_max_float(value: float, device) -> float:
    if not (dist.is_available() and dist.is_initialized()):
        return value
    t = torch.tensor([value], dtype=torch.float64, device=device if dist.get_backend() == "nccl" else "cpu")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())
