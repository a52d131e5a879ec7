"""Synthetic token loader.

Parity target: ``create_dummy_dataloader`` in both reference trainers
(ddp_trainer.py:460-487, fsdp_trainer.py:508-527): ``num_batches * batch_size`` random int64
sequences in a ``TensorDataset``, ``DistributedSampler`` when distributed, pinned batches.

Differences, on purpose (SURVEY section 5 "RNG / determinism"): generation is seeded, and each
rank generates only the rows its sampler will visit instead of the whole data set; batches come
back as 1-element lists exactly like the reference's ``TensorDataset`` loader so the training
loops' ``batch[0]`` handling is unchanged.
"""
from __future__ import annotations

from typing import Optional

import torch
from torch.utils.data import DataLoader, DistributedSampler, TensorDataset


def create_dummy_dataloader(batch_size: int, seq_len: int, vocab_size: int, num_batches: int = 1000,
                            distributed: bool = False, rank: int = 0, world_size: int = 1,
                            seed: Optional[int] = None, num_workers: int = 0,
                            per_rank_slice: bool = True) -> DataLoader:
    gen = torch.Generator()
    if seed is not None:
        gen.manual_seed(int(seed) * 1000003 + (rank if per_rank_slice else 0))
    else:
        gen.seed()
    use_sampler = distributed and not per_rank_slice
    rows = num_batches * batch_size
    if distributed and per_rank_slice:
        rows = max(batch_size, rows // max(world_size, 1))
    data = torch.randint(0, vocab_size, (rows, seq_len), generator=gen, dtype=torch.long)
    ds = TensorDataset(data)
    sampler = DistributedSampler(ds, num_replicas=world_size, rank=rank) if use_sampler else None
    shuffle_gen = torch.Generator()
    shuffle_gen.manual_seed(0 if seed is None else int(seed) + 17 * rank)
    return DataLoader(ds, batch_size=batch_size, sampler=sampler, shuffle=(sampler is None),
                      generator=shuffle_gen if sampler is None else None,
                      pin_memory=torch.cuda.is_available(), num_workers=num_workers)
