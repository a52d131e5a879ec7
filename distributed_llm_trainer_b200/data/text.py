"""Text datasets: TinyStories and OpenWebText (map-style and streaming).

Parity targets: reference ``src/data/tinystories.py`` (C33-C35) and ``src/data/openwebtext.py``
(C36).  The two reference files are clones that differ only in gzip support, so here one
implementation serves both and the public names of each are thin aliases:

* items are exactly ``seq_len`` tokens (no +1; the label shift happens in the model);
* map-style length is ``(N - 1) // seq_len``, items are 1-tuples (tinystories.py:44-50);
* streaming shards *lines*: shard id ``rank * num_workers + worker_id`` of
  ``world_size * num_workers`` by ``line_idx % num_shards`` (tinystories.py:87-99), with a
  per-shard ``max_tokens`` budget and an LRU token cache bounded by ``cache_max_tokens``;
* ``.gz`` paths are read transparently and ``path`` <-> ``path.gz`` fall back to each other
  (openwebtext.py:32-37, 147-155).

Differences, on purpose: the LRU cache is consulted *before* tokenising (the reference encodes
first, so its cache never saves work -- SURVEY C34), the tokenizer has an offline fallback, and
``DistributedSampler.set_epoch`` is driven by the training loop.
"""
from __future__ import annotations

import gzip
import os
from collections import OrderedDict
from dataclasses import dataclass
from typing import Iterator, List, Optional

import torch
from torch.utils.data import DataLoader, Dataset, DistributedSampler, IterableDataset, get_worker_info

from .tokenizer import load_tokenizer


@dataclass
class TextDataConfig:
    path: str
    seq_len: int
    tokenizer_name: str = "gpt2"
    max_tokens: Optional[int] = None
    streaming: bool = False
    cache_max_tokens: Optional[int] = None


# reference-compatible names
TinyStoriesConfig = TextDataConfig
OpenWebTextConfig = TextDataConfig


def _open_text(path: str):
    if path.endswith(".gz"):
        return gzip.open(path, "rt", encoding="utf-8")
    return open(path, "r", encoding="utf-8")


def resolve_path(path: str, what: str = "dataset") -> str:
    """``x`` <-> ``x.gz`` fallback (reference openwebtext.py:147-155)."""
    if os.path.exists(path):
        return path
    other = path[:-3] if path.endswith(".gz") else path + ".gz"
    if os.path.exists(other):
        return other
    raise FileNotFoundError(f"{what} file not found: {path}")


class TokenFileDataset(Dataset):
    """Whole file tokenised once, cut into fixed-length windows."""

    def __init__(self, cfg: TextDataConfig, tokenizer=None):
        self.cfg = cfg
        self.tokenizer = tokenizer or load_tokenizer(cfg.tokenizer_name)
        with _open_text(cfg.path) as fh:
            ids = self.tokenizer.encode(fh.read())
        if cfg.max_tokens is not None:
            ids = ids[: cfg.max_tokens]
        if len(ids) < cfg.seq_len:
            raise ValueError(f"Not enough tokens ({len(ids)}) for seq_len={cfg.seq_len} in {cfg.path}")
        self.tokens = torch.tensor(ids, dtype=torch.long)

    def __len__(self) -> int:
        return (len(self.tokens) - 1) // self.cfg.seq_len

    def __getitem__(self, idx: int):
        lo = idx * self.cfg.seq_len
        return (self.tokens[lo: lo + self.cfg.seq_len],)


class TokenStreamDataset(IterableDataset):
    """Line-sharded streaming reader with a bounded LRU tokenisation cache."""

    def __init__(self, cfg: TextDataConfig, rank: int = 0, world_size: int = 1, tokenizer=None):
        self.cfg = cfg
        self.rank = rank
        self.world_size = world_size
        self.tokenizer = tokenizer or load_tokenizer(cfg.tokenizer_name)
        self._lru: "OrderedDict[int, List[int]]" = OrderedDict()
        self._lru_tokens = 0
        self.cache_hits = 0

    def _tokens_for_line(self, line_idx: int, line: str) -> List[int]:
        budget = self.cfg.cache_max_tokens
        if not budget or budget <= 0:
            return self.tokenizer.encode(line)
        hit = self._lru.get(line_idx)
        if hit is not None:
            self._lru.move_to_end(line_idx)
            self.cache_hits += 1
            return hit
        toks = self.tokenizer.encode(line)
        n = len(toks)
        while self._lru and self._lru_tokens + n > budget:
            _, old = self._lru.popitem(last=False)
            self._lru_tokens -= len(old)
        if n <= budget:
            self._lru[line_idx] = toks
            self._lru_tokens += n
        return toks

    def __iter__(self) -> Iterator[torch.Tensor]:
        info = get_worker_info()
        wid = info.id if info is not None else 0
        nworkers = info.num_workers if info is not None else 1
        shard = self.rank * nworkers + wid
        nshards = self.world_size * nworkers
        S = self.cfg.seq_len
        limit = self.cfg.max_tokens
        pending: List[int] = []
        seen = 0
        with _open_text(self.cfg.path) as fh:
            for line_idx, line in enumerate(fh):
                if line_idx % nshards != shard:
                    continue
                toks = self._tokens_for_line(line_idx, line)
                if limit is not None:
                    room = limit - seen
                    if room <= 0:
                        break
                    toks = toks[:room]
                seen += len(toks)
                pending.extend(toks)
                while len(pending) >= S:
                    yield torch.tensor(pending[:S], dtype=torch.long)
                    del pending[:S]
                if limit is not None and seen >= limit:
                    break


def create_text_dataloader(path: str, batch_size: int, seq_len: int, distributed: bool = False, rank: int = 0,
                           world_size: int = 1, tokenizer_name: str = "gpt2", max_tokens: Optional[int] = None,
                           streaming: bool = False, cache_max_tokens: Optional[int] = None, num_workers: int = 2,
                           gz_fallback: bool = False, tokenizer=None) -> DataLoader:
    if gz_fallback:
        path = resolve_path(path, "OpenWebText")
    cfg = TextDataConfig(path, seq_len, tokenizer_name, max_tokens, streaming, cache_max_tokens)
    if streaming:
        ds = TokenStreamDataset(cfg, rank=rank, world_size=world_size, tokenizer=tokenizer)
        sampler, shuffle = None, False
    else:
        ds = TokenFileDataset(cfg, tokenizer=tokenizer)
        sampler = DistributedSampler(ds, num_replicas=world_size, rank=rank) if distributed else None
        shuffle = sampler is None
    return DataLoader(ds, batch_size=batch_size, sampler=sampler, shuffle=shuffle, drop_last=True,
                      pin_memory=torch.cuda.is_available(), num_workers=num_workers)


# ---- reference-named entry points ------------------------------------------------------------
TinyStoriesDataset = TokenFileDataset
TinyStoriesIterableDataset = TokenStreamDataset
OpenWebTextDataset = TokenFileDataset
OpenWebTextIterableDataset = TokenStreamDataset


def create_tinystories_dataloader(path, batch_size, seq_len, **kw) -> DataLoader:
    """Reference ``create_tinystories_dataloader`` (tinystories.py:122-161)."""
    return create_text_dataloader(path, batch_size, seq_len, gz_fallback=False, **kw)


def create_openwebtext_dataloader(path, batch_size, seq_len, **kw) -> DataLoader:
    """Reference ``create_openwebtext_dataloader`` (openwebtext.py:133-181)."""
    return create_text_dataloader(path, batch_size, seq_len, gz_fallback=True, **kw)
