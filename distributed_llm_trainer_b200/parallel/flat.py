"""Flat parameter / gradient / shadow storage.

Every trainable tensor of a ``GPT`` lives at a fixed offset of one flat fp32 *master* buffer, in
``named_parameters()`` order (the order the reference's optimizer state is keyed by,
ddp_trainer.py:214-234).  Alongside it:

* ``grad``   -- flat fp32 gradient accumulator, same offsets; ``param.grad`` are views of it, so the
  wgrad GEMM epilogues accumulate straight into the buffer the comm kernels later read.
* ``shadow`` -- flat bf16 compute copy, same offsets, rewritten by the fused AdamW kernel
  (replaces the per-forward fp32->bf16 autocast casts of the reference).

Offsets are padded to ``ALIGN`` elements so (a) every tensor base is 16-byte aligned for TMA in
both fp32 and bf16 views and (b) a ``ALIGN``-element chunk never straddles two tensors, which lets
one AdamW launch cover the whole buffer with a per-chunk weight-decay table.  Consecutive
``q/k/v_proj`` (and ``gate/up_proj``) land back to back *without* padding between them when their
sizes are ``ALIGN`` multiples, which is what gives the packed ``[3h, h]`` / ``[2I, h]`` GEMM operands
while the state dict still exposes separate tensors.

Units: ``root`` = embedding (+ tied head) and final norm, then one unit per transformer block --
the same partition the reference's FSDP wrap policy produces (fsdp_trainer.py:183-186).  DDP
buckets and FSDP shards are both defined on these units.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Dict, Iterable, List, Optional, Tuple

import torch
import torch.nn as nn

ALIGN = 128


@dataclass
class ParamSlot:
    name: str
    shape: Tuple[int, ...]
    offset: int            # element offset in the flat buffers
    numel: int
    decay: bool            # weight decay applies (reference rule: no "bias"/"norm" in the name)
    unit: int              # 0 = root, 1 + layer_idx otherwise
    index: int             # position in named_parameters() order


@dataclass
class UnitRange:
    name: str
    lo: int
    hi: int                # [lo, hi) element range, ALIGN-padded
    slots: List[int] = field(default_factory=list)


def _unit_of(name: str) -> int:
    if name.startswith("layers."):
        return 1 + int(name.split(".")[1])
    return 0


def plan_layout(named_params: Iterable[Tuple[str, torch.Tensor]]) -> Tuple[List[ParamSlot], List[UnitRange], int]:
    """Assign offsets.  Physical order is unit-major (root = embedding + final norm first, then
    block 0, 1, ...) so every unit is one contiguous range; ``ParamSlot.index`` keeps the
    ``named_parameters()`` position, which is what optimizer state dicts are keyed by.  The
    returned slot list is in ``named_parameters()`` order."""
    entries = [(idx, name, p) for idx, (name, p) in enumerate(named_params)]
    slots_by_index: Dict[int, ParamSlot] = {}
    cursor = 0
    for idx, name, p in sorted(entries, key=lambda e: (_unit_of(e[1]), e[0])):
        n = p.numel()
        slots_by_index[idx] = ParamSlot(name, tuple(p.shape), cursor, n,
                                        not ("bias" in name or "norm" in name), _unit_of(name), idx)
        cursor += (n + ALIGN - 1) // ALIGN * ALIGN
    slots = [slots_by_index[i] for i in range(len(entries))]
    total = cursor
    units: Dict[int, UnitRange] = {}
    for s in slots:
        hi = s.offset + (s.numel + ALIGN - 1) // ALIGN * ALIGN
        u = units.get(s.unit)
        if u is None:
            units[s.unit] = UnitRange("root" if s.unit == 0 else f"layers.{s.unit - 1}", s.offset, hi, [s.index])
        else:
            u.lo, u.hi = min(u.lo, s.offset), max(u.hi, hi)
            u.slots.append(s.index)
    return slots, [units[k] for k in sorted(units)], total


class FlatParamStore:
    """Owns the flat buffers and re-points the module's parameters at views of them."""

    def __init__(self, model: nn.Module, device: Optional[torch.device] = None, with_grad: bool = True,
                 with_shadow: bool = True, grad_buffer: Optional[torch.Tensor] = None,
                 shadow_buffer: Optional[torch.Tensor] = None):
        named = [(n, p) for n, p in model.named_parameters() if p.requires_grad]
        self.slots, self.units, self.total = plan_layout(named)
        self.by_name = {s.name: s for s in self.slots}
        dev = device or named[0][1].device
        self.device = dev
        self.master = torch.zeros(self.total, dtype=torch.float32, device=dev)
        if with_grad:
            self.grad = grad_buffer if grad_buffer is not None else torch.zeros(self.total, dtype=torch.float32, device=dev)
            assert self.grad.numel() >= self.total
        else:
            self.grad = None
        if with_shadow:
            self.shadow = shadow_buffer if shadow_buffer is not None else torch.zeros(self.total, dtype=torch.bfloat16, device=dev)
        else:
            self.shadow = None
        self.params: List[nn.Parameter] = []
        with torch.no_grad():
            for (name, p), s in zip(named, self.slots):
                view = self.master[s.offset: s.offset + s.numel].view(s.shape)
                view.copy_(p.detach().to(device=dev, dtype=torch.float32))
                p.data = view
                if self.grad is not None:
                    p.grad = self.grad[s.offset: s.offset + s.numel].view(s.shape)
                self.params.append(p)
        # per-chunk decay table for the single-launch AdamW kernel
        self.decay_chunks = torch.zeros(self.total // ALIGN, dtype=torch.uint8)
        for s in self.slots:
            if s.decay:
                self.decay_chunks[s.offset // ALIGN: (s.offset + s.numel + ALIGN - 1) // ALIGN] = 1
        self.decay_chunks = self.decay_chunks.to(dev)
        if self.shadow is not None:
            self.refresh_shadow()

    # ---- views -------------------------------------------------------------------------------
    def view(self, buf: torch.Tensor, name: str) -> torch.Tensor:
        s = self.by_name[name]
        return buf[s.offset: s.offset + s.numel].view(s.shape)

    def packed(self, buf: torch.Tensor, names: List[str]) -> torch.Tensor:
        """Row-concatenated view over consecutive 2-D tensors (e.g. q,k,v -> [3h, h])."""
        ss = [self.by_name[n] for n in names]
        cols = ss[0].shape[1]
        for a, b in zip(ss, ss[1:]):
            if a.offset + a.numel != b.offset or b.shape[1] != cols:
                raise ValueError(f"{[s.name for s in ss]} are not packable (padding or shape mismatch)")
        rows = sum(s.shape[0] for s in ss)
        return buf[ss[0].offset: ss[0].offset + rows * cols].view(rows, cols)

    def refresh_shadow(self) -> None:
        self.shadow.copy_(self.master)

    def zero_grad(self) -> None:
        if self.grad is not None:
            self.grad.zero_()

    def rebind_grads(self) -> None:
        """Restore ``param.grad`` views after something set them to ``None``."""
        for p, s in zip(self.params, self.slots):
            p.grad = self.grad[s.offset: s.offset + s.numel].view(s.shape)

    # ---- optimizer-state index maps -------------------------------------------------------------
    def group_indices(self) -> Tuple[List[int], List[int]]:
        """(decay, no_decay) slot indices in the order the reference builds its two param groups."""
        d = [s.index for s in self.slots if s.decay]
        nd = [s.index for s in self.slots if not s.decay]
        return d, nd
