"""Who owns which rows: the partition behind the GEMM -> reduce-scatter fusion.

The reference reduces gradients with collectives that know nothing about tensors: DDP all-reduces flat
buckets (ddp_trainer.py:329-342), FSDP reduce-scatters flat 1/N slices of a unit (fsdp_trainer.py:383).
A flat slice boundary falls in the middle of a matrix row, which a TMA box cannot follow, so a GEMM
epilogue could not deliver its tile to "the owner".  Here ownership is defined on what the wgrad GEMMs
produce instead:

* every weight matrix (q|k|v packed, o, gate|up packed, down, embedding) is cut into **128-row slabs** --
  exactly the rows one CTA of the pair GEMM holds in TMEM -- and slab ``s`` of ``S`` belongs to rank
  ``((s * world) // S + rot) % world``: one contiguous run of slabs per rank, ``rot`` (the group's ordinal)
  spreading the remainders and the matrices with fewer slabs than ranks;
* 1-D tensors (norm weights) are owned whole, round-robin.

The same rule is implemented in ``csrc/gemm2_tcgen05.cu`` (``b200_slab_owner`` /
``b200_owner_slab_range``); ``tests/test_ownership_cpu.py`` checks the two against each other.
An owner keeps the fp32 master copy, both AdamW moments and the reduced gradient of its rows; the wgrad
GEMMs of every rank reduce-add their tiles straight into the owner's gradient buffer over NVLink.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Dict, List, Sequence, Tuple

from .flat import ALIGN, ParamSlot

SLAB = 128


@dataclass
class SlabGroup:
    """One wgrad destination: consecutive 2-D parameters with the same column count, viewed as one matrix."""
    key: str
    offset: int          # flat element offset of row 0
    rows: int
    cols: int
    rot: int
    names: Tuple[str, ...]

    @property
    def slabs(self) -> int:
        return (self.rows + SLAB - 1) // SLAB

    def owner_of_slab(self, s: int, world: int) -> int:
        return ((s * world) // self.slabs + self.rot) % world

    def slab_range(self, owner: int, world: int) -> Tuple[int, int]:
        """Slabs [first, last) owned by ``owner`` (empty when first >= last)."""
        op = (owner - self.rot) % world
        S = self.slabs
        return (op * S + world - 1) // world, ((op + 1) * S + world - 1) // world

    def row_range(self, owner: int, world: int) -> Tuple[int, int]:
        a, b = self.slab_range(owner, world)
        return min(self.rows, a * SLAB), min(self.rows, b * SLAB)

    def flat_range(self, owner: int, world: int) -> Tuple[int, int]:
        a, b = self.row_range(owner, world)
        return self.offset + a * self.cols, self.offset + b * self.cols


@dataclass
class SmallTensor:
    name: str
    offset: int
    numel: int           # padded to ALIGN
    owner: int


def group_key(name: str) -> str:
    """Parameter name -> key of the slab group it belongs to ('' for 1-D tensors)."""
    if name == "embed_tokens.weight":
        return "emb"
    if name.startswith("layers."):
        _, l, rest = name.split(".", 2)
        for tag, parts in (("qkv", ("attention.q_proj", "attention.k_proj", "attention.v_proj")),
                           ("o", ("attention.o_proj",)), ("gu", ("mlp.gate_proj", "mlp.up_proj")),
                           ("d", ("mlp.down_proj",))):
            if any(rest == p + ".weight" for p in parts):
                return f"L{l}.{tag}"
    return ""


class OwnershipPlan:
    def __init__(self, slots: Sequence[ParamSlot], world: int):
        self.world = world
        self.groups: Dict[str, SlabGroup] = {}
        self.smalls: List[SmallTensor] = []
        order: List[str] = []
        for s in sorted(slots, key=lambda t: t.offset):
            key = group_key(s.name) if len(s.shape) == 2 else ""
            if not key:
                padded = (s.numel + ALIGN - 1) // ALIGN * ALIGN
                self.smalls.append(SmallTensor(s.name, s.offset, padded, len(self.smalls) % world))
                continue
            g = self.groups.get(key)
            if g is None:
                self.groups[key] = SlabGroup(key, s.offset, s.shape[0], s.shape[1], len(order) % world, (s.name,))
                order.append(key)
            else:
                if g.offset + g.rows * g.cols != s.offset or g.cols != s.shape[1]:
                    raise ValueError(f"{s.name} does not continue group {key} (padding or shape mismatch)")
                g.rows += s.shape[0]
                g.names = g.names + (s.name,)
        self.order = order

    def ranges(self, rank: int) -> List[Tuple[int, int]]:
        """Flat element ranges owned by ``rank``, ascending, non-empty."""
        out = [g.flat_range(rank, self.world) for g in self.groups.values()]
        out += [(t.offset, t.offset + t.numel) for t in self.smalls if t.owner == rank]
        return sorted(r for r in out if r[1] > r[0])

    def all_ranges(self) -> List[List[Tuple[int, int]]]:
        return [self.ranges(r) for r in range(self.world)]

    def small_ranges(self, rank: int) -> List[Tuple[int, int]]:
        return [(t.offset, t.offset + t.numel) for t in self.smalls if t.owner == rank]
