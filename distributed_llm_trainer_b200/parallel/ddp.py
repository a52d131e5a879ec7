"""Data parallelism.

Two implementations of what the reference gets from ``torch.nn.parallel.DistributedDataParallel``
(ddp_trainer.py:167-172, ``no_sync`` :329-332):

``NVLinkGradSync`` (CUDA / bf16 engine path)
    Gradients live in a flat fp32 buffer in peer-mapped memory.  After the last micro-step each rank
    *reduces only its 1/N slice of every unit* by loading the peers' gradients straight over NVLink
    (``b200_reduce_scatter_f32``, squared-norm partial folded in), the N partial norms are summed
    through the signal pads, and one fused kernel applies the clip coefficient + AdamW to the slice
    and stores the updated bf16 parameters into **every** rank's shadow buffer
    (``b200_adamw_bcast``).  That is the DDP all-reduce + ``clip_grad_norm_`` + ``optimizer.step()``
    of the reference in two passes over 1/N of the data, moving 4 B/param in and 2 B/param out per
    rank instead of 2 x 4 B/param, with optimizer state sharded (ZeRO-1) for free.  Numerically it is
    the same update on every rank (fp32 sum in a fixed rank order, so all replicas stay bit-identical).

``BucketedDDP`` (PyTorch path: CPU/gloo tests, fp32/fp16 runs)
    Our own bucketed all-reduce: per-unit flat buckets filled by post-accumulate-grad hooks in
    backward order, asynchronous ``dist.all_reduce`` per full bucket, ``no_sync()`` for accumulation
    micro-steps, rank-0 parameter broadcast at wrap time.  No RoPE buffer broadcast: the
    tables are deterministic.
"""
from __future__ import annotations

from contextlib import contextmanager
import os
from typing import Any, Dict, List, Optional, Tuple

import torch
import torch.distributed as dist
import torch.nn as nn

from .flat import ALIGN, FlatParamStore, plan_layout


# ================================================================================================ torch path
class BucketedDDP(nn.Module):
    def __init__(self, module: nn.Module, group=None, broadcast_init: bool = True):
        super().__init__()
        self.module = module
        self.group = group
        self.world = dist.get_world_size(group)
        named = [(n, p) for n, p in module.named_parameters() if p.requires_grad]
        slots, units, total = plan_layout(named)
        dev = named[0][1].device
        self._flat = torch.zeros(total, dtype=torch.float32, device=dev)
        self._slots = slots
        self._params = [p for _, p in named]
        # buckets = units, launched in the order backward completes them (last block first, root last)
        self._buckets = units
        self._pending = [0] * len(units)
        self._works: List[Any] = []
        self._sync = True
        if broadcast_init:
            with torch.no_grad():
                for p in module.parameters():
                    dist.broadcast(p.data, src=0, group=group)
        for p, s in zip(self._params, slots):
            p.register_post_accumulate_grad_hook(self._make_hook(s))

    def _make_hook(self, slot):
        def hook(p: torch.Tensor) -> None:
            if not self._sync:
                return
            self._flat[slot.offset: slot.offset + slot.numel].copy_(p.grad.reshape(-1))
            self._pending[slot.unit] += 1
            if self._pending[slot.unit] == len(self._buckets[slot.unit].slots):
                u = self._buckets[slot.unit]
                self._works.append(dist.all_reduce(self._flat[u.lo:u.hi], group=self.group, async_op=True))
        return hook

    @contextmanager
    def no_sync(self):
        prev, self._sync = self._sync, False
        try:
            yield
        finally:
            self._sync = prev

    def finish_grad_sync(self) -> None:
        """Wait for the bucket all-reduces and write the averaged gradients back."""
        for w in self._works:
            w.wait()
        self._works = []
        inv = 1.0 / self.world
        with torch.no_grad():
            for p, s in zip(self._params, self._slots):
                if p.grad is not None and self._pending[s.unit] > 0:
                    p.grad.copy_(self._flat[s.offset: s.offset + s.numel].view(s.shape) * inv)
        self._pending = [0] * len(self._buckets)

    def forward(self, *a, **k):
        return self.module(*a, **k)


# ================================================================================================ b200 path
def shard_ranges(units, world: int) -> List[List[Tuple[int, int]]]:
    """Per rank, the ``[lo, hi)`` slice it owns of every unit (ALIGN-multiple parts, last may be short)."""
    out: List[List[Tuple[int, int]]] = [[] for _ in range(world)]
    for u in units:
        n = u.hi - u.lo
        part = ((n + world - 1) // world + ALIGN - 1) // ALIGN * ALIGN
        for r in range(world):
            a = min(u.hi, u.lo + r * part)
            b = min(u.hi, u.lo + (r + 1) * part)
            out[r].append((a, b))
    return out


class NVLinkGradSync:
    def __init__(self, engine, rank: int, world: int, grad_buf, shadow_buf, pads):
        self.engine = engine
        self.store: FlatParamStore = engine.store
        self.rank, self.world = rank, world
        self.grad_buf, self.shadow_buf, self.pads = grad_buf, shadow_buf, pads
        self.all_ranges = shard_ranges(self.store.units, world)
        self.ranges = self.all_ranges[rank]
        dev = self.store.device
        self.sumsq = torch.zeros(8, dtype=torch.float32, device=dev)
        self.sumsq_total = torch.zeros(8, dtype=torch.float32, device=dev)
        self.optimizer = None
        self.comm_stream = torch.cuda.Stream(device=dev)
        self.max_ctas = 0
        # B200_DDP_OVERLAP=1 forks per-unit reductions under the last backward (captured in the CUDA graph of
        # that micro-step).  Default off: measured on 2xB200 with the small model the whole reduction is ~0.5 ms
        # of a 52 ms step, and running it under the backward GEMMs costs them more (52.8 ms with overlap and 32
        # comm CTAs, 62.4 ms with 4, vs 51.7 ms reducing after the backward) -- profiles/ddp_overlap_ab.md.
        self.overlap = os.environ.get("B200_DDP_OVERLAP", "0") == "1"
        # small grid: shares the SMs with the backward GEMMs it overlaps
        self.comm_ctas = int(os.environ.get("B200_DDP_COMM_CTAS", "32"))
        # B200_DDP_CE=1: peers' copies of the owned slice come over by copy engine into local scratch and the
        # reduction kernel reads local memory only (no SM time spent waiting on NVLink loads)
        self.ce = os.environ.get("B200_DDP_CE", "0") == "1"
        self._ce_tmp: Optional[List[torch.Tensor]] = None
        self._last = False
        self._overlapped = False
        self._units_done = 0

    def _reduce_range(self, lo: int, hi: int, ctas: int) -> None:
        """store.grad[lo:hi] = mean over ranks of their grad[lo:hi]; adds the slice's squared norm to sumsq[0]."""
        from ..ops import kernels as K

        if hi <= lo:
            return
        ptrs = self.grad_buf.ptrs
        if self.ce and self.world > 1:
            if self._ce_tmp is None:
                mx = max(b - a for a, b in self.ranges)
                self._ce_tmp = [torch.empty(mx, dtype=torch.float32, device=self.store.device)
                                for _ in range(self.world - 1)]
            tab, k = [], 0
            for r in range(self.world):
                if r == self.rank:
                    tab.append(self.grad_buf.raw_ptrs[r])
                else:
                    t = self._ce_tmp[k]
                    k += 1
                    K.memcpy_async(t.data_ptr(), self.grad_buf.raw_ptrs[r] + 4 * lo, 4 * (hi - lo))
                    tab.append(t.data_ptr() - 4 * lo)          # the kernel indexes every source at [lo, hi)
            ptrs = K.ptr_array(tab)
        K.reduce_scatter_f32(ptrs, self.store.grad, lo, hi, 1.0 / self.world, self.sumsq[0:1], self.rank, self.world,
                             ctas)

    def owned_ranges(self) -> List[Tuple[int, int]]:
        return self.ranges

    def attach_optimizer(self, opt) -> None:
        self.optimizer = opt

    def begin_micro_step(self, last: bool) -> None:
        """Before each micro-step.  On the last one (the only one whose gradients are final -- the
        ``no_sync`` rule of the reference, ddp_trainer.py:329-332) the engine reports every unit as soon as
        its backward is done, and that unit's slice reduction is forked onto the comm stream so it overlaps
        the rest of the backward pass (reference: DDP bucket hooks)."""
        self._last = last
        self._overlapped = False
        if last and self.overlap:
            self.sumsq.zero_()
            self._units_done = 0
            self.engine.grad_ready_hook = self._on_unit_ready
            self.engine.end_step_hook = self._join_comm
            self.engine.hooks_capturable = True
            self._overlapped = True

    def _on_unit_ready(self, unit: int) -> None:
        from ..ops import kernels as K

        cur = torch.cuda.current_stream()
        ev = torch.cuda.Event()
        ev.record(cur)
        self.comm_stream.wait_event(ev)
        lo, hi = self.ranges[unit]
        with torch.cuda.stream(self.comm_stream):
            self.pads.barrier()      # every rank has finished this unit's gradients
            self._reduce_range(lo, hi, self.comm_ctas)
        self._units_done += 1

    def _join_comm(self) -> None:
        """End of the micro-step: the compute stream waits for the forked reductions (inside a CUDA graph
        capture this is the join that closes the fork)."""
        torch.cuda.current_stream().wait_stream(self.comm_stream)

    # ---- the fused step --------------------------------------------------------------------------------
    def finish_and_step(self, grad_clip: float, norm_out: Optional[torch.Tensor] = None) -> None:
        from ..ops import kernels as K

        st, opt = self.store, self.optimizer
        g0 = opt.param_groups[0]
        lr, (b1, b2), eps, wd = g0["lr"], g0["betas"], g0["eps"], g0["weight_decay"]
        opt.step_count += 1
        if self._overlapped:
            # reductions were forked under the backward and already joined by end_step_hook
            self.engine.grad_ready_hook = None
            self.engine.end_step_hook = None
        else:
            self.sumsq.zero_()
            # every rank's backward has finished writing its gradients
            self.pads.barrier()
            for lo, hi in self.ranges:
                self._reduce_range(lo, hi, self.max_ctas)
        # global gradient norm: sum of the per-rank partials (identical summation order everywhere)
        self.pads.allreduce_scalars(self.sumsq, self.sumsq_total, 1)
        K.clip_coef(self.sumsq_total, float(grad_clip), 1.0, opt.grad_scale, norm_out)
        for (lo, hi), off in zip(self.ranges, opt.state_offsets):
            if hi > lo:
                K.adamw_bcast(st.master, st.grad, opt.exp_avg[off: off + hi - lo], opt.exp_avg_sq[off: off + hi - lo],
                              self.shadow_buf.ptrs, self.rank, self.world, opt._decay, lo, hi, opt.step_count, lr, b1,
                              b2, eps, wd, opt.grad_scale, self.max_ctas)
        # all peers have read my gradients (safe to zero) and all shadow stores have landed
        self.pads.barrier()
        st.grad.zero_()

    # ---- checkpoint support (cold path, NCCL) ----------------------------------------------------------------
    def _gather_flat(self, local_pieces: List[torch.Tensor], out_full: torch.Tensor) -> None:
        sizes = [sum(b - a for a, b in rr) for rr in self.all_ranges]
        mx = max(sizes)
        mine = torch.zeros(mx, dtype=out_full.dtype, device=out_full.device)
        cur = 0
        for t in local_pieces:
            mine[cur: cur + t.numel()].copy_(t)
            cur += t.numel()
        gathered = [torch.empty_like(mine) for _ in range(self.world)]
        dist.all_gather(gathered, mine)
        for r, rr in enumerate(self.all_ranges):
            cur = 0
            for a, b in rr:
                out_full[a:b].copy_(gathered[r][cur: cur + b - a])
                cur += b - a

    def gather_master_and_state(self) -> None:
        """fp32 master weights are only current inside each rank's owned slices; make them current everywhere."""
        st = self.store
        self._gather_flat([st.master[a:b] for a, b in self.ranges], st.master)

    def full_optimizer_state_dict(self) -> Dict[str, Any]:
        st, opt = self.store, self.optimizer
        full_m = torch.zeros(st.total, dtype=torch.float32, device=st.device)
        full_v = torch.zeros(st.total, dtype=torch.float32, device=st.device)
        self._gather_flat([opt.exp_avg[o: o + b - a] for (a, b), o in zip(opt.ranges, opt.state_offsets)], full_m)
        self._gather_flat([opt.exp_avg_sq[o: o + b - a] for (a, b), o in zip(opt.ranges, opt.state_offsets)], full_v)
        return opt.state_dict(full_m.cpu(), full_v.cpu())

    def load_full_optimizer_state_dict(self, sd: Dict[str, Any]) -> None:
        self.optimizer.load_state_dict(sd)   # keeps only the owned ranges
        with torch.no_grad():
            self.shadow_buf.local[: self.store.total].copy_(self.store.master)


def build_engine_for_ddp(model, distributed: bool, rank: int, world_size: int, seed: int = 0):
    """Create the fused engine; with >1 rank its gradient and shadow buffers are peer-mapped."""
    from ..ops.engine import FusedGPTEngine

    dev = next(model.parameters()).device
    if not distributed or world_size == 1:
        eng = FusedGPTEngine(model, seed=seed)
        return eng, None
    from .comm import SignalPads, SymmetricBuffer

    named = [(n, p) for n, p in model.named_parameters() if p.requires_grad]
    _, _, total = plan_layout(named)
    grad_buf = SymmetricBuffer(total, torch.float32, dev)
    shadow_buf = SymmetricBuffer(total, torch.bfloat16, dev)
    pads = SignalPads(dev)
    store = FlatParamStore(model, device=dev, grad_buffer=grad_buf.local, shadow_buffer=shadow_buf.local)
    # replicas must start identical (reference: DDP broadcasts rank 0's parameters)
    dist.broadcast(store.master, src=0)
    store.refresh_shadow()
    eng = FusedGPTEngine(model, store=store, seed=seed)
    sync = NVLinkGradSync(eng, rank, world_size, grad_buf, shadow_buf, pads)
    # first launches of the comm kernels happen here, eagerly, not inside a later CUDA-graph capture
    from ..ops import kernels as K

    pads.barrier()
    K.reduce_scatter_f32(grad_buf.ptrs, store.grad, 0, 128, 1.0, sync.sumsq[0:1], rank, world_size, 1)
    pads.barrier()
    torch.cuda.synchronize()
    dist.barrier()
    return eng, sync
