"""Data parallelism.

Two implementations of what the reference gets from ``torch.nn.parallel.DistributedDataParallel``
(ddp_trainer.py:167-172, ``no_sync`` :329-332):

``NVLinkGradSync`` (CUDA / bf16 engine path)
    Gradients live in flat fp32 buffers in peer-mapped memory and every weight matrix is owned by ranks
    in 128-row slabs (``parallel/ownership.py``).  The wgrad GEMMs deliver their tiles to the owners
    themselves -- a TMA reduce-add over NVLink from the epilogue (``b200_gemm2_rs_bf16``) -- so when the
    last backward finishes the reduce-scatter has already happened, tile by tile, under the GEMM main
    loops.  The step that follows is a barrier, the squared norm of the owned rows, a scalar all-reduce
    through the signal pads and one AdamW launch whose bf16 result is stored into **every** rank's
    compute copy (``b200_adamw_chunks``): the DDP all-reduce + ``clip_grad_norm_`` +
    ``optimizer.step()`` of the reference with 4 B/param of gradient out (hidden) and 2 B/param of
    parameters out per rank, optimizer work and state / N (ZeRO-1), and replicas that stay bit-identical
    because every rank receives the same bf16 values from the one owner that computed them.

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
    """Gradient synchronisation + sharded optimizer of the b200 DDP path.

    The gradient all-reduce of the reference (DDP buckets fired from backward, ddp_trainer.py:329-342) does not
    exist as a separate step here: every wgrad GEMM (and the embedding scatter) of every micro-step reduce-adds
    its 128-row output slabs straight into the buffer of the rank that OWNS those rows
    (``b200_gemm2_rs_bf16``: TMA reduce-add over NVLink from the GEMM's epilogue, under the main loop of the
    next tiles).  When the last backward is done each rank already holds the cross-rank sum of its rows, and
    the step is: barrier -> (owners pull the few norm-weight gradients) -> squared norm of the owned rows ->
    scalar all-reduce -> clip coefficient -> ONE AdamW launch over the owned rows that stores the new bf16
    parameters into every rank's compute copy and clears the gradient -> barrier.
    """

    def __init__(self, engine, rank: int, world: int, grad_buf, shadow_buf, pads, red_buf):
        from ..ops import kernels as K
        from .ownership import OwnershipPlan

        self.engine = engine
        self.store: FlatParamStore = engine.store
        self.rank, self.world = rank, world
        # grad_buf: this rank's own accumulator (micro-steps before the last add into it locally; peer-mapped only so
        # that owners can pull the few norm-weight gradients).  red_buf: where the owners' rows receive the
        # cross-rank sums (reduce-adds from every rank's last wgrad GEMMs); read and cleared by the AdamW launch.
        self.grad_buf, self.shadow_buf, self.pads, self.red_buf = grad_buf, shadow_buf, pads, red_buf
        self.plan = OwnershipPlan(self.store.slots, world)
        self.all_ranges = self.plan.all_ranges()
        self.ranges = self.all_ranges[rank]
        dev = self.store.device
        self.sumsq = torch.zeros(8, dtype=torch.float32, device=dev)
        self.sumsq_total = torch.zeros(8, dtype=torch.float32, device=dev)
        self.optimizer = None
        self.max_ctas = 0
        # key -> (RsArgs, keep-alive): base pointer of every owner's rows inside its (peer-mapped) gradient buffer
        self._rs: Dict[Tuple[str, bool], tuple] = {}
        for key, g in self.plan.groups.items():
            bases = []
            for o in range(world):
                a, b = g.row_range(o, world)
                bases.append(red_buf.raw_ptrs[o] + 4 * (g.offset + a * g.cols) if b > a else None)
            self._rs[(key, False)] = K.make_rs_args(bases, g.rot)
            self._rs[(key, True)] = K.make_rs_args(bases, g.rot, prev_ptr=grad_buf.raw_ptrs[rank] + 4 * g.offset)
        engine.rs_router = lambda key, prev: self._rs.get((key, bool(prev)))
        self._chunks = None

    def owned_ranges(self) -> List[Tuple[int, int]]:
        return self.ranges

    def close(self) -> None:
        """Collective: release the peer-mapped buffers (the engine and its store must not be used afterwards)."""
        self.engine.rs_router = None
        for b in (self.grad_buf, self.red_buf, self.shadow_buf, self.pads.buf):
            b.close()

    def attach_optimizer(self, opt) -> None:
        """Builds the chunk tables (they carry the optimizer-state offsets of the owned runs)."""
        from ..ops import kernels as K

        self.optimizer = opt
        dev = self.store.device
        assert opt.ranges == [(int(a), int(b)) for a, b in self.ranges]
        runs = [(lo, lo, off, lo, hi - lo, -1) for (lo, hi), off in zip(opt.ranges, opt.state_offsets)]
        self._chunks = K.make_chunk_table(runs, dev)
        small = [(lo, lo, 0, lo, hi - lo, -1) for lo, hi in self.plan.small_ranges(self.rank)]
        self._pull_chunks = K.make_chunk_table(small, dev)

    def begin_micro_step(self, index: int, count: int) -> None:
        """Micro-step ``index`` of ``count``: the first overwrites the local accumulator (no zero-fill pass), the
        middle ones add to it, and the LAST one's wgrad GEMMs send accumulator + their own product to the owners --
        the reference's ``no_sync()`` rule (ddp_trainer.py:329-332): gradients cross the wire once per step."""
        if os.environ.get("B200_DDP_DEBUG_NO_RS") == "1":   # timing attribution only (wrong gradients): never leave the GPU
            self.engine.grad_mode = "add"
            return
        self.engine.grad_mode = "rs1" if count == 1 else ("store" if index == 0 else ("rs" if index == count - 1 else "add"))

    # ---- the fused step --------------------------------------------------------------------------------
    def finish_and_step(self, grad_clip: float, norm_out: Optional[torch.Tensor] = None) -> None:
        from ..ops import kernels as K

        st, opt = self.store, self.optimizer
        g0 = opt.param_groups[0]
        lr, (b1, b2), eps, wd = g0["lr"], g0["betas"], g0["eps"], g0["weight_decay"]
        opt.step_count += 1
        mark = getattr(self, "trace", None) or (lambda name: None)   # benchmarks/bench_ddp_breakdown.py: event per phase
        self.sumsq.zero_()
        mark("start")
        # every rank's reduce-adds (the last micro-step's wgrad GEMMs) have landed in the owners' buffers
        self.pads.barrier()
        mark("barrier")
        table, n = self._pull_chunks
        red = self.red_buf.local
        K.pull_reduce_chunks(self.grad_buf.ptrs, red, table, n, self.rank, self.world, zero_remote=True,
                             multicast_ptr=getattr(self.grad_buf, "mc_ptr", 0))
        table, n = self._chunks
        mark("pull")
        K.sumsq_chunks(red, table, n, self.sumsq[0:1], 1.0)
        mark("sumsq")
        # global gradient norm: sum of the per-rank partials (identical summation order everywhere); the buffers
        # hold SUMS over ranks, the 1/world of the mean is folded into the clip coefficient
        self.pads.allreduce_scalars(self.sumsq, self.sumsq_total, 1)
        K.clip_coef(self.sumsq_total, float(grad_clip), 1.0 / self.world, opt.grad_scale, norm_out)
        mark("allreduce+clip")
        K.adamw_chunks(st.master, red, opt.exp_avg, opt.exp_avg_sq, self.shadow_buf.ptrs, self.rank, self.world,
                       opt._decay, table, n, opt.step_count, lr, b1, b2, eps, wd, opt.grad_scale, zero_grad=True,
                       max_ctas=self.max_ctas, multicast_ptr=getattr(self.shadow_buf, "mc_ptr", 0))
        mark("adamw")
        # all parameter stores have landed and every owned gradient is zero again: the next step may start
        self.pads.barrier()
        mark("barrier2")

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


def symmetric_or_multicast(numel: int, dtype, dev):
    """Peer-mapped buffer with an NVLS multicast mapping when the box offers one (``mc_ptr`` != 0: broadcasts and
    sums over ranks become single ``multimem`` instructions executed by the NVSwitch), plain CUDA-IPC mapping
    otherwise (``B200_NVLS=0`` forces it).  Every rank must take the same branch, so the outcome is agreed on."""
    from .comm import MulticastBuffer, SymmetricBuffer

    ok = MulticastBuffer.available()
    buf = None
    if ok:
        try:
            buf = MulticastBuffer(numel, dtype, dev)
        except Exception:      # no multicast support for this process model: every rank fails the same way
            buf = None
    flag = torch.tensor([1 if buf is not None else 0], device=dev)
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    if int(flag.item()) == 1:
        return buf
    if buf is not None:
        buf.close()
    return SymmetricBuffer(numel, dtype, dev)


def build_engine_for_ddp(model, distributed: bool, rank: int, world_size: int, seed: int = 0):
    """Create the fused engine; with >1 rank its gradient and shadow buffers are peer-mapped."""
    from ..ops.engine import FusedGPTEngine

    dev = next(model.parameters()).device
    if not distributed or world_size == 1:
        eng = FusedGPTEngine(model, seed=seed)
        return eng, None
    from .comm import SignalPads, SymmetricBuffer

    if world_size > 8:
        raise NotImplementedError("the NVLink engines address the ranks of ONE NVSwitch domain (<= 8 GPUs): the wgrad "
                                  f"GEMM carries one tensor map per owner; got world_size={world_size}")
    named = [(n, p) for n, p in model.named_parameters() if p.requires_grad]
    _, _, total = plan_layout(named)
    grad_buf, shadow_buf = symmetric_or_multicast(total, torch.float32, dev), symmetric_or_multicast(total, torch.bfloat16, dev)
    red_buf = SymmetricBuffer(total, torch.float32, dev)
    pads = SignalPads(dev)
    store = FlatParamStore(model, device=dev, grad_buffer=grad_buf.local, shadow_buffer=shadow_buf.local)
    # replicas must start identical (reference: DDP broadcasts rank 0's parameters)
    dist.broadcast(store.master, src=0)
    store.refresh_shadow()
    eng = FusedGPTEngine(model, store=store, seed=seed)
    sync = NVLinkGradSync(eng, rank, world_size, grad_buf, shadow_buf, pads, red_buf)
    # first launch of the pad kernels happens here, eagerly, not inside a later CUDA-graph capture
    pads.barrier()
    torch.cuda.synchronize()
    dist.barrier()
    return eng, sync
