"""Fully-sharded data parallelism (ZeRO-3 / ZeRO-2 / no-shard) on the flat-unit layout.

Parity target: what the reference gets from torch FSDP1 (fsdp_trainer.py:158-332): one unit per
``TransformerBlock`` plus a root unit (wrap policy :183-186), strategies FULL_SHARD / SHARD_GRAD_OP /
NO_SHARD (:269-274), bf16 parameters + communication with fp32 sharded master weights
(``MixedPrecision`` :223-234), backward prefetch, CPU offload flag, ``clip_grad_norm_`` across shards
(:388) and FULL_STATE_DICT checkpoints (:444-465).

``ShardedGPT`` (CUDA / bf16 engine path)
    * every rank keeps only its 1/N slice of each unit in bf16 (peer-mapped) + fp32 optimizer state;
    * forward/backward walk the units: the peers' slices are copied over NVLink (copy engines,
      ``b200_memcpy_async``; optionally the ``b200_all_gather_bf16`` pull kernel) into a pooled unit
      buffer on a gather stream (next unit prefetched while the current one computes; FULL_SHARD
      frees it after use and gathers again in backward, SHARD_GRAD_OP keeps it);
    * wgrad GEMMs write fp32 gradients into peer-mapped per-block buffers that accumulate over the
      micro-steps of a step; on the last micro-step each rank copies the peers' versions of *its*
      slice over and sums them into the sharded fp32 gradient (``b200_reduce_scatter_f32_accum``).
      ``B200_FSDP_ACCUM_LOCAL=0`` reduces every micro-step through two rotating staging buffers
      instead, as FSDP1 does (the reference has no ``no_sync``);
    * the step is: sharded norm^2 -> pad all-reduce -> clip coefficient -> fused AdamW on the shard,
      writing the bf16 slice peers will gather next step.

``TorchShardedModel`` (PyTorch path: CPU/gloo tests)
    same sharding arithmetic and optimizer on top of autograd + ``torch.distributed`` collectives.
"""
from __future__ import annotations

import os
from typing import Any, Dict, List, Optional, Tuple

import torch
import torch.distributed as dist
import torch.nn as nn

from .ddp import shard_ranges
from .flat import ALIGN, FlatParamStore


def _gather_ranges(all_ranges, world, local_pieces: List[torch.Tensor], out_full: torch.Tensor, group=None) -> None:
    """all-gather of ragged per-rank slices into a full flat buffer (cold path / torch path)."""
    if world == 1:
        for (a, b), t in zip(all_ranges[0], local_pieces):
            if out_full[a:b].data_ptr() != t.data_ptr():
                out_full[a:b].copy_(t.reshape(-1))
        return
    sizes = [sum(b - a for a, b in rr) for rr in all_ranges]
    mx = max(sizes)
    mine = torch.zeros(mx, dtype=out_full.dtype, device=out_full.device)
    cur = 0
    for t in local_pieces:
        mine[cur: cur + t.numel()].copy_(t.reshape(-1))
        cur += t.numel()
    gathered = [torch.empty_like(mine) for _ in range(world)]
    dist.all_gather(gathered, mine, group=group)
    for r, rr in enumerate(all_ranges):
        cur = 0
        for a, b in rr:
            out_full[a:b].copy_(gathered[r][cur: cur + b - a])
            cur += b - a


# ================================================================================================ torch path
class TorchShardedModel:
    """Sharded optimizer + gradient reduction around a plain PyTorch ``GPT`` (autograd computes full local
    gradients into the flat buffer; communication goes through ``torch.distributed``)."""

    def __init__(self, model: nn.Module, strategy: str, rank: int, world: int, group=None):
        self.model = model
        self.strategy = strategy
        self.rank, self.world, self.group = rank, world, group
        self.store = FlatParamStore(model, with_shadow=False)
        st = self.store
        if world > 1:
            dist.broadcast(st.master, src=0, group=group)   # identical start (FSDP sync_module_states=True)
        self.all_ranges = shard_ranges(st.units, world) if strategy != "NO_SHARD" else [[(0, st.total)]] * world
        self.ranges = self.all_ranges[rank]

    def owned_ranges(self):
        return self.ranges

    def reduce_gradients(self) -> None:
        """Average gradients across ranks (reduce-scatter semantics: only owned slices are used afterwards)."""
        st = self.store
        if self.world > 1:
            for u in st.units:
                dist.all_reduce(st.grad[u.lo:u.hi], group=self.group)
            st.grad.div_(self.world)

    def grad_sumsq(self) -> torch.Tensor:
        st = self.store
        acc = torch.zeros(1, dtype=torch.float32, device=st.device)
        for a, b in self.ranges:
            acc += st.grad[a:b].float().pow(2).sum()
        if self.world > 1 and self.strategy != "NO_SHARD":
            dist.all_reduce(acc, group=self.group)
        return acc

    def gather_params(self) -> None:
        """After the sharded step: make the full fp32 parameters current on every rank (all-gather)."""
        if self.world > 1 and self.strategy != "NO_SHARD":
            st = self.store
            _gather_ranges(self.all_ranges, self.world, [st.master[a:b] for a, b in self.ranges], st.master, self.group)

    def full_optimizer_state(self, opt) -> Tuple[torch.Tensor, torch.Tensor]:
        st = self.store
        if opt.is_full:
            return opt.exp_avg, opt.exp_avg_sq
        fm = torch.zeros(st.total, dtype=torch.float32, device=st.device)
        fv = torch.zeros_like(fm)
        pieces_m = [opt.exp_avg[o: o + b - a] for (a, b), o in zip(opt.ranges, opt.state_offsets)]
        pieces_v = [opt.exp_avg_sq[o: o + b - a] for (a, b), o in zip(opt.ranges, opt.state_offsets)]
        _gather_ranges(self.all_ranges, self.world, pieces_m, fm, self.group)
        _gather_ranges(self.all_ranges, self.world, pieces_v, fv, self.group)
        return fm, fv


# ================================================================================================ b200 path
class ShardedGPT:
    """FSDP engine over ``FusedGPTEngine`` hooks.  See the module docstring for the data flow."""

    def __init__(self, model, strategy: str, rank: int, world: int, seed: int = 0, cpu_offload: bool = False,
                 limit_all_gathers: bool = True, backward_prefetch: bool = True):
        from ..ops.engine import FusedGPTEngine, LayerWeights
        from .comm import SignalPads, SymmetricBuffer

        assert strategy in ("FULL_SHARD", "SHARD_GRAD_OP")
        self.strategy, self.rank, self.world = strategy, rank, world
        self.cpu_offload = cpu_offload
        dev = next(model.parameters()).device
        self.device = dev
        # flat layout; the full-size fp32 buffer only lives until the shards are cut (release_full_master below)
        self.store = FlatParamStore(model, device=dev, with_grad=False, with_shadow=False)
        st = self.store
        dist.broadcast(st.master, src=0)
        self.units = st.units
        self.L = len(self.units) - 1
        self.all_ranges = shard_ranges(self.units, world)
        self.ranges = self.all_ranges[rank]
        # per-unit shard geometry
        self.unit_part: List[int] = []
        self.shard_off: List[int] = []     # offset of unit u's slice inside this rank's concatenated shard
        cur = 0
        for u in self.units:
            n = u.hi - u.lo
            part = ((n + world - 1) // world + ALIGN - 1) // ALIGN * ALIGN
            self.unit_part.append(part)
            self.shard_off.append(cur)
            cur += part
        self.shard_total = cur
        self.pads = SignalPads(dev)
        # bf16 shard (peers pull from it) and sharded fp32 gradient accumulator
        self.shard_bf16 = SymmetricBuffer(self.shard_total, torch.bfloat16, dev)
        self.grad_shard = torch.zeros(self.shard_total, dtype=torch.float32, device=dev)
        self.master_shard = torch.zeros(self.shard_total, dtype=torch.float32, device=dev)
        self.keep_full_master = os.environ.get("B200_FSDP_FULL_MASTER", "0") == "1"
        self._refresh_shard_from_master()
        self.release_full_master()
        # pools
        blk = max(self.unit_part[1:]) * world if self.L > 0 else ALIGN
        root = self.unit_part[0] * world
        n_gather = self.L if strategy == "SHARD_GRAD_OP" else min(self.L, 3)
        slack = 64 * model.config.hidden_size   # the LM-head GEMM's last B tile may read past V rows
        self.root_w = torch.zeros(root + slack, dtype=torch.bfloat16, device=dev)
        self.gather_pool = [torch.zeros(blk, dtype=torch.bfloat16, device=dev) for _ in range(max(n_gather, 1))]
        self.root_g = SymmetricBuffer(root, torch.float32, dev)
        # Gradient accumulation across micro-steps.  accum_local (default when it fits): every block unit has its own
        # peer-mapped fp32 buffer that accumulates over the micro-steps of a step and is reduce-scattered ONCE, on
        # the last one -- 180 GB of HBM makes the reference's "reduce-scatter every micro-step" (FSDP1 without
        # no_sync, fsdp_trainer.py:369-384) an avoidable cost: GA x less traffic and GA x fewer barriers.
        # B200_FSDP_ACCUM_LOCAL=0 restores the per-micro-step reduction through two rotating staging buffers
        # (gradient memory / world, the reference's footprint).
        full_bytes = 4 * blk * self.L
        free_b, _tot = torch.cuda.mem_get_info(dev)
        self.accum_local = (os.environ.get("B200_FSDP_ACCUM_LOCAL", "1") != "0" and self.L > 0
                            and full_bytes < 0.25 * free_b)
        n_stage = self.L if self.accum_local else 2
        self.grad_pool = [SymmetricBuffer(blk, torch.float32, dev) for _ in range(n_stage)]
        self.comm_stream = torch.cuda.Stream(device=dev)     # gradient reductions (barrier / copy / sum / barrier)
        # Parameter gathers get their own stream: queued behind a reduction they would wait for that reduction's
        # barrier, i.e. for the slowest peer's backward, and stall the next block's weights_hook.
        self.gather_stream = torch.cuda.Stream(device=dev)
        self._params_ready: Optional[torch.cuda.Event] = None   # recorded after the optimizer step's barrier
        self.max_ctas = 0
        # all-gather fused into the QKV GEMM (B200_FSDP_FUSE_AG=0 gathers those rows with the rest of the unit)
        self._unit_ptr_cache: Dict[int, Any] = {}
        h = model.config.hidden_size
        self.qkv_len = 3 * h * h
        self.qkv_off = (st.by_name["layers.0.attention.q_proj.weight"].offset - self.units[1].lo) if self.L > 0 else 0
        # Unit all-gathers ride the copy engines by default (B200_FSDP_CE_GATHER=0: the P2P-pull SM kernel).
        self.ce_gather = os.environ.get("B200_FSDP_CE_GATHER", "1") != "0"
        # Same for the gradient reduce-scatter: copy-engine transfer + local sum (B200_FSDP_CE_RS=0: P2P-pull kernel)
        self.ce_rs = os.environ.get("B200_FSDP_CE_RS", "1") != "0"
        self._rs_tmp: Dict[int, List[torch.Tensor]] = {}
        self.use_graphs = (os.environ.get("B200_FSDP_GRAPHS", "1") != "0"
                           and os.environ.get("B200_NO_GRAPHS", "0") != "1")
        self._graphs: Dict[Tuple[int, int, float], Any] = {}
        self._plain_calls = 0
        # The fused all-gather GEMM is opt-in (B200_FSDP_FUSE_AG=1): a PREFETCHED gather hides completely under
        # the previous block's compute, the fused one only under its own GEMM -- measured +7 ms / step on the
        # 355M benchmark at 2 GPUs (profiles/fsdp_comm_breakdown.md).  It pays when nothing precedes the GEMM.
        self.fuse_ag = (os.environ.get("B200_FSDP_FUSE_AG", "0") != "0" and self.L > 0
                        and min(self.unit_part[1:]) >= 16384 and 3 * h >= 256)
        self.ag_pairs = int(os.environ.get("B200_FSDP_AG_PAIRS", "2"))
        self._ag_flags = [torch.zeros((3 * h + 127) // 128, dtype=torch.int32, device=dev) for _ in self.gather_pool]
        self._ag_epoch = 0
        # engine bound to hooks
        self.engine = FusedGPTEngine.__new__(FusedGPTEngine)
        self._init_engine(model, seed, LayerWeights)
        self.sumsq = torch.zeros(8, dtype=torch.float32, device=dev)
        self.sumsq_total = torch.zeros(8, dtype=torch.float32, device=dev)
        self._gathered: Dict[int, Tuple[int, torch.cuda.Event]] = {}   # layer -> (pool slot, ready event)
        self._slot_free_evt: List[Optional[torch.cuda.Event]] = [None] * len(self.gather_pool)
        self._slot_owner: List[Optional[int]] = [None] * len(self.gather_pool)
        self._gslot_free_evt: List[Optional[torch.cuda.Event]] = [None] * len(self.grad_pool)
        self._root_ready: Optional[torch.cuda.Event] = None
        self._phase = "idle"
        torch.cuda.synchronize()
        dist.barrier()

    # ------------------------------------------------------------------------------------------ setup helpers
    def _refresh_shard_from_master(self) -> None:
        """Full fp32 parameters (materialised) -> this rank's fp32 master shard and bf16 shard."""
        st = self.store
        with torch.no_grad():
            for u, (a, b) in zip(range(len(self.units)), self.ranges):
                off = self.shard_off[u]
                self.master_shard[off: off + b - a].copy_(st.master[a:b])
                self.shard_bf16.local[off: off + b - a].copy_(st.master[a:b])

    # The fp32 master weights are sharded like everything else: outside of checkpoint save / load only
    # ``master_shard`` (1/N of the model) exists and the module's parameters point at empty storage.
    # B200_FSDP_FULL_MASTER=1 keeps the full-size buffer resident (module parameters stay readable, owned
    # slices only are current).
    def release_full_master(self) -> None:
        st = self.store
        if self.keep_full_master or st.master.numel() == 0:
            return
        empty = torch.empty(0, dtype=torch.float32, device=self.device)
        for p in st.params:
            p.data = empty
        st.master = empty

    def materialize_full_master(self) -> None:
        """Re-create the full fp32 buffer with this rank's owned slices filled in (others zero) and point the
        module's parameters back at it."""
        st = self.store
        if st.master.numel() == st.total:
            if self.keep_full_master:
                for u, (a, b) in enumerate(self.ranges):
                    st.master[a:b].copy_(self.master_shard[self.shard_off[u]: self.shard_off[u] + b - a])
            return
        st.master = torch.zeros(st.total, dtype=torch.float32, device=self.device)
        for p, s in zip(st.params, st.slots):
            p.data = st.master[s.offset: s.offset + s.numel].view(s.shape)
        for u, (a, b) in enumerate(self.ranges):
            st.master[a:b].copy_(self.master_shard[self.shard_off[u]: self.shard_off[u] + b - a])

    def _views(self, buf: torch.Tensor, l: int, LayerWeights):
        """LayerWeights views into a gathered/staging unit buffer for block l (offsets relative to the unit)."""
        st = self.store
        u = self.units[1 + l]
        pre = f"layers.{l}."

        def v(name):
            s = st.by_name[pre + name]
            return buf[s.offset - u.lo: s.offset - u.lo + s.numel].view(s.shape)

        def packed(names):
            ss = [st.by_name[pre + n] for n in names]
            rows = sum(s.shape[0] for s in ss)
            cols = ss[0].shape[1]
            o = ss[0].offset - u.lo
            return buf[o: o + rows * cols].view(rows, cols)

        return LayerWeights(ln1=v("input_layernorm.weight"),
                            wqkv=packed([f"attention.{n}_proj.weight" for n in "qkv"]),
                            wo=v("attention.o_proj.weight"), ln2=v("post_attention_layernorm.weight"),
                            wgu=packed(["mlp.gate_proj.weight", "mlp.up_proj.weight"]), wd=v("mlp.down_proj.weight"))

    def _init_engine(self, model, seed, LayerWeights) -> None:
        from ..ops import kernels as K
        from ..ops.reference import rope_tables

        e = self.engine
        cfg = model.config
        e.model, e.cfg, e.store, e.device = model, cfg, self.store, self.device
        K._lib.lib()
        e.vocab = cfg.vocab_size
        e.vocab_padded = (cfg.vocab_size + 127) // 128 * 128
        e.checkpointing = bool(cfg.gradient_checkpointing)
        e.rope_cos, e.rope_sin = rope_tables(cfg.max_seq_len, 64, device=self.device)
        e.rng_state = torch.tensor([seed & 0x7FFFFFFFFFFFFFFF, 0], dtype=torch.int64, device=self.device)
        e._ws = {}
        e.use_graphs = False
        e.fuse_swiglu_bwd = True
        e.end_step_hook = None
        e.hooks_capturable = False
        e.rs_router = None
        e.max_ctas = 0
        e.wgrad_overwrite = not self.accum_local    # rotating staging buffers hold one contribution per use
        e.sm_count = torch.cuda.get_device_properties(self.device).multi_processor_count
        self._LayerWeights = LayerWeights
        st = self.store
        root_u = self.units[0]

        def rootv(buf, name):
            s = st.by_name[name]
            return buf[s.offset - root_u.lo: s.offset - root_u.lo + s.numel].view(s.shape)

        e.w_emb = rootv(self.root_w, "embed_tokens.weight")
        e.w_norm = rootv(self.root_w, "norm.weight")
        e.g_emb = rootv(self.root_g.local, "embed_tokens.weight")
        e.g_norm = rootv(self.root_g.local, "norm.weight")
        e.w_layers, e.g_layers = [], [None] * self.L
        e.ag_hook = self._ag_hook
        e.weights_hook = self._weights_hook
        e.release_hook = self._release_hook
        e.grad_ready_hook = self._grad_ready_hook

    # ------------------------------------------------------------------------------------------ gathers
    def _unit_ptrs(self, u: int):
        """Pointer table (one entry per rank) of unit u's slice inside every rank's bf16 shard."""
        tab = self._unit_ptr_cache.get(u)
        if tab is None:
            tab = self.shard_bf16.offset_ptrs(self.shard_off[u] * 2)
            self._unit_ptr_cache[u] = tab
        return tab

    def _gather_unit(self, u: int, dst: torch.Tensor, skip_qkv: bool = False) -> None:
        """All-gather of unit u.  ``skip_qkv``: leave the q|k|v rows alone -- the QKV GEMM of this block gathers
        them itself (all-gather fused into the first consuming GEMM, ``b200_gemm2_ag_bf16``)."""
        from ..ops import kernels as K

        excl = (self.qkv_off, self.qkv_off + self.qkv_len) if skip_qkv else (0, 0)
        if not self.ce_gather:
            K.all_gather_bf16(self._unit_ptrs(u), dst, self.unit_part[u], self.rank, self.world, self.max_ctas,
                              exclude=excl)
            return
        # Copy-engine all-gather: one stream-ordered P2P copy per (owner, segment).  No SM is used, so the
        # gather overlaps the persistent GEMM / attention CTAs instead of sharing their SMs (measured:
        # profiles/fsdp_comm_breakdown.md).
        part = self.unit_part[u]
        total = part * self.world
        segs = [(0, total)] if excl[1] <= excl[0] else [(0, excl[0]), (excl[1], total)]
        base = dst.data_ptr()
        src_off = self.shard_off[u]
        for r in range(self.world):
            lo, hi = r * part, (r + 1) * part
            for a, b in segs:
                a2, b2 = max(a, lo), min(b, hi)
                if b2 > a2:
                    K.memcpy_async(base + 2 * a2, self.shard_bf16.raw_ptrs[r] + 2 * (src_off + a2 - lo), 2 * (b2 - a2))

    def _fuse_qkv_now(self) -> bool:
        """Will the QKV GEMM run on the weights gathered in the current phase?  Forward: always.  Backward:
        only when the block forward is recomputed (activation checkpointing)."""
        return self.fuse_ag and (self._phase == "forward" or self.engine.checkpointing)

    def _ag_hook(self, l: int):
        """Fused all-gather arguments for block l's QKV GEMM, or None if its rows are already in place."""
        ent = self._gathered.get(l)
        if ent is None or not ent[2]:
            return None
        slot = ent[0]
        self._gathered[l] = (slot, ent[1], False)
        self._ag_epoch += 1
        return (self._unit_ptrs(1 + l), self.unit_part[1 + l], self.qkv_off, self._ag_flags[slot], self._ag_epoch,
                self.rank, self.world, self.ag_pairs)

    def _acquire_slot(self, l: int) -> int:
        for i, owner in enumerate(self._slot_owner):
            if owner is None:
                self._slot_owner[i] = l
                return i
        raise RuntimeError("gather pool exhausted")

    def _prefetch(self, l: int) -> None:
        if l < 0 or l >= self.L or l in self._gathered:
            return
        if all(o is not None for o in self._slot_owner):
            return
        slot = self._acquire_slot(l)
        cs = self.gather_stream
        if self._slot_free_evt[slot] is not None:
            cs.wait_event(self._slot_free_evt[slot])
        if self._params_ready is not None:
            cs.wait_event(self._params_ready)       # never read a peer's shard while its AdamW may be writing it
        fused = self._fuse_qkv_now()
        with torch.cuda.stream(cs):
            self._gather_unit(1 + l, self.gather_pool[slot], skip_qkv=fused)
            ev = torch.cuda.Event()
            ev.record(cs)
        self._gathered[l] = (slot, ev, fused)      # fused: the q|k|v rows are still missing

    def _weights_hook(self, l: int):
        if l not in self._gathered:
            self._prefetch(l)
        slot, ev, _ = self._gathered[l]
        torch.cuda.current_stream().wait_event(ev)
        # prefetch the unit the walk needs next
        nxt = l + 1 if self._phase == "forward" else l - 1
        self._prefetch(nxt)
        return self._views(self.gather_pool[slot], l, self._LayerWeights)

    def _free_gathered(self, l: int) -> None:
        slot = self._gathered.pop(l)[0]
        ev = torch.cuda.Event()
        ev.record(torch.cuda.current_stream())
        self._slot_free_evt[slot] = ev
        self._slot_owner[slot] = None

    # ------------------------------------------------------------------------------------------ grad staging
    def _stage_for(self, l: int):
        slot = l if self.accum_local else (l & 1)
        ev = self._gslot_free_evt[slot]
        if ev is not None:
            torch.cuda.current_stream().wait_event(ev)
        # no zero-fill: wgrad GEMMs overwrite (engine.wgrad_overwrite), the two norm gradients are cleared by
        # the caller, and the alignment padding between parameters is never written after allocation
        return slot, self.grad_pool[slot]

    def _reduce_unit(self, u: int, staging) -> torch.cuda.Event:
        """comm stream: barrier (all ranks produced unit u) -> reduce own slice -> barrier (all done reading)."""
        from ..ops import kernels as K

        cs = self.comm_stream
        done = torch.cuda.Event()
        produced = torch.cuda.Event()
        produced.record(torch.cuda.current_stream())
        cs.wait_event(produced)
        with torch.cuda.stream(cs):
            self.pads.barrier()
            part = self.unit_part[u]
            lo = self.rank * part
            n = min(part, max(0, (self.units[u].hi - self.units[u].lo) - lo))
            n = (n + 3) // 4 * 4
            dst = self.grad_shard[self.shard_off[u]: self.shard_off[u] + part]
            if n > 0 and self.ce_rs and self.world > 1:
                # Copy engines bring the peers' copies of this rank's slice next to the local one; the sum is
                # then a local streaming kernel.  The SMs never stall on NVLink loads and the transfer does not
                # occupy the SMs the persistent GEMM / attention CTAs are using.
                tmps = self._rs_tmp.get(part)      # shared by all units of this size: reductions are stream-ordered
                if tmps is None:
                    tmps = [torch.empty(part, dtype=torch.float32, device=self.device) for _ in range(self.world - 1)]
                    self._rs_tmp[part] = tmps
                tab, k = [], 0
                for r in range(self.world):
                    if r == self.rank:
                        tab.append(staging.raw_ptrs[r])
                    else:
                        K.memcpy_async(tmps[k].data_ptr(), staging.raw_ptrs[r] + 4 * lo, 4 * n)
                        tab.append(tmps[k].data_ptr() - 4 * lo)     # the kernel indexes every source at `lo`
                        k += 1
                K.reduce_scatter_f32_accum(K.ptr_array(tab), dst, lo, n, 1.0 / self.world, self.rank, self.world,
                                           self.max_ctas)
            elif n > 0:
                K.reduce_scatter_f32_accum(staging.ptrs, dst, lo, n, 1.0 / self.world, self.rank, self.world,
                                           self.max_ctas)
            self.pads.barrier()
            if self.accum_local:
                staging.local.zero_()      # every peer has read its slice: start the next step's accumulation at 0
            done.record(cs)
        return done

    def _release_hook(self, l: int, phase: str) -> None:
        if phase == "forward":
            if self.strategy == "FULL_SHARD":
                self._free_gathered(l)
            return
        # backward of block l finished: its staging buffer holds this rank's full-unit gradient
        if self.accum_local:
            if self._last_micro:
                self._gslot_free_evt[l] = self._reduce_unit(1 + l, self.grad_pool[l])
        else:
            slot = l & 1
            self._gslot_free_evt[slot] = self._reduce_unit(1 + l, self.grad_pool[slot])
        self._free_gathered(l)

    def _grad_ready_hook(self, unit: int) -> None:
        if unit == 0 and (self._last_micro or not self.accum_local):
            self._root_done = self._reduce_unit(0, self.root_g)

    # ------------------------------------------------------------------------------------------ micro-step
    def forward_backward(self, input_ids: torch.Tensor, labels: torch.Tensor, loss_scale: float,
                         last: bool = True) -> torch.Tensor:
        """One micro-step.  ``last``: this is the final micro-step before ``step()`` -- with local accumulation the
        gradient reduce-scatter only happens then (callers that never pass it reduce every micro-step).

        Micro-steps without a reduction are pure compute + copy-engine gathers with a fixed host-side schedule
        (pool slots, events), so from the third one on they are replayed from a CUDA graph that spans the main
        and the gather stream (B200_FSDP_GRAPHS=0 / B200_NO_GRAPHS=1: always eager)."""
        from ..ops import kernels as K

        self._last_micro = bool(last)
        e = self.engine
        B, S = input_ids.shape
        ws = e.workspace(B, S, True)
        e._load_inputs(ws, input_ids, labels)
        cur = torch.cuda.current_stream()
        # root unit: gathered once per micro-step, stays resident (as FSDP's root does)
        if getattr(self, "_root_done", None) is not None:
            cur.wait_event(self._root_done)
            self._root_done = None
        if getattr(self, "_root_stale", True):
            self._gather_unit(0, self.root_w)       # parameters only change at the optimizer step
            self._root_stale = False
        if not self.accum_local:
            self.root_g.local.zero_()               # accum_local: zeroed on the comm stream after each reduction
        if not (self.use_graphs and self.accum_local and not last and not self.fuse_ag):
            self._micro_body(ws, loss_scale)
            return ws.loss_sum
        key = (B, S, float(loss_scale))
        entry = self._graphs.get(key)
        self._plain_calls += 1
        if entry is None and self._plain_calls > 2:
            # every dependency inside the capture must be on captured work: join the side streams and drop the
            # events of earlier (eager) micro-steps -- stream order of the graph launch covers them
            cur.wait_stream(self.gather_stream)
            cur.wait_stream(self.comm_stream)
            torch.cuda.synchronize()
            self._drop_events()
            g = torch.cuda.CUDAGraph()
            before = K.launch_count
            with torch.cuda.graph(g):
                self._micro_body(ws, loss_scale)
                torch.cuda.current_stream().wait_stream(self.gather_stream)     # join the forked stream
            self._drop_events()                     # events recorded during capture mean nothing outside it
            entry = (g, K.launch_count - before)
            K.launch_count = before
            self._graphs[key] = entry
        if entry is None:
            self._micro_body(ws, loss_scale)
        else:
            entry[0].replay()
            K.launch_count += entry[1]
        return ws.loss_sum

    def _drop_events(self) -> None:
        self._slot_free_evt = [None] * len(self.gather_pool)
        self._gslot_free_evt = [None] * len(self.grad_pool)
        self._params_ready = None

    def _micro_body(self, ws, loss_scale: float) -> None:
        from ..ops import kernels as K

        e = self.engine
        # fork: gathers issued from here on are ordered after everything already on the main stream (and, under
        # capture, this is what pulls the gather stream into the graph)
        self.gather_stream.wait_stream(torch.cuda.current_stream())
        ws.loss_sum.zero_()
        self._phase = "forward"
        self._prefetch(0)
        e._forward_body(ws, True)
        self._phase = "backward"
        self._prefetch(self.L - 1)                  # BACKWARD_PRE: overlaps the LM head
        e._lm_head_loss(ws, loss_scale, with_grad=True)
        # engine backward with per-block staging buffers
        L = self.L
        e._final_norm_bwd(ws)
        for l in reversed(range(L)):
            slot, staging = self._stage_for(l)
            g = self._views(staging.local, l, self._LayerWeights)
            if not self.accum_local:
                g.ln1.zero_()       # RMSNorm weight gradients accumulate with atomics
                g.ln2.zero_()
            e.g_layers[l] = g
            if e.checkpointing:
                e._layer_fwd(l, ws, ws.x[l], None, True, skip_last=True)
            e._layer_bwd(l, ws, ws.x[l])
            self._release_hook(l, "backward")
        K.embed_bwd(ws.ids, ws.dx, e.g_emb)
        self._grad_ready_hook(0)
        K.rng_advance(e.rng_state, 1)
        self._phase = "idle"

    # ------------------------------------------------------------------------------------------ optimizer step
    def step(self, opt_state, lr: float, betas, eps: float, wd: float, grad_clip: float,
             norm_out: Optional[torch.Tensor] = None) -> None:
        """Sharded clip + AdamW.  ``opt_state`` = dict(exp_avg, exp_avg_sq, step, grad_scale, decay)."""
        from ..ops import kernels as K

        cur = torch.cuda.current_stream()
        cur.wait_stream(self.comm_stream)
        cur.wait_stream(self.gather_stream)         # no gather may still be reading the shards AdamW rewrites
        self.sumsq.zero_()
        K.sumsq(self.grad_shard, 0, self.shard_total, self.sumsq[0:1])
        self.pads.allreduce_scalars(self.sumsq, self.sumsq_total, 1)
        K.clip_coef(self.sumsq_total, float(grad_clip), 1.0, opt_state["grad_scale"], norm_out)
        opt_state["step"] += 1
        st = self.store
        b1, b2 = betas
        if self.cpu_offload:
            self._step_offloaded(opt_state, lr, b1, b2, eps, wd)
            self.grad_shard.zero_()
            self.pads.barrier()
            self._params_ready = torch.cuda.Event()
            self._params_ready.record(cur)
            self._root_done = None
            self._root_stale = True
            return
        for u, (a, b) in enumerate(self.ranges):
            if b <= a:
                continue
            off = self.shard_off[u]
            # every buffer is addressed so that flat element index a lands on this unit's slice of the shard
            K.adamw_step_ptr(self.master_shard.data_ptr() + (off - a) * 4, self.grad_shard.data_ptr() + (off - a) * 4,
                             opt_state["exp_avg"].data_ptr() + off * 4, opt_state["exp_avg_sq"].data_ptr() + off * 4,
                             self.shard_bf16.local.data_ptr() + (off - a) * 2, opt_state["decay"].data_ptr(), a,
                             a + (b - a + 3) // 4 * 4, opt_state["step"], lr, b1, b2, eps, wd,
                             opt_state["grad_scale"].data_ptr())
        self.grad_shard.zero_()
        # peers may gather my new slices only after every rank finished its update
        self.pads.barrier()
        self._params_ready = torch.cuda.Event()
        self._params_ready.record(cur)
        self._root_done = None
        self._root_stale = True

    # ------------------------------------------------------------------------------------------ CPU offload
    def _step_offloaded(self, opt_state, lr, b1, b2, eps, wd) -> None:
        """``--cpu_offload`` (reference: ``CPUOffload(offload_params=True)``, fsdp_trainer.py:300-301): the fp32
        master shard and both AdamW moments live in pinned host memory; the clipped gradient shard goes
        device -> host, AdamW runs on the CPU, the new bf16 shard comes back host -> device."""
        from ..ops.optim import adamw_reference_

        if not hasattr(self, "_host"):
            pin = dict(dtype=torch.float32, pin_memory=True)
            master = torch.empty(self.shard_total, **pin)
            master.copy_(self.master_shard)
            self._host = {"master": master, "grad": torch.empty(self.shard_total, **pin),
                          "bf16": torch.empty(self.shard_total, dtype=torch.bfloat16, pin_memory=True)}
        host = self._host
        scaled = self.grad_shard * opt_state["grad_scale"]          # clip coefficient applied on the device
        host["grad"].copy_(scaled, non_blocking=True)
        torch.cuda.current_stream().synchronize()
        m, v = opt_state["exp_avg"], opt_state["exp_avg_sq"]       # host tensors in this mode
        ones = torch.ones(1)
        adamw_reference_(host["master"], host["grad"], m, v, ones, opt_state["step"], lr, b1, b2, eps, wd, 1.0)
        host["bf16"].copy_(host["master"])
        self.shard_bf16.local.copy_(host["bf16"], non_blocking=True)
        self.master_shard.copy_(host["master"], non_blocking=True)   # device copy of the shard (checkpoints read it)

    # ------------------------------------------------------------------------------------------ checkpoints
    def gather_full_master(self) -> None:
        """Collective: materialise the full fp32 parameters on every rank (module parameters become readable).
        Call ``release_full_master()`` when done with them."""
        st = self.store
        self.materialize_full_master()
        pieces = [self.master_shard[self.shard_off[u]: self.shard_off[u] + b - a] for u, (a, b) in enumerate(self.ranges)]
        _gather_ranges(self.all_ranges, self.world, pieces, st.master)

    def gather_flat(self, shard: torch.Tensor) -> torch.Tensor:
        """shard-layout tensor (concatenated per-unit parts) -> full flat tensor."""
        st = self.store
        full = torch.zeros(st.total, dtype=shard.dtype, device=shard.device)
        pieces = [shard[self.shard_off[u]: self.shard_off[u] + b - a] for u, (a, b) in enumerate(self.ranges)]
        _gather_ranges(self.all_ranges, self.world, pieces, full)
        return full

    def scatter_flat(self, full: torch.Tensor, shard: torch.Tensor) -> None:
        for u, (a, b) in enumerate(self.ranges):
            shard[self.shard_off[u]: self.shard_off[u] + b - a].copy_(full[a:b])
