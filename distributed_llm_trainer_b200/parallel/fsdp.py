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
    * gradients and optimizer state follow a second partition, 128-row slabs of every weight matrix
      (``parallel/ownership.py``): micro-steps before the last accumulate into this rank's own flat fp32
      buffer, and the LAST micro-step's wgrad GEMMs reduce-add accumulator + product straight into the
      owners' buffers over NVLink (``b200_gemm2_rs_bf16``) -- the reduce-scatter is the epilogue of the
      GEMM that produces the gradient, one crossing per step (``B200_FSDP_ACCUM_LOCAL=0``: every
      micro-step delivers, as FSDP1 does; the reference has no ``no_sync``);
    * the step is: barrier -> norm^2 of the owned rows -> pad all-reduce -> clip coefficient -> ONE
      fused AdamW launch that writes the new bf16 values into the parameter shards of whichever
      ranks hold them (``b200_adamw_chunks``) -> barrier.  A micro-step contains no cross-rank
      synchronisation, so all of them replay from CUDA graphs.

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
        if world > 8:
            raise NotImplementedError("the NVLink engines address the ranks of ONE NVSwitch domain (<= 8 GPUs): the wgrad "
                                      f"GEMM carries one tensor map per owner; got world_size={world}")
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
        # bf16 parameter shards (what peers gather from): flat 1/N slices of every unit
        self.shard_bf16 = SymmetricBuffer(self.shard_total, torch.bfloat16, dev)
        # Gradient / optimizer ownership is a different partition: 128-row slabs of every weight matrix
        # (parallel/ownership.py), so that the wgrad GEMMs can reduce-add their tiles straight into the owner's
        # buffer.  An owner keeps the fp32 master copy, both AdamW moments and the reduced gradient of its rows,
        # stored back to back in flat order ("owner layout"); its AdamW launch writes the new bf16 values into
        # whichever rank's parameter shard holds them.
        from .ownership import OwnershipPlan

        self.plan = OwnershipPlan(st.slots, world)
        self.own_ranges_all = self.plan.all_ranges()
        self.own_ranges = self.own_ranges_all[rank]
        self.own_off: List[int] = []
        cur = 0
        for a_, b_ in self.own_ranges:
            self.own_off.append(cur)
            cur += b_ - a_
        self.own_total = cur
        own_max = max(sum(b_ - a_ for a_, b_ in rr) for rr in self.own_ranges_all)
        self.red_buf = SymmetricBuffer(max(own_max, ALIGN), torch.float32, dev)    # owners' rows: cross-rank sums land here
        self.master_shard = torch.zeros(max(self.own_total, ALIGN), dtype=torch.float32, device=dev)
        self.keep_full_master = os.environ.get("B200_FSDP_FULL_MASTER", "0") == "1"
        self._refresh_shard_from_master()
        self.release_full_master()
        # pools
        blk = max(self.unit_part[1:]) * world if self.L > 0 else ALIGN
        root = self.unit_part[0] * world
        # limit_all_gathers (reference fsdp_trainer.py:306): True = at most one gather in flight beyond the unit
        # being computed (3 pooled buffers); False = prefetch two units ahead (more buffers, more NVLink overlap)
        self.limit_all_gathers = bool(limit_all_gathers)
        self.backward_prefetch = bool(backward_prefetch)
        depth = 3 if self.limit_all_gathers else 5
        n_gather = self.L if strategy == "SHARD_GRAD_OP" else min(self.L, depth)
        slack = 64 * model.config.hidden_size   # the LM-head GEMM's last B tile may read past V rows
        self.root_w = torch.zeros(root + slack, dtype=torch.bfloat16, device=dev)
        self.gather_pool = [torch.zeros(blk, dtype=torch.bfloat16, device=dev) for _ in range(max(n_gather, 1))]
        # This rank's own fp32 gradient accumulator (flat, full size): the micro-steps before the last add into it
        # locally; the last micro-step's wgrad GEMMs send accumulator + product to the owners, so a step's gradient
        # crosses NVLink once (the reference reduce-scatters every micro-step, fsdp_trainer.py:369-384;
        # B200_FSDP_ACCUM_LOCAL=0 does the same here: every micro-step delivers to the owners).  Peer-mapped only so
        # that owners can pull the few norm-weight gradients.
        from .ddp import symmetric_or_multicast

        self.acc_buf = symmetric_or_multicast(st.total, torch.float32, dev)
        self.accum_local = os.environ.get("B200_FSDP_ACCUM_LOCAL", "1") != "0"
        # Parameter gathers run on their own stream, prefetched ahead of the compute walk.
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
        self.use_graphs = (os.environ.get("B200_FSDP_GRAPHS", "1") != "0"
                           and os.environ.get("B200_NO_GRAPHS", "0") != "1")
        self._graphs: Dict[Tuple[int, int, float], Any] = {}
        self._plain_calls: Dict[Any, int] = {}
        # The fused all-gather GEMM is opt-in (B200_FSDP_FUSE_AG=1): a PREFETCHED gather hides completely under
        # the previous block's compute, the fused one only under its own GEMM.  Round 2 (graph-replayable, device-side
        # epoch): 355M benchmark at 2 GPUs 161.2-163.5 ms without, 172.6 / 165.3 / 167.1 / 167.9 ms with 2 / 6 / 10 / 16
        # gather pairs (profiles/fsdp_comm_breakdown.md) -- still 1-2 % behind, so it stays opt-in.
        self.fuse_ag = (os.environ.get("B200_FSDP_FUSE_AG", "0") != "0" and self.L > 0
                        and min(self.unit_part[1:]) >= 16384 and 3 * h >= 256)
        self.ag_pairs = int(os.environ.get("B200_FSDP_AG_PAIRS", "6"))
        self._ag_flags = [torch.zeros((3 * h + 127) // 128, dtype=torch.int32, device=dev) for _ in self.gather_pool]
        # per pool slot: {unused, epoch} as two int64 words; the fused all-gather GEMM reads the low half of word 1,
        # bumped on the stream before every launch (device-side, so the launch replays from a CUDA graph)
        self._ag_epoch_state = [torch.zeros(2, dtype=torch.int64, device=dev) for _ in self.gather_pool]
        # engine bound to hooks
        self.engine = FusedGPTEngine.__new__(FusedGPTEngine)
        self._init_engine(model, seed, LayerWeights)
        self.sumsq = torch.zeros(8, dtype=torch.float32, device=dev)
        self.sumsq_total = torch.zeros(8, dtype=torch.float32, device=dev)
        self._gathered: Dict[int, Tuple[int, torch.cuda.Event]] = {}   # layer -> (pool slot, ready event)
        self._slot_free_evt: List[Optional[torch.cuda.Event]] = [None] * len(self.gather_pool)
        self._slot_owner: List[Optional[int]] = [None] * len(self.gather_pool)
        self._root_ready: Optional[torch.cuda.Event] = None
        self._phase = "idle"
        self._build_owner_tables()
        torch.cuda.synchronize()
        dist.barrier()

    def close(self) -> None:
        """Collective: release the peer-mapped buffers (the object must not be used afterwards)."""
        torch.cuda.synchronize()
        self._graphs.clear()
        self.engine.rs_router = None
        for b in (self.acc_buf, self.red_buf, self.shard_bf16, self.pads.buf):
            b.close()

    # ------------------------------------------------------------------------------------------ setup helpers
    def _refresh_shard_from_master(self) -> None:
        """Full fp32 parameters (materialised) -> this rank's fp32 master rows (owner layout) and its bf16 parameter
        shard (flat-slice layout)."""
        st = self.store
        with torch.no_grad():
            for (a, b), off in zip(self.own_ranges, self.own_off):
                self.master_shard[off: off + b - a].copy_(st.master[a:b])
            for u, (a, b) in zip(range(len(self.units)), self.ranges):
                off = self.shard_off[u]
                self.shard_bf16.local[off: off + b - a].copy_(st.master[a:b])

    def _build_owner_tables(self) -> None:
        """Owner tables for the wgrad GEMMs (where every rank keeps its rows of each weight gradient) and the chunk
        tables of the optimizer launch (owner layout -> flat-slice parameter shards)."""
        from ..ops import kernels as K

        rank, world = self.rank, self.world
        # compact offset of every owned range, per rank
        off_of = []
        for rr in self.own_ranges_all:
            d, cur = {}, 0
            for a, b in rr:
                d[a] = cur
                cur += b - a
            off_of.append(d)
        self._rs: Dict[Tuple[str, bool], tuple] = {}
        for key, g in self.plan.groups.items():
            bases = []
            for o in range(world):
                lo, hi = g.flat_range(o, world)
                bases.append(self.red_buf.raw_ptrs[o] + 4 * off_of[o][lo] if hi > lo else None)
            self._rs[(key, False)] = K.make_rs_args(bases, g.rot)
            self._rs[(key, True)] = K.make_rs_args(bases, g.rot, prev_ptr=self.acc_buf.raw_ptrs[rank] + 4 * g.offset)
        self.engine.rs_router = lambda key, prev: self._rs.get((key, bool(prev)))
        # AdamW runs: split every owned range where the flat-slice holder of the bf16 parameters changes
        runs = []
        unit_of = []
        for u, un in enumerate(self.units):
            unit_of.append((un.lo, un.hi, u))
        for (a, b), off in zip(self.own_ranges, self.own_off):
            u = next(i for lo, hi, i in unit_of if lo <= a < hi)
            ulo, part = self.units[u].lo, self.unit_part[u]
            f = a
            while f < b:
                dst = (f - ulo) // part
                end = min(b, ulo + (dst + 1) * part)
                runs.append((f, off + (f - a), off + (f - a), self.shard_off[u] + (f - ulo - dst * part), end - f, dst))
                f = end
        self._chunks = K.make_chunk_table(runs, self.device)
        small = {lo for lo, _ in self.plan.small_ranges(rank)}
        pulls = [(a, off, 0, 0, b - a, -1) for (a, b), off in zip(self.own_ranges, self.own_off) if a in small]
        self._pull_chunks = K.make_chunk_table(pulls, self.device)

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
                for (a, b), off in zip(self.own_ranges, self.own_off):
                    st.master[a:b].copy_(self.master_shard[off: off + b - a])
            return
        st.master = torch.zeros(st.total, dtype=torch.float32, device=self.device)
        for p, s in zip(st.params, st.slots):
            p.data = st.master[s.offset: s.offset + s.numel].view(s.shape)
        for (a, b), off in zip(self.own_ranges, self.own_off):
            st.master[a:b].copy_(self.master_shard[off: off + b - a])

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
        e.wgrad_overwrite = False
        e.grad_mode = "add"
        e.sm_count = torch.cuda.get_device_properties(self.device).multi_processor_count
        self._LayerWeights = LayerWeights
        st = self.store
        root_u = self.units[0]

        def rootv(buf, name):
            s = st.by_name[name]
            return buf[s.offset - root_u.lo: s.offset - root_u.lo + s.numel].view(s.shape)

        e.w_emb = rootv(self.root_w, "embed_tokens.weight")
        e.w_norm = rootv(self.root_w, "norm.weight")
        acc = self.acc_buf.local
        e.g_emb = st.view(acc, "embed_tokens.weight")
        e.g_norm = st.view(acc, "norm.weight")
        e.w_layers = []
        e.g_layers = []
        for l in range(self.L):        # gradient views into this rank's flat accumulator (as the DDP engine's)
            pre = f"layers.{l}."
            e.g_layers.append(LayerWeights(
                ln1=st.view(acc, pre + "input_layernorm.weight"),
                wqkv=st.packed(acc, [pre + f"attention.{n}_proj.weight" for n in "qkv"]),
                wo=st.view(acc, pre + "attention.o_proj.weight"),
                ln2=st.view(acc, pre + "post_attention_layernorm.weight"),
                wgu=st.packed(acc, [pre + "mlp.gate_proj.weight", pre + "mlp.up_proj.weight"]),
                wd=st.view(acc, pre + "mlp.down_proj.weight")))
        e.ag_hook = self._ag_hook
        e.weights_hook = self._weights_hook
        e.release_hook = self._release_hook
        e.grad_ready_hook = None

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
        from ..ops import kernels as K

        K.rng_advance(self._ag_epoch_state[slot], 1)
        return (self._unit_ptrs(1 + l), self.unit_part[1 + l], self.qkv_off, self._ag_flags[slot],
                self._ag_epoch_state[slot][1:2],
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
        # prefetch the unit(s) the walk needs next: forward always (FSDP's forward_prefetch), backward only with
        # backward_prefetch=BACKWARD_PRE (reference fsdp_trainer.py:291-307); limit_all_gathers=False looks two ahead
        if self._phase == "forward":
            self._prefetch(l + 1)
            if not self.limit_all_gathers:
                self._prefetch(l + 2)
        elif self.backward_prefetch:
            self._prefetch(l - 1)
            if not self.limit_all_gathers:
                self._prefetch(l - 2)
        return self._views(self.gather_pool[slot], l, self._LayerWeights)

    def _free_gathered(self, l: int) -> None:
        slot = self._gathered.pop(l)[0]
        ev = torch.cuda.Event()
        ev.record(torch.cuda.current_stream())
        self._slot_free_evt[slot] = ev
        self._slot_owner[slot] = None

    def _release_hook(self, l: int, phase: str) -> None:
        if phase == "forward":
            if self.strategy == "FULL_SHARD":
                self._free_gathered(l)
            return
        self._free_gathered(l)         # backward of block l finished: its gathered weights are dead

    # ------------------------------------------------------------------------------------------ micro-step
    def forward_backward(self, input_ids: torch.Tensor, labels: torch.Tensor, loss_scale: float,
                         last: bool = True, first: Optional[bool] = None) -> torch.Tensor:
        """One micro-step.  ``first`` / ``last``: position inside the optimizer step -- the first micro-step overwrites
        this rank's gradient accumulator, the middle ones add to it, and the last one's wgrad GEMMs send accumulator +
        product to the rows' owners (callers that pass neither deliver every micro-step).

        A micro-step is pure compute + copy-engine gathers with a fixed host-side schedule (pool slots, events) and
        contains no cross-rank synchronisation -- the reduce-adds into the owners are ordinary TMA stores of the wgrad
        kernels -- so from the third call on every kind of micro-step is replayed from a CUDA graph that spans the
        main and the gather stream (B200_FSDP_GRAPHS=0 / B200_NO_GRAPHS=1: always eager)."""
        from ..ops import kernels as K

        e = self.engine
        if first is None:
            first = bool(last)
        if not self.accum_local or (first and last):
            e.grad_mode = "rs1"
        else:
            e.grad_mode = "store" if first else ("rs" if last else "add")
        B, S = input_ids.shape
        ws = e.workspace(B, S, True)
        e._load_inputs(ws, input_ids, labels)
        cur = torch.cuda.current_stream()
        # root unit: gathered once per optimizer step, stays resident (as FSDP's root does)
        if getattr(self, "_root_stale", True):
            if self._params_ready is not None:
                cur.wait_event(self._params_ready)
            self._gather_unit(0, self.root_w)       # parameters only change at the optimizer step
            self._root_stale = False
        if not self.use_graphs:
            self._micro_body(ws, loss_scale)
            return ws.loss_sum
        key = (B, S, float(loss_scale), e.grad_mode)
        entry = self._graphs.get(key)
        self._plain_calls[key] = self._plain_calls.get(key, 0) + 1
        if entry is None and self._plain_calls[key] > 2:
            # every dependency inside the capture must be on captured work: join the side stream and drop the
            # events of earlier (eager) micro-steps -- stream order of the graph launch covers them
            cur.wait_stream(self.gather_stream)
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
            if self._params_ready is not None:      # first replay after an optimizer step: peers' new shards are ready
                cur.wait_event(self._params_ready)
                self._params_ready = None
            entry[0].replay()
            K.launch_count += entry[1]
        return ws.loss_sum

    def _drop_events(self) -> None:
        self._slot_free_evt = [None] * len(self.gather_pool)
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
        if self.backward_prefetch:
            self._prefetch(self.L - 1)              # BACKWARD_PRE: overlaps the LM head
        e._lm_head_loss(ws, loss_scale, with_grad=True)
        e._final_norm_bwd(ws)
        for l in reversed(range(self.L)):
            if e.checkpointing:
                e._layer_fwd(l, ws, ws.x[l], None, True, skip_last=True)
            e._layer_bwd(l, ws, ws.x[l])
            self._release_hook(l, "backward")
        e._embed_bwd(ws)
        K.rng_advance(e.rng_state, 1)
        self._phase = "idle"

    # ------------------------------------------------------------------------------------------ optimizer step
    def step(self, opt_state, lr: float, betas, eps: float, wd: float, grad_clip: float,
             norm_out: Optional[torch.Tensor] = None) -> None:
        """Sharded clip + AdamW over the rows this rank owns.  ``opt_state`` = dict(exp_avg, exp_avg_sq, step,
        grad_scale, decay), moments in owner layout.  The reduce-scatter has already happened inside the last
        micro-step's wgrad GEMMs; what is left is: barrier (all reduce-adds landed) -> owners pull the norm-weight
        gradients -> squared norm -> scalar all-reduce -> clip coefficient -> ONE AdamW launch that writes the new
        bf16 values into the parameter shards of whichever ranks hold them and clears the gradient -> barrier."""
        from ..ops import kernels as K

        cur = torch.cuda.current_stream()
        cur.wait_stream(self.gather_stream)         # no gather may still be reading the shards AdamW rewrites
        self.sumsq.zero_()
        self.pads.barrier()
        red = self.red_buf.local
        table, n = self._pull_chunks
        K.pull_reduce_chunks(self.acc_buf.ptrs, red, table, n, self.rank, self.world, zero_remote=True,
                             multicast_ptr=getattr(self.acc_buf, "mc_ptr", 0))
        table, n = self._chunks
        K.sumsq_chunks(red, table, n, self.sumsq[0:1], 1.0)
        self.pads.allreduce_scalars(self.sumsq, self.sumsq_total, 1)
        # the buffers hold SUMS over ranks: the 1/world of the mean is folded into the clip coefficient
        K.clip_coef(self.sumsq_total, float(grad_clip), 1.0 / self.world, opt_state["grad_scale"], norm_out)
        opt_state["step"] += 1
        b1, b2 = betas
        if self.cpu_offload:
            self._step_offloaded(opt_state, lr, b1, b2, eps, wd)
        else:
            K.adamw_chunks(self.master_shard, red, opt_state["exp_avg"], opt_state["exp_avg_sq"], self.shard_bf16.ptrs,
                           self.rank, self.world, opt_state["decay"], table, n, opt_state["step"], lr, b1, b2, eps, wd,
                           opt_state["grad_scale"], zero_grad=True, max_ctas=self.max_ctas)
        # peers may gather the new shards (and send the next step's gradients) only after every rank finished
        self.pads.barrier()
        self._params_ready = torch.cuda.Event()
        self._params_ready.record(cur)
        self._root_stale = True

    # ------------------------------------------------------------------------------------------ CPU offload
    def _step_offloaded(self, opt_state, lr, b1, b2, eps, wd) -> None:
        """``--cpu_offload`` (reference: ``CPUOffload(offload_params=True)``, fsdp_trainer.py:300-301): the fp32
        master rows and both AdamW moments live in pinned host memory.  The owned gradient goes device -> host in
        pieces on a side stream while the host runs AdamW on the previous piece and the updated master rows of the
        piece before that travel back (pinned double buffering, three stages in flight); one device launch then
        casts the new values into the holders' bf16 parameter shards and clears the gradient."""
        from ..ops import kernels as K
        from ..ops.optim import adamw_reference_

        n = self.own_total
        if not hasattr(self, "_host"):
            pin = dict(dtype=torch.float32, pin_memory=True)
            master = torch.empty(n, **pin)
            master.copy_(self.master_shard[:n])
            self._host = {"master": master, "grad": torch.empty(n, **pin), "stream": torch.cuda.Stream(device=self.device)}
        host = self._host
        side = host["stream"]
        red = self.red_buf.local
        red[:n].mul_(opt_state["grad_scale"])                     # clip coefficient (and 1/world) applied on the device
        side.wait_stream(torch.cuda.current_stream())
        pieces = max(1, min(8, n // (1 << 20)))
        bounds = [n * i // pieces // 4 * 4 for i in range(pieces)] + [n]
        d2h = []
        with torch.cuda.stream(side):
            for i in range(pieces):
                a, b = bounds[i], bounds[i + 1]
                host["grad"][a:b].copy_(red[a:b], non_blocking=True)
                ev = torch.cuda.Event()
                ev.record(side)
                d2h.append(ev)
        m, v = opt_state["exp_avg"], opt_state["exp_avg_sq"]       # host tensors in this mode
        ones = torch.ones(1)
        for i in range(pieces):
            a, b = bounds[i], bounds[i + 1]
            d2h[i].synchronize()                                    # piece i is on the host; pieces > i still in flight
            adamw_reference_(host["master"][a:b], host["grad"][a:b], m[a:b], v[a:b], ones, opt_state["step"], lr, b1, b2,
                             eps, wd, 1.0)
            with torch.cuda.stream(side):                           # H2D of piece i overlaps the host AdamW of piece i+1
                self.master_shard[a:b].copy_(host["master"][a:b], non_blocking=True)
        torch.cuda.current_stream().wait_stream(side)
        table, nch = self._chunks
        K.cast_scatter_chunks(self.master_shard, red, self.shard_bf16.ptrs, self.rank, self.world, table, nch)

    # ------------------------------------------------------------------------------------------ checkpoints
    def gather_full_master(self) -> None:
        """Collective: materialise the full fp32 parameters on every rank (module parameters become readable).
        Call ``release_full_master()`` when done with them."""
        st = self.store
        self.materialize_full_master()
        pieces = [self.master_shard[off: off + b - a] for (a, b), off in zip(self.own_ranges, self.own_off)]
        _gather_ranges(self.own_ranges_all, self.world, pieces, st.master)

    def gather_flat(self, shard: torch.Tensor) -> torch.Tensor:
        """owner-layout tensor (this rank's owned rows back to back) -> full flat tensor."""
        st = self.store
        full = torch.zeros(st.total, dtype=shard.dtype, device=shard.device)
        pieces = [shard[off: off + b - a] for (a, b), off in zip(self.own_ranges, self.own_off)]
        _gather_ranges(self.own_ranges_all, self.world, pieces, full)
        return full

    def scatter_flat(self, full: torch.Tensor, shard: torch.Tensor) -> None:
        for (a, b), off in zip(self.own_ranges, self.own_off):
            shard[off: off + b - a].copy_(full[a:b])
