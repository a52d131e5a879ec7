"""FusedGPTEngine: the whole GPT forward/backward as a fixed sequence of sm_100a kernels over static
workspaces (no autograd graph, no per-step allocation, CUDA-graph capturable).

Per transformer block the forward is 7 launches (the reference's eager graph is ~45 ATen kernels,
SURVEY G2-G8):

    rmsnorm -> [QKV GEMM + RoPE epilogue] -> flash attention (+dropout) -> [o_proj GEMM + dropout +
    residual epilogue] -> rmsnorm -> [gate|up GEMM + SwiGLU epilogue] -> [down GEMM + dropout +
    residual epilogue]

and the backward mirrors it with dgrad/wgrad GEMMs reading the same activations through MN-major
UMMA descriptors (no transposes), wgrad accumulating straight into the flat fp32 gradient buffer
(``red.global.add``, split-K), SwiGLU' fused into the down-proj dgrad epilogue.

The LM head never materialises ``[B*S, V]`` logits (823 MB per micro-batch in the reference,
SURVEY G9/G10): rows are processed in L2-sized chunks -- logits chunk GEMM -> fused softmax-CE
(loss + dlogits in place) -> dX and dW GEMMs -- so the tied-embedding gradient from the head and
from the embedding scatter land in the same buffer.

Residual stream is fp32 (as under the reference's DDP autocast), GEMM operands bf16, accumulation
fp32.  Dropout is counter-based Philox keyed by (seed, step offset, layer/op id): backward and
activation recompute regenerate the forward masks instead of storing them.

Parity targets: reference ``GPT.forward`` gpt.py:388-455, ``TransformerBlock.forward`` :299-316,
gradient checkpointing :441-442 (here: save only the block input and re-run the fused forward).
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Callable, Dict, List, Optional, Tuple

import os
import torch
import torch.nn.functional as F

from ..parallel.flat import FlatParamStore
from . import kernels as K
from .reference import rope_tables

# bf16 logits chunk.  Measured (benchmarks/bench_lmchunk.py, small model, 8192 rows): 24 MB 18.3 ms/micro-step,
# 96 MB 13.2, 192 MB 12.6, 288-420 MB 12.5, one 824 MB chunk 12.7 -- GEMM efficiency at large M beats L2 residency.
_LM_CHUNK_BYTES = 416 << 20


@dataclass
class LayerWeights:
    ln1: torch.Tensor
    wqkv: torch.Tensor
    wo: torch.Tensor
    ln2: torch.Tensor
    wgu: torch.Tensor
    wd: torch.Tensor


class _LayerActs:
    """Activations one block saves for its backward."""

    def __init__(self, M, h, I, B, H, S, dev):
        bf, f32 = torch.bfloat16, torch.float32
        self.xn1 = torch.empty(M, h, dtype=bf, device=dev)
        self.rstd1 = torch.empty(M, dtype=f32, device=dev)
        self.qkv = torch.empty(M, 3 * h, dtype=bf, device=dev)
        self.attn = torch.empty(M, h, dtype=bf, device=dev)
        self.lse = torch.empty(B, H, S, dtype=f32, device=dev)
        self.xmid = torch.empty(M, h, dtype=f32, device=dev)
        self.rstd2 = torch.empty(M, dtype=f32, device=dev)
        self.xn2 = torch.empty(M, h, dtype=bf, device=dev)
        self.gu = torch.empty(M, 2 * I, dtype=bf, device=dev)
        self.act = torch.empty(M, I, dtype=bf, device=dev)


class _Workspace:
    def __init__(self, eng: "FusedGPTEngine", B: int, S: int, training: bool):
        cfg, dev = eng.cfg, eng.device
        h, I, L, H = cfg.hidden_size, cfg.intermediate_size, cfg.num_layers, cfg.num_heads
        M = B * S
        bf, f32 = torch.bfloat16, torch.float32
        self.B, self.S, self.M = B, S, M
        self.ids = torch.zeros(M, dtype=torch.int64, device=dev)
        self.labels = torch.full((B, S), -100, dtype=torch.int64, device=dev)
        self.x = [torch.empty(M, h, dtype=f32, device=dev) for _ in range(L + 1)]
        n_saved = 1 if (eng.checkpointing or not training) else L
        self.acts = [_LayerActs(M, h, I, B, H, S, dev) for _ in range(n_saved)]
        self.xnf = torch.empty(M, h, dtype=bf, device=dev)
        self.rstdf = torch.empty(M, dtype=f32, device=dev)
        rows = max(128, min(M, (_LM_CHUNK_BYTES // (eng.vocab_padded * 2)) // 128 * 128))
        self.lm_rows = rows
        self.logits = torch.empty(rows, eng.vocab_padded, dtype=bf, device=dev)
        self.loss_sum = torch.zeros(1, dtype=f32, device=dev)
        if training:
            self.dxnf = torch.empty(M, h, dtype=bf, device=dev)
            self.dx = torch.empty(M, h, dtype=f32, device=dev)
            self.dy = torch.empty(M, h, dtype=bf, device=dev)
            self.dxn = torch.empty(M, h, dtype=bf, device=dev)
            self.dattn = torch.empty(M, h, dtype=bf, device=dev)
            self.dgu = torch.empty(M, 2 * I, dtype=bf, device=dev)
            self.dact = torch.empty(M, I, dtype=bf, device=dev)
            self.dqkv = torch.empty(M, 3 * h, dtype=bf, device=dev)
            self.delta = torch.empty(B, H, S, dtype=f32, device=dev)
            self.dq_acc = torch.empty(M, h, dtype=f32, device=dev)
        self.graphs: Dict[tuple, tuple] = {}     # (loss_scale, max_ctas, hooked) -> (CUDAGraph, launches per replay)
        self.calls = 0
        self.n_valid: Optional[int] = None       # number of non-ignored targets when the caller masks labels

    def acts_for(self, l: int) -> _LayerActs:
        return self.acts[l if len(self.acts) > 1 else 0]


class FusedGPTEngine:
    def __init__(self, model, store: Optional[FlatParamStore] = None, seed: int = 0):
        cfg = model.config
        self.model = model
        self.cfg = cfg
        if cfg.hidden_size // cfg.num_heads != 64:
            raise ValueError("the sm_100a attention kernel is specialised for head_dim 64 "
                             f"(got {cfg.hidden_size // cfg.num_heads})")
        if cfg.hidden_size % 64 or cfg.intermediate_size % 128:
            raise ValueError("hidden_size must be a multiple of 64 and intermediate_size of 128")
        self.store = store if store is not None else FlatParamStore(model)
        self.device = self.store.device
        if self.device.type != "cuda":
            raise RuntimeError("FusedGPTEngine needs a CUDA device")
        K._lib.lib()  # fail loudly now if the native library is missing
        self.vocab = cfg.vocab_size
        self.vocab_padded = (cfg.vocab_size + 127) // 128 * 128
        self.checkpointing = bool(cfg.gradient_checkpointing)
        self.rope_cos, self.rope_sin = rope_tables(cfg.max_seq_len, 64, device=self.device)
        self.rng_state = torch.tensor([seed & 0x7FFFFFFFFFFFFFFF, 0], dtype=torch.int64, device=self.device)
        self._ws: Dict[Tuple[int, int, bool], _Workspace] = {}
        self.use_graphs = os.environ.get("B200_NO_GRAPHS", "0") != "1"   # B200_NO_GRAPHS=1: eager launches (profiling)
        # epilogue-fusion switch (A/B-able): SwiGLU' inside the down-proj dgrad GEMM (CTA-pair kernel: gate|up
        # prefetched by TMA into the staging ring) vs GEMM + streaming kernel
        self.fuse_swiglu_bwd = True
        self.max_ctas = 0            # cap for GEMM grids when a comm kernel needs resident SMs
        self.wgrad_overwrite = False  # True: every wgrad destination is written once per use (FSDP staging)
        self.sm_count = torch.cuda.get_device_properties(self.device).multi_processor_count
        # hooks used by the DDP / FSDP engines
        self.weights_hook: Optional[Callable[[int], LayerWeights]] = None   # FSDP: gathered weights for layer l
        self.release_hook: Optional[Callable[[int, str], None]] = None      # FSDP: free / reduce after use
        self.grad_ready_hook: Optional[Callable[[int], None]] = None        # DDP: unit's grads are final
        self.end_step_hook: Optional[Callable[[], None]] = None             # DDP: join forked streams
        self.ag_hook: Optional[Callable[[int], Optional[tuple]]] = None     # FSDP: fused all-gather args for QKV
        self.hooks_capturable = False     # True when grad_ready_hook/end_step_hook only launch stream-ordered work
        # DDP / FSDP: key ("L3.qkv", "emb", ...) -> where every rank keeps its rows of that weight gradient.  When
        # set, the wgrad GEMMs reduce-add their tiles into the owners' buffers (GEMM -> reduce-scatter, one kernel)
        self.rs_router: Optional[Callable[[str], Optional[tuple]]] = None
        # what a micro-step does with its weight gradients: "add" accumulates into the local fp32 buffer, "store"
        # overwrites it (first micro-step of a step: no zero-fill needed), "rs" sends local accumulation + this
        # micro-step's product to the owners (last micro-step under DDP / FSDP; needs rs_router)
        self.grad_mode = "add"
        self._bind_views()

    # ------------------------------------------------------------------------------------------ parameter views
    def _bind_views(self) -> None:
        st = self.store
        L = self.cfg.num_layers

        def layer(buf, l):
            pre = f"layers.{l}."
            return LayerWeights(
                ln1=st.view(buf, pre + "input_layernorm.weight"),
                wqkv=st.packed(buf, [pre + f"attention.{n}_proj.weight" for n in "qkv"]),
                wo=st.view(buf, pre + "attention.o_proj.weight"),
                ln2=st.view(buf, pre + "post_attention_layernorm.weight"),
                wgu=st.packed(buf, [pre + "mlp.gate_proj.weight", pre + "mlp.up_proj.weight"]),
                wd=st.view(buf, pre + "mlp.down_proj.weight"),
            )

        self.w_layers: List[LayerWeights] = [layer(st.shadow, l) for l in range(L)]
        self.g_layers: List[LayerWeights] = [layer(st.grad, l) for l in range(L)]
        self.w_emb = st.view(st.shadow, "embed_tokens.weight")
        self.g_emb = st.view(st.grad, "embed_tokens.weight")
        self.w_norm = st.view(st.shadow, "norm.weight")
        self.g_norm = st.view(st.grad, "norm.weight")

    def set_seed(self, seed: int) -> None:
        self.rng_state[0] = seed & 0x7FFFFFFFFFFFFFFF
        self.rng_state[1] = 0

    def _grow_rope(self, S: int) -> None:
        if S > self.rope_cos.shape[0]:
            self.rope_cos, self.rope_sin = rope_tables(S, 64, device=self.device)
            self._ws.clear()

    def workspace(self, B: int, S: int, training: bool) -> _Workspace:
        key = (B, S, training)
        ws = self._ws.get(key)
        if ws is None:
            self._grow_rope(S)
            if not training:
                # inference shapes come and go (generation grows S by one per token): keep the two most recent
                stale = [k for k in self._ws if not k[2]]
                for k in stale[:-1]:
                    del self._ws[k]
            ws = _Workspace(self, B, S, training)
            self._ws[key] = ws
        return ws

    # ------------------------------------------------------------------------------------------ GEMM helpers
    def _wgrad(self, dy: torch.Tensor, x: torch.Tensor, gw: torch.Tensor, key: Optional[str] = None) -> None:
        """gw[N_out, K_in] (fp32) += dy[M, N_out]^T @ x[M, K_in]  -- both operands MN-major.  With an ``rs_router``
        the sum goes to the rows' owners instead of ``gw`` (which then only describes the layout)."""
        n_out, k_in = gw.shape
        bn = 256 if k_in % 256 == 0 else 128
        tiles = ((n_out + 127) // 128) * ((k_in + bn - 1) // bn)
        split = max(1, min(8, self.sm_count // max(tiles, 1)))
        epi = K.EPI_F32_ATOMIC
        rs = self._rs(key)
        if rs is not None:
            # every K split sends its partial product over NVLink: cap the split (traffic) at 4
            K.gemm(dy, x, gw, a_mn=True, b_mn=True, epilogue=epi, split_k=min(split, 4), tile_n=bn,
                   max_ctas=self.max_ctas, M=n_out, N=k_in, K=dy.shape[0], rs=rs)
            return
        if self.wgrad_overwrite or self.grad_mode == "store":   # ("rs1" never gets here: it always has an owner table)
            # the destination holds exactly one contribution (FSDP per-block staging): store instead of
            # reduce-add, so the buffer needs no zero-fill; split-K still accumulates, over a zeroed matrix
            if split > 1:
                gw.zero_()
            else:
                epi = K.EPI_F32
        K.gemm(dy, x, gw, a_mn=True, b_mn=True, epilogue=epi, split_k=split, tile_n=bn,
               max_ctas=self.max_ctas, M=n_out, N=k_in, K=dy.shape[0])

    def _rs(self, key: Optional[str], prev: Optional[bool] = None):
        """Owner table for ``key`` when this micro-step delivers to the owners ("rs": with the local accumulation of
        the earlier micro-steps; "rs1": a step of ONE micro-step, the local buffer holds nothing valid)."""
        if self.grad_mode not in ("rs", "rs1") or self.rs_router is None or key is None:
            return None
        return self.rs_router(key, (self.grad_mode == "rs") if prev is None else prev)

    def _weights(self, l: int) -> LayerWeights:
        return self.weights_hook(l) if self.weights_hook is not None else self.w_layers[l]

    # ------------------------------------------------------------------------------------------ block forward
    def _layer_fwd(self, l: int, ws: _Workspace, x_in: torch.Tensor, x_out: Optional[torch.Tensor], train: bool,
                   skip_last: bool = False) -> None:
        cfg = self.cfg
        W = self._weights(l)
        a = ws.acts_for(l)
        h, I = cfg.hidden_size, cfg.intermediate_size
        p_res = cfg.dropout if train else 0.0
        p_att = cfg.attention_dropout if train else 0.0
        rng = self.rng_state
        K.rmsnorm_fwd(x_in, W.ln1, a.xn1, a.rstd1, 1e-6)
        # FSDP: the QKV projection is the first consumer of a block's gathered weights -- its launch also
        # gathers the q|k|v rows from the peers (all-gather fused into the GEMM) when ag_hook says so
        ag = self.ag_hook(l) if self.ag_hook is not None else None
        K.gemm(a.xn1, W.wqkv, a.qkv, epilogue=K.EPI_ROPE_BF16, rope_cos=self.rope_cos, rope_sin=self.rope_sin,
               seq_len=ws.S, rope_cols=2 * h, max_ctas=self.max_ctas, ag=ag)
        K.attn_fwd(a.qkv, a.attn, a.lse, ws.B, ws.S, cfg.num_heads, p_att, rng, 8 * l)
        K.gemm(a.attn, W.wo, a.xmid, epilogue=K.EPI_RESID_DROPOUT, residual=x_in, dropout_p=p_res, rng_state=rng,
               rng_stream=8 * l + 1, max_ctas=self.max_ctas)
        K.rmsnorm_fwd(a.xmid, W.ln2, a.xn2, a.rstd2, 1e-6)
        K.gemm(a.xn2, W.wgu, a.gu, epilogue=K.EPI_SWIGLU, aux=a.act, half_n=I, tile_n=256, max_ctas=self.max_ctas)
        if not skip_last:
            K.gemm(a.act, W.wd, x_out, epilogue=K.EPI_RESID_DROPOUT, residual=a.xmid, dropout_p=p_res, rng_state=rng,
                   rng_stream=8 * l + 2, max_ctas=self.max_ctas)

    # ------------------------------------------------------------------------------------------ block backward
    def _layer_bwd(self, l: int, ws: _Workspace, x_in: torch.Tensor) -> None:
        """ws.dx holds dL/d(block output) on entry and dL/d(block input) on exit."""
        cfg = self.cfg
        W = self._weights(l)
        G = self.g_layers[l]
        a = ws.acts_for(l)
        h, I = cfg.hidden_size, cfg.intermediate_size
        rng = self.rng_state
        mc = self.max_ctas
        # ---- MLP branch: x_out = xmid + dropout(down(act)).  ws.dy = dropout'(ws.dx) for stream 8l+2 was
        # written by the norm backward that produced ws.dx (fused dropout_bwd_cast).
        self._wgrad(ws.dy, a.act, G.wd, f"L{l}.d")
        if self.fuse_swiglu_bwd:
            K.gemm(ws.dy, W.wd, ws.dgu, b_mn=True, epilogue=K.EPI_SWIGLU_BWD, aux=a.gu, half_n=I, N=I, max_ctas=mc)
        else:
            # d_act reuses the (dead by now) act buffer of this layer; SwiGLU' as a separate streaming kernel
            K.gemm(ws.dy, W.wd, ws.dact, b_mn=True, max_ctas=mc)
            K.swiglu_bwd(ws.dact, a.gu, ws.dgu)
        self._wgrad(ws.dgu, a.xn2, G.wgu, f"L{l}.gu")
        K.gemm(ws.dgu, W.wgu, ws.dxn, b_mn=True, max_ctas=mc)
        # ---- attention branch: xmid = x_in + dropout(o_proj(attn)); its dY comes out of the same pass
        K.rmsnorm_bwd(ws.dxn, a.xmid, W.ln2, a.rstd2, ws.dx, G.ln2, dy_next=ws.dy, p_next=cfg.dropout, rng_state=rng,
                      rng_stream=8 * l + 1)
        self._wgrad(ws.dy, a.attn, G.wo, f"L{l}.o")
        K.gemm(ws.dy, W.wo, ws.dattn, b_mn=True, max_ctas=mc)
        K.attn_bwd(a.qkv, a.attn, ws.dattn, a.lse, ws.dqkv, ws.delta, ws.dq_acc, ws.B, ws.S, cfg.num_heads,
                   cfg.attention_dropout, rng, 8 * l, rope_cos=self.rope_cos, rope_sin=self.rope_sin)
        self._wgrad(ws.dqkv, a.xn1, G.wqkv, f"L{l}.qkv")
        K.gemm(ws.dqkv, W.wqkv, ws.dxn, b_mn=True, max_ctas=mc)
        if l > 0:   # ... and the dY of the previous block's MLP branch
            K.rmsnorm_bwd(ws.dxn, x_in, W.ln1, a.rstd1, ws.dx, G.ln1, dy_next=ws.dy, p_next=cfg.dropout, rng_state=rng,
                          rng_stream=8 * (l - 1) + 2)
        else:
            K.rmsnorm_bwd(ws.dxn, x_in, W.ln1, a.rstd1, ws.dx, G.ln1)

    # ------------------------------------------------------------------------------------------ LM head
    def _lm_head_loss(self, ws: _Workspace, loss_scale: float, with_grad: bool) -> None:
        """Chunked tied LM head + CE.  ``ws.loss_sum += loss_scale * mean CE``; with_grad also fills
        ``ws.dxnf`` and accumulates the head's share of the tied-embedding gradient."""
        M, V = ws.M, self.vocab
        # trainers label every position (labels = input_ids): B*(S-1) targets.  GPT.forward(labels=...) callers may
        # mask targets with -100; module_forward counts them (ws.n_valid) so the mean is over the real targets
        n_valid = ws.n_valid if ws.n_valid is not None else ws.B * (ws.S - 1)
        scale = loss_scale / max(n_valid, 1)
        labels = ws.labels.view(-1)
        w = self._emb_weight()
        for r0 in range(0, M, ws.lm_rows):
            r1 = min(M, r0 + ws.lm_rows)
            logits = ws.logits[: r1 - r0]
            K.gemm(ws.xnf[r0:r1], w, logits, tile_n=256, max_ctas=self.max_ctas, N=self.vocab_padded)
            K.ce_fwd_bwd(logits, labels[r0:r1], ws.loss_sum, V, scale, scale)
            if with_grad:
                # dX chunk: [rows, V] x [V, h], one pass over the 50k-long contraction per output tile, written as
                # bf16 directly.  (Measured, benchmarks/bench_lmhead.py, 4224 rows: split-K 1 254.8 us, split-K 3 with
                # the fp32 reduce-add epilogue 287.6 us -- and split 1 needs no zero-fill and no cast pass.)
                K.gemm(logits, w, ws.dxnf[r0:r1], b_mn=True, tile_n=256, max_ctas=self.max_ctas, K=V)
                # tied-embedding gradient: the first chunk of a "store" micro-step overwrites, the last chunk of an
                # "rs" micro-step carries everything accumulated so far to the owners, the rest accumulate locally
                first, final = r0 == 0, r1 == M
                fresh = self.grad_mode in ("store", "rs1")       # nothing valid in the local buffer yet
                rs = self._rs("emb", prev=(self.grad_mode == "rs" or not first)) if final else None
                K.gemm(logits, ws.xnf[r0:r1], self.g_emb, a_mn=True, b_mn=True,
                       epilogue=K.EPI_F32 if (first and fresh and rs is None) else K.EPI_F32_ATOMIC, tile_n=256,
                       max_ctas=self.max_ctas, M=V, K=r1 - r0, rs=rs)

    def _emb_weight(self) -> torch.Tensor:
        return self.w_emb

    # ------------------------------------------------------------------------------------------ full passes
    def _forward_body(self, ws: _Workspace, train: bool) -> None:
        L = self.cfg.num_layers
        K.embed_fwd(ws.ids, self._emb_weight(), ws.x[0])
        for l in range(L):
            self._layer_fwd(l, ws, ws.x[l], ws.x[l + 1], train)
            if self.release_hook is not None:
                self.release_hook(l, "forward")
        K.rmsnorm_fwd(ws.x[L], self.w_norm, ws.xnf, ws.rstdf, 1e-6)

    def _final_norm_bwd(self, ws: _Workspace) -> None:
        """Start of the backward pass: ws.dx = d loss / d x[L] through the final norm, and ws.dy = its
        dropout-masked bf16 copy for the last block's MLP branch (every _layer_bwd expects ws.dy on entry and
        leaves the previous block's behind)."""
        L = self.cfg.num_layers
        ws.dx.zero_()
        K.rmsnorm_bwd(ws.dxnf, ws.x[L], self.w_norm, ws.rstdf, ws.dx, self.g_norm, dy_next=ws.dy,
                      p_next=self.cfg.dropout, rng_state=self.rng_state, rng_stream=8 * (L - 1) + 2)

    def _backward_body(self, ws: _Workspace) -> None:
        L = self.cfg.num_layers
        self._final_norm_bwd(ws)
        for l in reversed(range(L)):
            if self.checkpointing:
                self._layer_fwd(l, ws, ws.x[l], None, True, skip_last=True)
            self._layer_bwd(l, ws, ws.x[l])
            if self.release_hook is not None:
                self.release_hook(l, "backward")
            if self.grad_ready_hook is not None:
                self.grad_ready_hook(1 + l)
        self._embed_bwd(ws)
        if self.grad_ready_hook is not None:
            self.grad_ready_hook(0)

    def _embed_bwd(self, ws: _Workspace) -> None:
        rs = self._rs("emb")
        if rs is not None:
            K.embed_bwd_rs(ws.ids, ws.dx, rs, self.vocab)
        else:
            K.embed_bwd(ws.ids, ws.dx, self.g_emb)

    def _step_body(self, ws: _Workspace, loss_scale: float) -> None:
        ws.loss_sum.zero_()
        self._forward_body(ws, True)
        self._lm_head_loss(ws, loss_scale, with_grad=True)
        self._backward_body(ws)
        K.rng_advance(self.rng_state, 1)
        if self.end_step_hook is not None:
            self.end_step_hook()

    def _load_inputs(self, ws: _Workspace, input_ids: torch.Tensor, labels: Optional[torch.Tensor]) -> None:
        ws.ids.copy_(input_ids.reshape(-1), non_blocking=True)
        if labels is not None:
            ws.labels[:, :-1].copy_(labels[:, 1:], non_blocking=True)

    # ---- hot path used by the trainers -----------------------------------------------------------------------
    def forward_backward(self, input_ids: torch.Tensor, labels: torch.Tensor, loss_scale: float = 1.0) -> torch.Tensor:
        """One micro-step: accumulates gradients into the flat fp32 buffer and returns the device
        scalar ``loss_scale * mean CE`` (no host sync).  ``labels`` are unshifted, as in the reference."""
        B, S = input_ids.shape
        ws = self.workspace(B, S, True)
        ws.n_valid = None
        self._load_inputs(ws, input_ids, labels)
        ws.calls += 1
        # FSDP hooks allocate pool slots on the host and cannot be replayed; the DDP gradient hook only forks
        # kernels onto a second stream (and joins in end_step_hook), which a captured graph records faithfully
        blocking = (self.release_hook is not None or self.weights_hook is not None
                    or (self.grad_ready_hook is not None and not self.hooks_capturable))
        key = (float(loss_scale), self.max_ctas, self.grad_ready_hook is not None, self.grad_mode)
        if self.use_graphs and not blocking and ws.calls > 2:
            entry = ws.graphs.get(key)
            if entry is None:
                g = torch.cuda.CUDAGraph()
                torch.cuda.synchronize()
                before = K.launch_count
                with torch.cuda.graph(g):
                    self._step_body(ws, loss_scale)
                entry = (g, K.launch_count - before)
                K.launch_count = before          # capture recorded the launches, it did not run them
                ws.graphs[key] = entry
            entry[0].replay()
            K.launch_count += entry[1]
        else:
            self._step_body(ws, loss_scale)
        return ws.loss_sum

    # ---- inference ---------------------------------------------------------------------------------------------
    @torch.no_grad()
    def forward_logits(self, input_ids: torch.Tensor, last_only: bool = False) -> torch.Tensor:
        B, S = input_ids.shape
        ws = self.workspace(B, S, False)
        self._load_inputs(ws, input_ids, None)
        self._forward_body(ws, False)
        w = self._emb_weight()
        if last_only:
            rows = ws.xnf.view(B, S, -1)[:, -1, :].contiguous()
            pad = (-B) % 8
            out = torch.empty(B + pad, self.vocab_padded, dtype=torch.bfloat16, device=self.device)
            K.gemm(rows, w, out, N=self.vocab_padded, M=B)
            return out[:B, : self.vocab].reshape(B, 1, self.vocab)
        out = torch.empty(ws.M, self.vocab_padded, dtype=torch.bfloat16, device=self.device)
        K.gemm(ws.xnf, w, out, tile_n=256, N=self.vocab_padded)
        return out[:, : self.vocab].reshape(B, S, self.vocab)

    @torch.no_grad()
    def forward_loss(self, input_ids: torch.Tensor, labels: torch.Tensor) -> torch.Tensor:
        """Evaluation loss (dropout off, no gradients)."""
        B, S = input_ids.shape
        ws = self.workspace(B, S, False)
        self._load_inputs(ws, input_ids, labels)
        ws.loss_sum.zero_()
        self._forward_body(ws, False)
        self._lm_head_loss(ws, 1.0, with_grad=False)
        return ws.loss_sum.clone()

    def _check_labels(self, input_ids, labels, training: bool) -> None:
        """Public-API label handling (F.cross_entropy semantics, reference gpt.py:449-453): -100 targets are ignored
        and the loss is the mean over the remaining ones; a label >= vocab raises.  One host sync, on this path only."""
        B, S = input_ids.shape
        ws = self.workspace(B, S, training)
        ws.n_valid = None
        if labels is None:
            return
        tgt = labels[:, 1:]
        stats = torch.stack([(tgt >= 0).sum(), tgt.max() if tgt.numel() else tgt.new_zeros(())]).tolist()
        if stats[1] >= self.vocab:
            raise IndexError(f"label {int(stats[1])} is out of range for vocab_size {self.vocab}")
        if stats[0] != tgt.numel():
            ws.n_valid = int(stats[0])

    def _sync_with_module(self) -> None:
        """``loss.backward()`` callers drive the model with stock torch code (``torch.optim.AdamW``,
        ``optimizer.zero_grad()``, ``load_state_dict``), which only sees the fp32 master views: the bf16 compute copy
        is refreshed whenever a parameter changed since the last sync, and dropped ``.grad`` tensors
        (``zero_grad(set_to_none=True)``) are re-bound to the zeroed flat gradient buffer."""
        st = self.store
        versions = tuple(p._version for p in st.params)
        if versions != getattr(self, "_param_versions", None):
            st.refresh_shadow()
            self._param_versions = versions
        if st.grad is not None and any(p.grad is None for p in st.params):
            st.grad.zero_()
            st.rebind_grads()

    # ---- nn.Module-compatible entry (GPT.forward) ------------------------------------------------------------
    def module_forward(self, input_ids: torch.Tensor, labels: Optional[torch.Tensor]):
        training = self.model.training and torch.is_grad_enabled()
        self._check_labels(input_ids, labels, training)
        if not training:
            logits = self.forward_logits(input_ids)
            loss = None
            if labels is not None:
                loss = self.forward_loss(input_ids, labels).squeeze(0)
            return logits, loss
        if labels is None:
            raise NotImplementedError("training-mode forward on the fused engine returns the loss only; "
                                      "pass labels (logits are never materialised), or call model.eval()")
        self._sync_with_module()
        loss = _EngineLoss.apply(self, input_ids, labels, *self.store.params)
        return None, loss


class _EngineLoss(torch.autograd.Function):
    """Autograd bridge for ``loss.backward()`` callers.  Forward computes the loss; backward re-runs the
    LM head with the incoming scale (so ``(loss / GA).backward()`` is exact) and the fused backward,
    accumulating into ``param.grad`` (views of the flat gradient buffer)."""

    @staticmethod
    def forward(ctx, eng: FusedGPTEngine, input_ids, labels, *params):
        B, S = input_ids.shape
        ws = eng.workspace(B, S, True)
        eng._load_inputs(ws, input_ids, labels)
        ws.loss_sum.zero_()
        eng._forward_body(ws, True)
        eng._lm_head_loss(ws, 1.0, with_grad=False)
        ctx.eng, ctx.ws = eng, ws
        return ws.loss_sum.clone().squeeze(0)

    @staticmethod
    def backward(ctx, grad_out):
        eng, ws = ctx.eng, ctx.ws
        scale = float(grad_out.item())
        keep = ws.loss_sum.clone()
        eng._lm_head_loss(ws, scale, with_grad=True)
        ws.loss_sum.copy_(keep)
        eng._backward_body(ws)
        K.rng_advance(eng.rng_state, 1)
        return (None, None, None) + tuple(None for _ in eng.store.params)
