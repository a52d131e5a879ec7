"""GPT (LLaMA-style decoder: RMSNorm + RoPE + SwiGLU, tied LM head).

Parity target: reference ``src/models/gpt.py`` -- module and parameter names are
identical so state dicts interchange with the reference in both directions
(SURVEY.md section 5.4):

    embed_tokens.weight, layers.{i}.input_layernorm.weight,
    layers.{i}.attention.{q,k,v,o}_proj.weight,
    layers.{i}.attention.rotary_emb.{inv_freq,cos_cached,sin_cached},
    layers.{i}.post_attention_layernorm.weight,
    layers.{i}.mlp.{gate,up,down}_proj.weight, norm.weight, lm_head.weight (tied)

Two execution paths share these modules:

* ``backend="torch"`` -- plain PyTorch restatement of the reference math
  (gpt.py:66-67 RMSNorm, :116-147 RoPE, :230-234 attention, :278-283 SwiGLU,
  :438-455 forward/loss).  It is the numerical oracle for every CUDA kernel and
  the path CPU-only processes run.
* ``backend="b200"`` -- the hand-written sm_100a engine
  (``distributed_llm_trainer_b200.ops.engine.FusedGPTEngine``): tcgen05 GEMMs,
  fused RMSNorm / RoPE / SwiGLU / dropout-residual kernels, flash attention,
  chunked LM-head + cross-entropy.  Selected automatically once the model lives
  on a CUDA device; it raises if the native library is missing instead of
  silently falling back.
"""
from __future__ import annotations

import math
from typing import Optional, Tuple

import torch
import torch.nn as nn
import torch.nn.functional as F
from torch.utils.checkpoint import checkpoint as _torch_checkpoint

from .config import GPTConfig


# ----------------------------------------------------------------------------- layers
class RMSNorm(nn.Module):
    """``x * rsqrt(mean(x^2) + eps) * weight`` (reference gpt.py:22-67), eps 1e-6, no bias."""

    def __init__(self, hidden_size: int, eps: float = 1e-6):
        super().__init__()
        self.weight = nn.Parameter(torch.ones(hidden_size))
        self.eps = eps

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        inv = torch.rsqrt(x.pow(2).mean(dim=-1, keepdim=True) + self.eps)
        return x * inv * self.weight


class RotaryPositionEmbedding(nn.Module):
    """NeoX-style (half-split) rotary tables, reference gpt.py:70-98.

    ``inv_freq``, ``cos_cached`` and ``sin_cached`` are persistent buffers so they
    appear in the state dict exactly like the reference's.  The CUDA engine does
    not read them per layer (it keeps one fp32 table for the whole model -- the
    tables are deterministic, which also removes the reference's per-step DDP
    buffer broadcast).
    """

    def __init__(self, dim: int, max_seq_len: int = 2048, base: int = 10000):
        super().__init__()
        self.dim = dim
        self.max_seq_len = max_seq_len
        self.base = base
        inv_freq = 1.0 / (base ** (torch.arange(0, dim, 2).float() / dim))
        self.register_buffer("inv_freq", inv_freq)
        self._build_cache(max_seq_len)

    def _build_cache(self, seq_len: int) -> None:
        pos = torch.arange(seq_len, device=self.inv_freq.device)
        ang = torch.outer(pos.to(self.inv_freq.dtype), self.inv_freq)
        ang = torch.cat((ang, ang), dim=-1)
        self.register_buffer("cos_cached", ang.cos())
        self.register_buffer("sin_cached", ang.sin())
        self.max_seq_len = max(self.max_seq_len, seq_len)

    def forward(self, x: torch.Tensor, seq_len: int) -> Tuple[torch.Tensor, torch.Tensor]:
        if seq_len > self.cos_cached.shape[0]:
            self._build_cache(seq_len)
        return self.cos_cached[:seq_len], self.sin_cached[:seq_len]


def rotate_half(x: torch.Tensor) -> torch.Tensor:
    """[a, b, c, d] -> [-c, -d, a, b]  (reference gpt.py:101-118)."""
    half = x.shape[-1] // 2
    return torch.cat((-x[..., half:], x[..., :half]), dim=-1)


def apply_rotary_pos_emb(q, k, cos, sin):
    """Reference gpt.py:122-147."""
    return q * cos + rotate_half(q) * sin, k * cos + rotate_half(k) * sin


class CausalSelfAttention(nn.Module):
    """Reference gpt.py:150-242.  ``attention_mask`` is accepted and ignored, as there."""

    def __init__(self, config: GPTConfig):
        super().__init__()
        self.config = config
        self.num_heads = config.num_heads
        self.head_dim = config.hidden_size // config.num_heads
        h = config.hidden_size
        self.q_proj = nn.Linear(h, h, bias=False)
        self.k_proj = nn.Linear(h, h, bias=False)
        self.v_proj = nn.Linear(h, h, bias=False)
        self.o_proj = nn.Linear(h, h, bias=False)
        self.attn_dropout = nn.Dropout(config.attention_dropout)
        self.resid_dropout = nn.Dropout(config.dropout)
        self.rotary_emb = RotaryPositionEmbedding(self.head_dim, config.max_seq_len)
        self.use_flash = bool(config.use_flash_attention)

    def forward(self, hidden_states: torch.Tensor, attention_mask: Optional[torch.Tensor] = None):
        b, s, _ = hidden_states.shape
        shp = (b, s, self.num_heads, self.head_dim)
        q = self.q_proj(hidden_states).view(shp).transpose(1, 2)
        k = self.k_proj(hidden_states).view(shp).transpose(1, 2)
        v = self.v_proj(hidden_states).view(shp).transpose(1, 2)
        cos, sin = self.rotary_emb(q, s)
        q, k = apply_rotary_pos_emb(q, k, cos[None, None], sin[None, None])
        if self.use_flash:
            ctx = F.scaled_dot_product_attention(
                q, k, v, attn_mask=None,
                dropout_p=self.config.attention_dropout if self.training else 0.0,
                is_causal=True,
            )
        else:
            scores = torch.matmul(q, k.transpose(-2, -1)) / math.sqrt(self.head_dim)
            future = torch.ones(s, s, device=scores.device, dtype=torch.bool).triu(1)
            scores = scores.masked_fill(future, float("-inf"))
            probs = F.softmax(scores, dim=-1, dtype=torch.float32).to(scores.dtype)
            ctx = self.attn_dropout(probs) @ v
        ctx = ctx.transpose(1, 2).contiguous().view(b, s, -1)
        return self.resid_dropout(self.o_proj(ctx))


class MLP(nn.Module):
    """SwiGLU feed-forward: ``down(silu(gate(x)) * up(x))`` + dropout (reference gpt.py:245-283)."""

    def __init__(self, config: GPTConfig):
        super().__init__()
        h, i = config.hidden_size, config.intermediate_size
        self.gate_proj = nn.Linear(h, i, bias=False)
        self.up_proj = nn.Linear(h, i, bias=False)
        self.down_proj = nn.Linear(i, h, bias=False)
        self.dropout = nn.Dropout(config.dropout)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        return self.dropout(self.down_proj(F.silu(self.gate_proj(x)) * self.up_proj(x)))


class TransformerBlock(nn.Module):
    """Pre-norm block (reference gpt.py:286-316)."""

    def __init__(self, config: GPTConfig, layer_idx: int):
        super().__init__()
        self.layer_idx = layer_idx
        self.input_layernorm = RMSNorm(config.hidden_size)
        self.attention = CausalSelfAttention(config)
        self.post_attention_layernorm = RMSNorm(config.hidden_size)
        self.mlp = MLP(config)

    def forward(self, hidden_states: torch.Tensor, attention_mask: Optional[torch.Tensor] = None):
        hidden_states = hidden_states + self.attention(self.input_layernorm(hidden_states), attention_mask)
        hidden_states = hidden_states + self.mlp(self.post_attention_layernorm(hidden_states))
        return hidden_states


# ----------------------------------------------------------------------------- model
class GPT(nn.Module):
    """Causal LM.  ``forward(input_ids, attention_mask=None, labels=None) -> (logits, loss)``
    with the label shift done inside (reference gpt.py:388-455)."""

    def __init__(self, config: GPTConfig, backend: str = "auto"):
        super().__init__()
        self.config = config
        self.embed_tokens = nn.Embedding(config.vocab_size, config.hidden_size)
        self.layers = nn.ModuleList(TransformerBlock(config, layer_idx=i) for i in range(config.num_layers))
        self.norm = RMSNorm(config.hidden_size)
        self.lm_head = nn.Linear(config.hidden_size, config.vocab_size, bias=False)
        self.lm_head.weight = self.embed_tokens.weight  # weight tying (reference gpt.py:340-343)
        self.gradient_checkpointing = config.gradient_checkpointing
        self.apply(self._init_weights)
        self.backend = backend
        self._engine = None  # FusedGPTEngine, created lazily on CUDA

    def _init_weights(self, module: nn.Module) -> None:
        """N(0, initializer_range) for Linear and Embedding weights (reference gpt.py:380-385)."""
        if isinstance(module, (nn.Linear, nn.Embedding)):
            nn.init.normal_(module.weight, mean=0.0, std=self.config.initializer_range)
            if getattr(module, "bias", None) is not None:
                nn.init.zeros_(module.bias)

    # -- backend selection -----------------------------------------------------------------
    def resolved_backend(self) -> str:
        if self.backend != "auto":
            return self.backend
        return "b200" if self.embed_tokens.weight.is_cuda else "torch"

    def engine(self):
        """The sm_100a engine bound to this model's parameters (CUDA only)."""
        if self._engine is None:
            from ..ops.engine import FusedGPTEngine

            self._engine = FusedGPTEngine(self)
        return self._engine

    # -- forward ---------------------------------------------------------------------------
    def forward(self, input_ids: torch.Tensor, attention_mask: Optional[torch.Tensor] = None,
                labels: Optional[torch.Tensor] = None):
        if self.resolved_backend() == "b200" and not torch.is_autocast_enabled():
            return self.engine().module_forward(input_ids, labels)
        return self.forward_torch(input_ids, attention_mask, labels)

    def forward_torch(self, input_ids, attention_mask=None, labels=None):
        x = self.embed_tokens(input_ids)
        for layer in self.layers:
            if self.gradient_checkpointing and self.training:
                x = _torch_checkpoint(layer, x, attention_mask, use_reentrant=False)
            else:
                x = layer(x, attention_mask)
        logits = self.lm_head(self.norm(x))
        loss = None
        if labels is not None:
            v = self.config.vocab_size
            loss = F.cross_entropy(logits[..., :-1, :].reshape(-1, v), labels[..., 1:].reshape(-1))
        return logits, loss

    # -- generation ------------------------------------------------------------------------
    @torch.no_grad()
    def generate(self, input_ids: torch.Tensor, max_new_tokens: int = 100, temperature: float = 1.0,
                 top_k: int = 50, greedy: bool = False, generator: Optional[torch.Generator] = None):
        """Autoregressive sampling (reference gpt.py:457-484: top-k=50, T=1 multinomial, no KV cache).

        Additions: ``greedy=True`` (or ``top_k=1``) gives the deterministic arg-max
        path used by parity tests; ``generator`` seeds the multinomial draw.
        """
        self.eval()
        limit = self.config.max_seq_len
        for _ in range(max_new_tokens):
            window = input_ids if input_ids.size(1) <= limit else input_ids[:, -limit:]
            if self.resolved_backend() == "b200" and not torch.is_autocast_enabled():
                logits = self.engine().forward_logits(window, last_only=True)   # only the last position's row of the head
            else:
                logits, _ = self(window)
            logits = logits[:, -1, :].float()
            if greedy or top_k == 1:
                nxt = logits.argmax(dim=-1, keepdim=True)
            else:
                logits = logits / temperature
                if top_k > 0:
                    kth = torch.topk(logits, min(top_k, logits.size(-1))).values[:, [-1]]
                    logits = logits.masked_fill(logits < kth, float("-inf"))
                nxt = torch.multinomial(F.softmax(logits, dim=-1), num_samples=1, generator=generator)
            input_ids = torch.cat((input_ids, nxt), dim=1)
        return input_ids


def count_parameters(model: nn.Module) -> int:
    """Trainable parameter count (reference gpt.py:487-489)."""
    return sum(p.numel() for p in model.parameters() if p.requires_grad)


if __name__ == "__main__":  # smoke script, same role as reference gpt.py:492-508
    cfg = GPTConfig.gpt2_small()
    net = GPT(cfg)
    print(f"Model config: {cfg}")
    print(f"Estimated parameters: {cfg.num_parameters():,}")
    print(f"Actual parameters: {count_parameters(net):,}")
    ids = torch.randint(0, cfg.vocab_size, (2, 128))
    out, ls = net(ids, labels=ids.clone())
    print(f"Logits shape: {out.shape}")
    print(f"Loss: {ls.item():.4f}")
