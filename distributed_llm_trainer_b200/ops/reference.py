"""Plain-PyTorch (fp32) statements of every fused op -- the oracles the CUDA kernels are tested
against, each a restatement of the cited reference lines (``/root/reference/src/models/gpt.py``)."""
from __future__ import annotations

import torch
import torch.nn.functional as F


def rope_tables(seq_len: int, head_dim: int = 64, base: float = 10000.0, device="cpu"):
    """cos/sin [S, head_dim/2] fp32 (gpt.py:80-93; the reference stores the two identical halves)."""
    inv = 1.0 / (base ** (torch.arange(0, head_dim, 2, dtype=torch.float32) / head_dim))
    ang = torch.outer(torch.arange(seq_len, dtype=torch.float32), inv)
    return ang.cos().contiguous().to(device), ang.sin().contiguous().to(device)


def rope_packed(qkv: torch.Tensor, cos: torch.Tensor, sin: torch.Tensor, seq_len: int, rope_cols: int,
                inverse: bool = False) -> torch.Tensor:
    """Rotary on the first ``rope_cols`` columns of a packed [rows, 3h] tensor (gpt.py:116-147)."""
    rows = qkv.shape[0]
    out = qkv.clone().float()
    pos = torch.arange(rows, device=qkv.device) % seq_len
    c, s = cos[pos][:, None, :], sin[pos][:, None, :]
    if inverse:
        s = -s
    part = out[:, :rope_cols].reshape(rows, rope_cols // 64, 64)
    x1, x2 = part[..., :32], part[..., 32:]
    out[:, :rope_cols] = torch.cat((x1 * c - x2 * s, x2 * c + x1 * s), dim=-1).reshape(rows, rope_cols)
    return out


def rmsnorm(x: torch.Tensor, w: torch.Tensor, eps: float = 1e-6):
    """gpt.py:66-67."""
    r = torch.rsqrt(x.float().pow(2).mean(-1, keepdim=True) + eps)
    return x.float() * r * w.float(), r.squeeze(-1)


def causal_attention(q, k, v, dropout_mask=None, p: float = 0.0):
    """Manual path of gpt.py:230-234 on [B,H,S,D] fp32 tensors; ``dropout_mask`` is a keep-mask."""
    d = q.shape[-1]
    s = q.shape[-2]
    scores = (q @ k.transpose(-2, -1)) / (d ** 0.5)
    future = torch.ones(s, s, dtype=torch.bool, device=q.device).triu(1)
    probs = F.softmax(scores.masked_fill(future, float("-inf")), dim=-1)
    if dropout_mask is not None:
        probs = probs * dropout_mask / (1.0 - p)
    return probs @ v


def swiglu(g, u):
    """gpt.py:278-281."""
    return F.silu(g) * u


def cross_entropy_shifted(logits: torch.Tensor, labels: torch.Tensor):
    """gpt.py:449-453: predict token t+1 from position t, mean over B*(S-1) positions."""
    v = logits.shape[-1]
    return F.cross_entropy(logits[..., :-1, :].reshape(-1, v).float(), labels[..., 1:].reshape(-1))
