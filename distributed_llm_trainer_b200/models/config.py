"""Model hyper-parameters.

Parity target: reference ``src/models/config.py:6-102`` (``GPTConfig`` and the
``gpt2_small/medium/large/xl`` presets).  Field names and defaults are kept so a
reference user can construct the same objects; what is new here:

* ``exact_num_parameters()`` -- the true trainable count of the architecture the
  reference actually builds (RMSNorm + RoPE + SwiGLU, tied LM head).  The
  reference's ``num_parameters()`` (config.py:81-102) is a GPT-2 style estimate
  that counts a position embedding which does not exist; it is preserved
  verbatim in behaviour for API parity.
* ``from_dict`` / ``to_dict`` so checkpoints can carry plain dicts.
* ``preset(name)`` lookup used by every CLI.
"""
from __future__ import annotations

from dataclasses import asdict, dataclass, fields
from typing import Any, Dict, Optional

_PRESETS = {
    # name: (hidden, layers, heads)
    "small": (768, 12, 12),
    "medium": (1024, 24, 16),
    "large": (1280, 36, 20),
    "xl": (1600, 48, 25),
}


@dataclass
class GPTConfig:
    vocab_size: int = 50257
    hidden_size: int = 768
    num_layers: int = 12
    num_heads: int = 12
    intermediate_size: Optional[int] = None
    max_seq_len: int = 1024
    dropout: float = 0.1
    attention_dropout: float = 0.1
    initializer_range: float = 0.02
    activation: str = "gelu"  # dead field in the reference (config.py:28); SwiGLU is what runs
    use_flash_attention: bool = False
    gradient_checkpointing: bool = False

    def __post_init__(self) -> None:
        if self.intermediate_size is None:
            self.intermediate_size = 4 * self.hidden_size
        if self.hidden_size % self.num_heads != 0:
            raise AssertionError(
                f"hidden_size ({self.hidden_size}) must be divisible by num_heads ({self.num_heads})"
            )

    # ---- presets (reference config.py:41-79) ---------------------------------
    @classmethod
    def preset(cls, name: str, **overrides: Any) -> "GPTConfig":
        if name not in _PRESETS:
            raise KeyError(f"unknown model size {name!r}; choose from {sorted(_PRESETS)}")
        h, l, a = _PRESETS[name]
        return cls(hidden_size=h, num_layers=l, num_heads=a, **overrides)

    @classmethod
    def gpt2_small(cls) -> "GPTConfig":
        return cls.preset("small")

    @classmethod
    def gpt2_medium(cls) -> "GPTConfig":
        return cls.preset("medium")

    @classmethod
    def gpt2_large(cls) -> "GPTConfig":
        return cls.preset("large")

    @classmethod
    def gpt2_xl(cls) -> "GPTConfig":
        return cls.preset("xl")

    # ---- derived quantities --------------------------------------------------
    @property
    def head_dim(self) -> int:
        return self.hidden_size // self.num_heads

    def num_parameters(self) -> int:
        """GPT-2 style estimate, bug-compatible with reference config.py:81-102."""
        h, i = self.hidden_size, self.intermediate_size
        per_layer = 4 * h * h + 2 * h * i + 4 * h
        return self.vocab_size * h + self.max_seq_len * h + self.num_layers * per_layer + 2 * h

    def exact_num_parameters(self) -> int:
        """Trainable parameters of the model that is really built (tied head counted once)."""
        h, i = self.hidden_size, self.intermediate_size
        per_layer = 4 * h * h + 3 * h * i + 2 * h
        return self.vocab_size * h + self.num_layers * per_layer + h

    def flops_per_token(self, seq_len: Optional[int] = None, causal: bool = True,
                        recompute: bool = False) -> float:
        """Training FLOPs per token (fwd + bwd), used for roofline reporting."""
        s = seq_len or self.max_seq_len
        h, i, l, v = self.hidden_size, self.intermediate_size, self.num_layers, self.vocab_size
        gemm = 2 * (l * (4 * h * h + 3 * h * i) + v * h)          # fwd MACs*2
        attn = l * 4 * s * h * (0.5 if causal else 1.0)            # QK^T + PV fwd
        fwd = gemm + attn
        return fwd * (4.0 if recompute else 3.0)

    # ---- (de)serialisation ---------------------------------------------------
    def to_dict(self) -> Dict[str, Any]:
        return asdict(self)

    @classmethod
    def from_dict(cls, d: Dict[str, Any]) -> "GPTConfig":
        known = {f.name for f in fields(cls)}
        return cls(**{k: v for k, v in d.items() if k in known})

    @classmethod
    def from_any(cls, obj: Any) -> "GPTConfig":
        """Accept our config, a dict, or a (possibly foreign) object with the same attributes."""
        if isinstance(obj, cls):
            return obj
        if isinstance(obj, dict):
            return cls.from_dict(obj)
        return cls.from_dict({f.name: getattr(obj, f.name) for f in fields(cls) if hasattr(obj, f.name)})
