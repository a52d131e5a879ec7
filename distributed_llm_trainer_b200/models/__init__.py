from .config import GPTConfig
from .gpt import (GPT, MLP, CausalSelfAttention, RMSNorm, RotaryPositionEmbedding, TransformerBlock,
                  apply_rotary_pos_emb, count_parameters, rotate_half)

__all__ = [n for n in dir() if not n.startswith("_")]
