"""Tokenizer access with an offline fallback.

The reference hard-codes ``GPT2TokenizerFast.from_pretrained("gpt2")``
(tinystories.py:27, openwebtext.py:29, infer.py:60).  On a machine with no HF cache that call
either raises or silently returns an *empty* tokenizer, which crashes
generation.  ``load_tokenizer`` tries the HF tokenizer first and, if it is unusable, falls back
to a byte-level tokenizer whose ids (0..255) are valid rows of any GPT-2 sized embedding.
"""
from __future__ import annotations

from typing import List, Sequence


class ByteTokenizer:
    """UTF-8 bytes as token ids; stand-in when GPT-2 vocab files are not available offline."""

    name = "byte-fallback"
    vocab_size = 256
    model_max_length = 10**9

    def encode(self, text: str, return_tensors: str | None = None):
        ids = list(text.encode("utf-8"))
        if return_tensors == "pt":
            import torch

            return torch.tensor([ids], dtype=torch.long)
        return ids

    def decode(self, ids: Sequence[int]) -> str:
        out: List[int] = []
        for t in ids:
            t = int(t)
            out.append(t if t < 256 else 63)  # '?' for ids outside the byte range
        return bytes(out).decode("utf-8", errors="replace")


def load_tokenizer(name: str = "gpt2", allow_fallback: bool = True):
    try:
        from transformers import GPT2TokenizerFast

        tok = GPT2TokenizerFast.from_pretrained(name)
        if len(tok.encode("hello world")) > 0:
            tok.model_max_length = 10**9
            return tok
    except Exception:
        pass
    if not allow_fallback:
        raise RuntimeError(f"tokenizer {name!r} is not available offline")
    return ByteTokenizer()
