"""Generate text from a trained checkpoint.

Parity target: reference ``src/eval/infer.py`` (flags ``--checkpoint --model_size --prompt
--max_new_tokens --device``; top-k=50 / T=1 multinomial sampling through ``GPT.generate``).
Additions: ``--greedy --top_k --temperature --seed --prompt_ids``, model shape read from
the checkpoint's ``model_config`` when present, both DDP and FSDP checkpoints load (the reference's
own FSDP checkpoints do not load in its infer.py), offline tokenizer fallback.
"""
from __future__ import annotations

import argparse
from typing import List, Optional

import torch

from ..data.tokenizer import load_tokenizer
from ..models import GPT, GPTConfig
from ..utils.checkpoint import load_checkpoint_file


def pick_device(arg: str) -> torch.device:
    if arg != "auto":
        return torch.device(arg)
    if torch.cuda.is_available():
        return torch.device("cuda")
    return torch.device("cpu")


def load_model(checkpoint: str, model_size: Optional[str], device: torch.device) -> GPT:
    ckpt = load_checkpoint_file(checkpoint, map_location="cpu")
    cfg = None
    if model_size is None and ckpt.get("model_config") is not None:
        cfg = GPTConfig.from_any(ckpt["model_config"])
    if cfg is None:
        cfg = GPTConfig.preset(model_size or "small")
    model = GPT(cfg)
    model.load_state_dict(ckpt["model"])   # strict, as the reference
    model = model.to(device)
    if device.type == "cuda":
        model.engine()                      # binds the bf16 shadow to the loaded weights
    model.eval()
    return model


def main(argv: Optional[List[str]] = None) -> None:
    p = argparse.ArgumentParser()
    p.add_argument("--checkpoint", type=str, required=True)
    p.add_argument("--model_size", type=str, default=None, choices=["small", "medium", "large", "xl"])
    p.add_argument("--prompt", type=str, default="Once upon a time,")
    p.add_argument("--prompt_ids", type=str, default=None, help="comma separated token ids (skips the tokenizer)")
    p.add_argument("--max_new_tokens", type=int, default=50)
    p.add_argument("--device", type=str, default="auto")
    p.add_argument("--greedy", action="store_true")
    p.add_argument("--top_k", type=int, default=50)
    p.add_argument("--temperature", type=float, default=1.0)
    p.add_argument("--seed", type=int, default=None)
    args = p.parse_args(argv)

    device = pick_device(args.device)
    model = load_model(args.checkpoint, args.model_size, device)
    tok = load_tokenizer("gpt2")
    if args.prompt_ids:
        ids = torch.tensor([[int(t) for t in args.prompt_ids.split(",")]], dtype=torch.long, device=device)
    else:
        ids = tok.encode(args.prompt, return_tensors="pt").to(device)
    gen = None
    if args.seed is not None:
        gen = torch.Generator(device=device)
        gen.manual_seed(args.seed)
    out = model.generate(ids, max_new_tokens=args.max_new_tokens, temperature=args.temperature, top_k=args.top_k,
                         greedy=args.greedy, generator=gen)
    print("token_ids:", out[0].tolist())
    print(tok.decode(out[0].tolist()))


if __name__ == "__main__":
    main()
