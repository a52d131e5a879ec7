#!/usr/bin/env python3
"""Second headline config: tokens/sec of the GPT-2 "355M" (medium preset, 454 M params) FSDP training step --
bf16, micro-batch 4 x seq 1024, gradient accumulation 8 (32,768 tokens / GPU / step), FULL_SHARD, activation
checkpointing on (both trainers' default).  Same timing rules and JSON shape as bench.py.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        benchmarks/bench_fsdp.py --gpus N [--impl reference] [--sharding SHARD_GRAD_OP] [--no_activation_checkpointing]
"""
from __future__ import annotations

import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from bench import barrier_sync, dist_info, max_over_ranks, timed_steps  # noqa: E402

MICRO_BATCH, SEQ, GA = 4, 1024, 8


def host_batches(n, vocab, seed):
    import torch

    g = torch.Generator().manual_seed(seed)
    return [torch.randint(0, vocab, (MICRO_BATCH * GA, SEQ), generator=g, dtype=torch.long).pin_memory() for _ in range(n)]


def main():
    global MICRO_BATCH
    p = argparse.ArgumentParser()
    p.add_argument("--gpus", type=int, default=1)
    p.add_argument("--steps", type=int, default=5)
    p.add_argument("--warmup", type=int, default=3)
    p.add_argument("--impl", default="b200", choices=["b200", "reference"])
    p.add_argument("--model_size", default="medium")
    p.add_argument("--sharding", default="FULL_SHARD")
    p.add_argument("--no_activation_checkpointing", action="store_true")
    p.add_argument("--batch_size", type=int, default=MICRO_BATCH)
    args = p.parse_args()
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    rank, world, local = dist_info()
    import torch

    MICRO_BATCH = args.batch_size
    if args.impl == "reference":
        sys.path.insert(0, os.path.join(ROOT, "baseline", "_ref"))
        from models.config import GPTConfig
        from training.fsdp_trainer import FSDPConfig, FSDPTrainer, TrainingConfig

        mc = {"small": GPTConfig.gpt2_small, "medium": GPTConfig.gpt2_medium, "large": GPTConfig.gpt2_large,
              "xl": GPTConfig.gpt2_xl}[args.model_size]()
        tc = TrainingConfig(batch_size=MICRO_BATCH, max_steps=1000)
        fc = FSDPConfig(sharding_strategy=args.sharding, activation_checkpointing=not args.no_activation_checkpointing)
        trainer = FSDPTrainer(mc, tc, fc)
    else:
        from distributed_llm_trainer_b200.models import GPTConfig
        from distributed_llm_trainer_b200.training.config import FSDPConfig, TrainingConfig
        from distributed_llm_trainer_b200.training.fsdp_trainer import FSDPTrainer

        mc = GPTConfig.preset(args.model_size)
        tc = TrainingConfig.fsdp_defaults(batch_size=MICRO_BATCH, max_steps=1000, seed=1234)
        fc = FSDPConfig(sharding_strategy=args.sharding, activation_checkpointing=not args.no_activation_checkpointing)
        trainer = FSDPTrainer(mc, tc, fc)
        assert trainer.path == "b200"
    dev = trainer.device
    host = host_batches(4, mc.vocab_size, 7 + rank)
    losses = []

    def step(i):
        losses.append(trainer.train_step({"input_ids": host[i % len(host)]})["loss"])

    from distributed_llm_trainer_b200.utils.timing import ClockSampler

    for i in range(max(args.warmup, 3)):
        step(i)
    torch.cuda.reset_peak_memory_stats()
    launches = 0
    if args.impl == "b200":
        from distributed_llm_trainer_b200.ops import kernels as K

        before = K.launch_count
    with ClockSampler(gpu_index=local) as clocks:
        ms = timed_steps(step, args.steps, dev)
    if args.impl == "b200":
        launches = K.launch_count - before
    tokens = MICRO_BATCH * GA * SEQ * world
    value = tokens * args.steps / (ms / 1e3)
    peak_gb = max_over_ranks(torch.cuda.max_memory_allocated() / 1e9, dev)
    res = {
        "metric": "tokens/sec GPT-2 355M FSDP (bf16, batch 4 x seq 1024, grad-accum 8)", "value": value,
        "unit": "tokens/s", "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
        "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "bf16", "data": "synthetic", "impl": args.impl,
        "config": {"model": f"gpt2-{args.model_size}", "global_batch": MICRO_BATCH * GA * world, "micro_batch": MICRO_BATCH,
                   "grad_accum": GA, "seq_len": SEQ, "parallelism": f"fsdp{world}:{args.sharding}",
                   "activation_checkpointing": not args.no_activation_checkpointing},
        "clocks": clocks.summary(), "gpu_launches": launches, "peak_mem_gb_torch_allocator": peak_gb,
        "e2e": {"value": value, "unit": "tokens/s", "h2d_bytes_per_step": MICRO_BATCH * GA * SEQ * 8,
                "d2h_bytes_per_step": 4 if args.impl == "b200" else 4 * GA,
                "note": "train_step copies the pinned host batch and reads the loss back"},
        "final_loss": losses[-1],
    }
    import torch.distributed as dist

    if dist.is_initialized():
        dist.barrier()
        dist.destroy_process_group()
    if rank == 0:
        print(json.dumps(res))


if __name__ == "__main__":
    main()
