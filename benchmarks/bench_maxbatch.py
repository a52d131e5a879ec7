#!/usr/bin/env python3
"""BASELINE config 4: GPT-2 "355M" FSDP bf16 + gradient (activation) checkpointing at the LARGEST per-GPU micro-batch
that fits, gradient accumulation 1.  The micro-batch doubles from --start until a step fails with an out-of-memory
error (caught; the process keeps running), then the last size that worked is timed.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        benchmarks/bench_maxbatch.py [--impl reference] [--limit 512]
"""
import argparse
import gc
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from bench import SEQ, dist_info, max_over_ranks, timed_steps  # noqa: E402


def build(impl, micro):
    if impl == "reference":
        if os.path.join(ROOT, "baseline", "_ref") not in sys.path:
            sys.path.insert(0, os.path.join(ROOT, "baseline", "_ref"))
        from models.config import GPTConfig
        from training.fsdp_trainer import FSDPConfig, FSDPTrainer, TrainingConfig

        mc = GPTConfig.gpt2_medium()
        tc = TrainingConfig(batch_size=micro, max_steps=1000, gradient_accumulation_steps=1)
        return FSDPTrainer(mc, tc, FSDPConfig(sharding_strategy="FULL_SHARD", activation_checkpointing=True)), mc
    from distributed_llm_trainer_b200.models import GPTConfig
    from distributed_llm_trainer_b200.training.config import FSDPConfig, TrainingConfig
    from distributed_llm_trainer_b200.training.fsdp_trainer import FSDPTrainer

    mc = GPTConfig.preset("medium")
    tc = TrainingConfig.fsdp_defaults(batch_size=micro, max_steps=1000, seed=1234, gradient_accumulation_steps=1)
    return FSDPTrainer(mc, tc, FSDPConfig(sharding_strategy="FULL_SHARD", activation_checkpointing=True)), mc


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--start", type=int, default=16)
    ap.add_argument("--limit", type=int, default=128)   # 128 x 1024 rows keeps every row*column product < 2^31
    ap.add_argument("--steps", type=int, default=3)
    args = ap.parse_args()
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    import torch
    import torch.distributed as dist

    rank, world, local = dist_info()
    if "RANK" not in os.environ:
        os.environ.update(RANK="0", WORLD_SIZE="1", LOCAL_RANK="0", MASTER_PORT="29655")
    tried, best = [], None
    micro = args.start
    while micro <= args.limit:
        trainer = None
        ok = 1
        try:
            trainer, mc = build(args.impl, micro)
            g = torch.Generator().manual_seed(3 + rank)
            host = torch.randint(0, mc.vocab_size, (micro, SEQ), generator=g, dtype=torch.long).pin_memory()
            for _ in range(2):
                trainer.train_step({"input_ids": host})
            torch.cuda.synchronize()
        except Exception as e:  # noqa: BLE001  (torch OOM, or cudaMalloc failing inside the native library)
            msg = str(e).lower()
            if "out of memory" not in msg and "memory" not in msg:
                raise
            ok = 0
        flag = torch.tensor([ok], device=f"cuda:{local}")
        if dist.is_initialized():
            dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        ok = int(flag.item())
        tried.append((micro, bool(ok)))
        if ok:
            torch.cuda.reset_peak_memory_stats()
            ms = timed_steps(lambda i: trainer.train_step({"input_ids": host}), args.steps, trainer.device)
            free_b, total_b = torch.cuda.mem_get_info()
            best = {"micro_batch": micro, "ms_per_step": ms / args.steps,
                    "tokens_per_s": micro * SEQ * world * args.steps / (ms / 1e3),
                    "device_mem_used_gb": max_over_ranks((total_b - free_b) / 1e9, trainer.device),
                    "peak_allocator_gb": max_over_ranks(torch.cuda.max_memory_allocated() / 1e9, trainer.device)}
        if trainer is not None:
            sh = getattr(trainer, "sharded", None)
            if sh is not None and ok:
                sh.close()
            for k in list(vars(trainer)):
                setattr(trainer, k, None)
        trainer = None
        gc.collect()
        torch.cuda.empty_cache()
        if not ok:
            break
        micro *= 2
    if rank == 0:
        print(json.dumps({"impl": args.impl, "n_gpus": world, "model": "gpt2-medium FSDP FULL_SHARD + activation checkpointing",
                          "seq_len": SEQ, "grad_accum": 1, "tried": tried, "search_limit": args.limit, "best": best}))
    if dist.is_initialized():
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
