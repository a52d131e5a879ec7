#!/usr/bin/env python3
"""Headline benchmark: tokens/sec of the GPT-2 "124M" DDP training step (bf16, micro-batch 8 x seq 1024,
gradient accumulation 4 => 32,768 tokens per GPU per optimizer step), synthetic tokens, random init.

    python bench.py --gpus 1 --steps 10 --warmup 3
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        bench.py --gpus N --steps K --warmup W
    python bench.py --impl reference ...      # the unmodified reference from baseline/_ref, same metric/config

Prints ONE JSON line on rank 0.  ``value`` is device-timed (CUDA events around exactly K optimizer steps,
barrier + synchronize on both sides, max over ranks) with inputs already on the device; ``e2e.value`` is the
same metric through the public API (``DistributedTrainer.train_step``) with every step's batch copied from
pinned host memory and the step's loss read back to the host.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))

MICRO_BATCH, SEQ, GA = 8, 1024, 4


def parse():
    p = argparse.ArgumentParser()
    p.add_argument("--gpus", type=int, default=1)
    p.add_argument("--steps", type=int, default=10)
    p.add_argument("--warmup", type=int, default=3)
    p.add_argument("--impl", type=str, default="b200", choices=["b200", "reference"])
    p.add_argument("--model_size", type=str, default="small")
    p.add_argument("--no_graphs", action="store_true")
    return p.parse_args()


def dist_info():
    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    return rank, world, local


def max_over_ranks(x: float, device) -> float:
    import torch
    import torch.distributed as dist

    if dist.is_available() and dist.is_initialized():
        t = torch.tensor([x], dtype=torch.float64, device=device)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())
    return x


def barrier_sync():
    import torch
    import torch.distributed as dist

    if dist.is_available() and dist.is_initialized():
        dist.barrier()
    torch.cuda.synchronize()


def timed_steps(step_fn, steps: int, device) -> float:
    """ms for exactly `steps` calls of step_fn, device-timed, max over ranks."""
    import torch

    barrier_sync()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(steps):
        step_fn(i)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    barrier_sync()
    return max_over_ranks(ms, device)


def make_host_batches(n: int, vocab: int, seed: int):
    import torch

    g = torch.Generator().manual_seed(seed)
    return [torch.randint(0, vocab, (MICRO_BATCH * GA, SEQ), generator=g, dtype=torch.long).pin_memory()
            for _ in range(n)]


# ------------------------------------------------------------------------------------------------ our arm
def run_b200(args) -> dict:
    sys.path.insert(0, ROOT)
    import torch

    from distributed_llm_trainer_b200.models import GPTConfig
    from distributed_llm_trainer_b200.ops import _lib
    from distributed_llm_trainer_b200.ops import kernels as K
    from distributed_llm_trainer_b200.training.config import TrainingConfig
    from distributed_llm_trainer_b200.training.ddp_trainer import DistributedTrainer
    from distributed_llm_trainer_b200.utils.timing import ClockSampler

    rank, world, local = dist_info()
    if not _lib.available():
        if rank == 0:
            _lib.build()
    model_cfg = GPTConfig.preset(args.model_size)          # dropout 0.1 / attention dropout 0.1 as the reference runs
    train_cfg = TrainingConfig(batch_size=MICRO_BATCH, gradient_accumulation_steps=GA, max_steps=1000,
                               mixed_precision="bf16", seed=1234)
    trainer = DistributedTrainer(model_cfg, train_cfg)
    assert trainer.path == "b200", "bench must run the native sm_100a path"
    if args.no_graphs:
        trainer.engine.use_graphs = False
    dev = trainer.device
    host = make_host_batches(4, model_cfg.vocab_size, 99 + rank)
    dev_batches = [b.to(dev) for b in host]

    def step_device(i):   # inputs resident on the device; loss stays on the device (synced once at the end)
        trainer.train_step({"input_ids": dev_batches[i % len(dev_batches)]}, sync=False)

    def step_e2e(i):      # public API: pinned host batch -> H2D -> step -> loss D2H
        out = trainer.train_step({"input_ids": host[i % len(host)]})
        losses.append(out["loss"])

    losses = []
    for i in range(max(args.warmup, 3)):
        step_device(i)
    before = K.launch_count
    with ClockSampler(gpu_index=local) as clocks:
        ms = timed_steps(step_device, args.steps, dev)
    launches = K.launch_count - before
    for i in range(2):
        step_e2e(i)
    ms_e2e = timed_steps(step_e2e, args.steps, dev)
    tokens_per_step = MICRO_BATCH * GA * SEQ * world
    value = tokens_per_step * args.steps / (ms / 1e3)
    e2e = tokens_per_step * args.steps / (ms_e2e / 1e3)
    flops_tok = model_cfg.flops_per_token(SEQ, causal=True)
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    sustained = peaks.get("bf16_tflops_sustained", 1400.0) * 1e12
    return {
        "metric": "tokens/sec GPT-2 124M DDP (bf16, batch 8 x seq 1024, grad-accum 4)",
        "value": value, "unit": "tokens/s", "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
        "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "bf16", "data": "synthetic", "impl": "b200",
        "config": {"model": "gpt2-small (151.9M params: RMSNorm+RoPE+SwiGLU, tied head; reference preset 'small')",
                   "global_batch": MICRO_BATCH * GA * world, "micro_batch": MICRO_BATCH, "grad_accum": GA,
                   "seq_len": SEQ, "parallelism": f"dp{world}", "dropout": model_cfg.dropout,
                   "l2": "no flush: one step streams ~4 GB (weights, grads, activations) >> 126 MB L2",
                   "cuda_graphs": bool(trainer.engine.use_graphs)},
        "clocks": clocks.summary(),
        "e2e": {"value": e2e, "unit": "tokens/s", "ms_per_step": ms_e2e / args.steps,
                "h2d_bytes_per_step": MICRO_BATCH * GA * SEQ * 8, "d2h_bytes_per_step": 4},
        "gpu_launches": launches,
        "mfu_of_measured_sustained": value / world * flops_tok / sustained,
        "final_loss": losses[-1] if losses else None,
    }


# ------------------------------------------------------------------------------------------------ reference arm
def run_reference(args) -> dict:
    ref = os.path.join(ROOT, "baseline", "_ref")
    if not os.path.isdir(os.path.join(ref, "training")):
        return {"impl": "reference", "unavailable": "baseline/_ref is missing (pip --target install of /root/reference "
                                                    "was not done on this checkout)"}
    sys.path.insert(0, ref)
    import torch

    try:
        from models.config import GPTConfig          # the reference's own modules, unmodified
        from training.ddp_trainer import DistributedTrainer, TrainingConfig
    except Exception as e:  # pragma: no cover
        return {"impl": "reference", "unavailable": f"import failed: {type(e).__name__}: {e}"[:200]}
    rank, world, local = dist_info()
    sizes = {"small": GPTConfig.gpt2_small, "medium": GPTConfig.gpt2_medium, "large": GPTConfig.gpt2_large}
    model_cfg = sizes[args.model_size]()
    # exactly what its CLI builds: TrainingConfig(batch_size, max_steps, mixed_precision), everything else default
    train_cfg = TrainingConfig(batch_size=MICRO_BATCH, max_steps=1000, mixed_precision="bf16")
    torch.manual_seed(1234)
    trainer = DistributedTrainer(model_cfg, train_cfg)
    dev = trainer.device
    host = make_host_batches(4, model_cfg.vocab_size, 99 + rank)
    losses = []

    def step(i):
        out = trainer.train_step({"input_ids": host[i % len(host)]})
        losses.append(out["loss"])

    sys.path.insert(0, ROOT)
    from distributed_llm_trainer_b200.utils.timing import ClockSampler   # clock sampling only; not on the timed path

    for i in range(max(args.warmup, 3)):
        step(i)
    with ClockSampler(gpu_index=local) as clocks:
        ms = timed_steps(step, args.steps, dev)
    tokens_per_step = MICRO_BATCH * GA * SEQ * world
    value = tokens_per_step * args.steps / (ms / 1e3)
    return {
        "metric": "tokens/sec GPT-2 124M DDP (bf16, batch 8 x seq 1024, grad-accum 4)",
        "value": value, "unit": "tokens/s", "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
        "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "bf16", "data": "synthetic", "impl": "reference",
        "config": {"model": "gpt2-small (reference GPTConfig.gpt2_small(), as run from its CLI: manual attention)",
                   "global_batch": MICRO_BATCH * GA * world, "micro_batch": MICRO_BATCH, "grad_accum": GA,
                   "seq_len": SEQ, "parallelism": f"dp{world}", "path": "torch DDP + autocast + fused AdamW (stock)"},
        "clocks": clocks.summary(),
        "e2e": {"value": value, "unit": "tokens/s", "h2d_bytes_per_step": MICRO_BATCH * GA * SEQ * 8,
                "d2h_bytes_per_step": 4 * GA,
                "note": "the reference's train_step already copies from the host batch and reads the loss back"},
        "gpu_launches": 0,
        "final_loss": losses[-1] if losses else None,
    }


def main():
    args = parse()
    rank, world, _ = dist_info()
    if world != args.gpus and world == 1 and args.gpus > 1:
        print(json.dumps({"error": f"--gpus {args.gpus} needs torchrun (WORLD_SIZE={world})"}))
        sys.exit(2)
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    res = run_reference(args) if args.impl == "reference" else run_b200(args)
    import torch.distributed as dist

    if dist.is_available() and dist.is_initialized():
        dist.barrier()
        dist.destroy_process_group()
    if rank == 0:
        print(json.dumps(res))


if __name__ == "__main__":
    main()
