#!/usr/bin/env python3
"""Headline benchmark: tokens/sec of the GPT-2 "124M" DDP training step (bf16, micro-batch 8 x seq 1024,
gradient accumulation 4 => 32,768 tokens per GPU per optimizer step), synthetic tokens, random init.

    python bench.py --gpus 1 --steps 10 --warmup 3
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        bench.py --gpus N --steps K --warmup W
    python bench.py --impl reference ...      # the unmodified reference from baseline/_ref, same metric/config

After the DDP measurement the same process (group) measures the second half of the BASELINE.json metric, GPT-2
"355M" FSDP FULL_SHARD (bf16, micro-batch 4 x seq 1024, grad-accum 8, activation checkpointing), through the
arm's own ``FSDPTrainer.train_step`` and reports it as the ``fsdp`` sub-record of the same JSON line
(``--skip_fsdp`` leaves it out).

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
    p.add_argument("--skip_fsdp", action="store_true", help="DDP headline only (no 355M FSDP sub-record)")
    p.add_argument("--fsdp_model_size", type=str, default="medium")
    p.add_argument("--ref_sdpa", action="store_true",
                   help="reference arm only: build its model with use_flash_attention=True (F.scaled_dot_product_attention); "
                        "a second, non-headline reference row -- the reference's CLI never sets this flag")
    return p.parse_args()


MODEL_NAMES = {"small": "gpt2-small", "medium": "gpt2-medium", "large": "gpt2-large", "xl": "gpt2-xl"}
L2_NOTE = "no flush: every step streams far more than the 126 MB L2 (weights, gradients, activations)"


def ddp_config(model_size: str, n_params: int, dropout: float, world: int) -> dict:
    """The headline config, key-for-key identical in both arms (prose goes into ``notes``)."""
    return {"model": MODEL_NAMES[model_size], "params": n_params, "global_batch": MICRO_BATCH * GA * world,
            "micro_batch": MICRO_BATCH, "grad_accum": GA, "seq_len": SEQ, "parallelism": f"dp{world}",
            "dropout": dropout, "l2": L2_NOTE}


FSDP_MICRO, FSDP_GA = 4, 8


def fsdp_config_dict(model_size: str, n_params: int, dropout: float, world: int) -> dict:
    return {"model": MODEL_NAMES[model_size], "params": n_params, "global_batch": FSDP_MICRO * FSDP_GA * world,
            "micro_batch": FSDP_MICRO, "grad_accum": FSDP_GA, "seq_len": SEQ,
            "parallelism": f"fsdp{world}:FULL_SHARD", "activation_checkpointing": True, "dropout": dropout,
            "l2": L2_NOTE}


def device_mem_used_gb(device) -> float:
    import torch

    free_b, total_b = torch.cuda.mem_get_info(device)
    return (total_b - free_b) / 1e9


def measure_fsdp(trainer, vocab: int, steps: int, warmup: int, rank: int, local: int, impl: str) -> dict:
    """Times ``trainer.train_step`` (pinned host batch in, loss float out: the public API already contains the H2D copy
    and the D2H read), device-timed, max over ranks."""
    import torch

    sys.path.insert(0, ROOT)
    from distributed_llm_trainer_b200.utils.timing import ClockSampler

    g = torch.Generator().manual_seed(7 + rank)
    host = [torch.randint(0, vocab, (FSDP_MICRO * FSDP_GA, SEQ), generator=g, dtype=torch.long).pin_memory()
            for _ in range(4)]
    losses = []

    def step(i):
        losses.append(trainer.train_step({"input_ids": host[i % len(host)]})["loss"])

    for i in range(warmup):
        step(i)
    torch.cuda.reset_peak_memory_stats()
    launches = 0
    if impl == "b200":
        from distributed_llm_trainer_b200.ops import kernels as K

        before = K.launch_count
    with ClockSampler(gpu_index=local) as clocks:
        ms = timed_steps(step, steps, trainer.device)
    if impl == "b200":
        launches = K.launch_count - before
    world = int(os.environ.get("WORLD_SIZE", 1))
    tokens = FSDP_MICRO * FSDP_GA * SEQ * world
    value = tokens * steps / (ms / 1e3)
    return {
        "metric": "tokens/sec GPT-2 355M FSDP (bf16, batch 4 x seq 1024, grad-accum 8, FULL_SHARD, activation checkpointing)",
        "value": value, "unit": "tokens/s", "steps": steps, "warmup": warmup, "ms_per_step": ms / steps,
        "e2e": {"value": value, "unit": "tokens/s", "h2d_bytes_per_step": FSDP_MICRO * FSDP_GA * SEQ * 8,
                "d2h_bytes_per_step": 4 if impl == "b200" else 4 * FSDP_GA,
                "note": "train_step copies the pinned host batch to the device and reads the loss back"},
        "peak_mem_gb_torch_allocator": max_over_ranks(torch.cuda.max_memory_allocated() / 1e9, trainer.device),
        "device_mem_used_gb": max_over_ranks(device_mem_used_gb(trainer.device), trainer.device),
        "clocks": clocks.summary(), "gpu_launches": launches, "final_loss": losses[-1],
    }


def release(trainer) -> None:
    """Drop the DDP trainer's device memory before the FSDP measurement."""
    import gc

    import torch

    for owner in ("grad_sync", "sharded"):
        close = getattr(getattr(trainer, owner, None), "close", None)
        if close is not None:
            close()
    for k in list(vars(trainer)):
        setattr(trainer, k, None)
    gc.collect()
    torch.cuda.empty_cache()


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
    n_params = sum(p.numel() for p in trainer.raw_model.parameters())
    cuda_graphs = bool(trainer.engine.use_graphs)
    fsdp = None
    if not args.skip_fsdp:
        release(trainer)
        from distributed_llm_trainer_b200.training.config import FSDPConfig
        from distributed_llm_trainer_b200.training.fsdp_trainer import FSDPTrainer

        fmc = GPTConfig.preset(args.fsdp_model_size)
        ftc = TrainingConfig.fsdp_defaults(batch_size=FSDP_MICRO, max_steps=1000, seed=1234)
        ftr = FSDPTrainer(fmc, ftc, FSDPConfig(sharding_strategy="FULL_SHARD", activation_checkpointing=True))
        assert ftr.path == "b200", "bench must run the native sm_100a path"
        fsdp = measure_fsdp(ftr, fmc.vocab_size, max(3, args.steps // 2), 3, rank, local, "b200")
        fsdp["config"] = fsdp_config_dict(args.fsdp_model_size, fmc.exact_num_parameters(), fmc.dropout, world)
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    sustained = peaks.get("bf16_tflops_sustained", 1400.0) * 1e12
    if fsdp is not None:
        fsdp["mfu_of_measured_sustained"] = (fsdp["value"] / world * fmc.flops_per_token(SEQ, causal=True, recompute=True)
                                             / sustained)
    return {
        "metric": "tokens/sec GPT-2 124M DDP (bf16, batch 8 x seq 1024, grad-accum 4)",
        "value": value, "unit": "tokens/s", "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
        "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "bf16", "data": "synthetic", "impl": "b200",
        "config": ddp_config(args.model_size, n_params, model_cfg.dropout, world),
        "notes": {"model": "reference preset: RMSNorm + RoPE + SwiGLU, tied head", "cuda_graphs": cuda_graphs,
                  "path": "hand-written sm_100a engine; wgrad GEMMs reduce-add into the owners over NVLink"},
        "clocks": clocks.summary(),
        "e2e": {"value": e2e, "unit": "tokens/s", "ms_per_step": ms_e2e / args.steps,
                "h2d_bytes_per_step": MICRO_BATCH * GA * SEQ * 8, "d2h_bytes_per_step": 4},
        "gpu_launches": launches,
        "mfu_of_measured_sustained": value / world * flops_tok / sustained,
        "final_loss": losses[-1] if losses else None,
        "fsdp": fsdp,
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
    if args.ref_sdpa:
        model_cfg.use_flash_attention = True      # the stronger library baseline (SDPA), never set by its CLI
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
    raw = trainer.model.module if hasattr(trainer.model, "module") else trainer.model
    n_params = sum(p.numel() for p in raw.parameters())
    fsdp = None
    if not args.skip_fsdp:
        fsdp = reference_fsdp(args, sizes, rank, world, local, trainer)
    return {
        "metric": "tokens/sec GPT-2 124M DDP (bf16, batch 8 x seq 1024, grad-accum 4)",
        "value": value, "unit": "tokens/s", "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
        "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "bf16", "data": "synthetic", "impl": "reference",
        "config": ddp_config(args.model_size, n_params, model_cfg.dropout, world),
        "notes": {"model": "reference GPTConfig preset, unmodified",
                  "attention": "F.scaled_dot_product_attention (--ref_sdpa)" if args.ref_sdpa
                  else "manual O(S^2) attention, as its CLI runs it",
                  "path": "torch DDP + autocast + fused AdamW (stock)"},
        "clocks": clocks.summary(),
        "e2e": {"value": value, "unit": "tokens/s", "h2d_bytes_per_step": MICRO_BATCH * GA * SEQ * 8,
                "d2h_bytes_per_step": 4 * GA,
                "note": "the reference's train_step already copies from the host batch and reads the loss back"},
        "gpu_launches": 0,
        "final_loss": losses[-1] if losses else None,
        "fsdp": fsdp,
    }


def reference_fsdp(args, sizes, rank, world, local, ddp_trainer):
    """The reference's own FSDPTrainer (torch FSDP1 + NCCL), unmodified, on the 355M preset."""
    import socket

    import torch

    try:
        from training.fsdp_trainer import FSDPConfig, FSDPTrainer
        from training.fsdp_trainer import TrainingConfig as FTrainingConfig
    except Exception as e:  # pragma: no cover
        return {"unavailable": f"import failed: {type(e).__name__}: {e}"[:200]}
    release(ddp_trainer)
    if "RANK" not in os.environ:
        # single process: the reference's FSDP trainer always calls init_process_group("nccl"); give it the
        # rendezvous variables torchrun would have set
        with socket.socket() as sk:
            sk.bind(("127.0.0.1", 0))
            port = sk.getsockname()[1]
        os.environ.update(RANK="0", WORLD_SIZE="1", LOCAL_RANK="0", MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    try:
        mc = sizes[args.fsdp_model_size]()
        if args.ref_sdpa:
            mc.use_flash_attention = True
        tc = FTrainingConfig(batch_size=FSDP_MICRO, max_steps=1000)
        fc = FSDPConfig(sharding_strategy="FULL_SHARD", activation_checkpointing=True)
        torch.manual_seed(1234)
        tr = FSDPTrainer(mc, tc, fc)
        rec = measure_fsdp(tr, mc.vocab_size, max(3, args.steps // 2), 3, rank, local, "reference")
        n_params = sum(p.numel() for p in tr.model.parameters()) if world == 1 else None
        sys.path.insert(0, ROOT)
        from distributed_llm_trainer_b200.models import GPTConfig as OurCfg

        rec["config"] = fsdp_config_dict(args.fsdp_model_size, OurCfg.preset(args.fsdp_model_size).exact_num_parameters(),
                                         mc.dropout, world)
        rec["notes"] = {"path": "torch FSDP1 + NCCL + autocast (stock)", "local_numel": n_params}
        return rec
    except Exception as e:
        return {"unavailable": f"{type(e).__name__}: {e}"[:300]}


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
