#!/usr/bin/env python3
"""Where the DDP step's time goes at N GPUs: the GA micro-steps (CUDA-graph replays whose wgrad GEMMs reduce-add
into the owners over NVLink) vs the end-of-step sequence (barrier, norm, scalar all-reduce, AdamW + parameter
broadcast, barrier).  CUDA events, max over ranks.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        benchmarks/bench_ddp_breakdown.py [--iters 10]
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from bench import GA, MICRO_BATCH, SEQ, barrier_sync, dist_info, max_over_ranks  # noqa: E402


def nvlink_kib(index: int):
    """(tx, rx) payload KiB summed over this GPU's NVLinks since boot (NVML field values 138 / 139), or None."""
    try:
        import pynvml as nv

        nv.nvmlInit()
        h = nv.nvmlDeviceGetHandleByIndex(index)
        vals = nv.nvmlDeviceGetFieldValues(h, [(138, 0xFFFFFFFF), (139, 0xFFFFFFFF)])
        out = []
        for v in vals:
            if v.nvmlReturn != 0:
                return None
            out.append(int(v.value.ullVal))
        return tuple(out)
    except Exception:  # noqa: BLE001
        return None


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--model_size", default="small")
    args = ap.parse_args()
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    import torch

    from distributed_llm_trainer_b200.models import GPTConfig
    from distributed_llm_trainer_b200.training.config import TrainingConfig
    from distributed_llm_trainer_b200.training.ddp_trainer import DistributedTrainer

    rank, world, local = dist_info()
    mc = GPTConfig.preset(args.model_size)
    tc = TrainingConfig(batch_size=MICRO_BATCH, gradient_accumulation_steps=GA, max_steps=1000, mixed_precision="bf16",
                        seed=1234)
    tr = DistributedTrainer(mc, tc)
    dev = tr.device
    g = torch.Generator().manual_seed(5 + rank)
    ids = torch.randint(0, mc.vocab_size, (MICRO_BATCH * GA, SEQ), generator=g).to(dev)
    for _ in range(4):
        tr.train_step({"input_ids": ids}, sync=False)
    eng, sync = tr.engine, tr.grad_sync
    micro = MICRO_BATCH
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
    t_micro, t_step = [], []
    nv0 = nvlink_kib(local)
    phases = {}
    marks = []
    if sync is not None:
        def trace(name):
            e = torch.cuda.Event(enable_timing=True)
            e.record()
            marks.append((name, e))
        sync.trace = trace
    for it in range(args.iters):
        marks.clear()
        barrier_sync()
        ev[0].record()
        for m in range(GA):
            if sync is not None:
                sync.begin_micro_step(m, GA)
            eng.forward_backward(ids[m * micro:(m + 1) * micro], ids[m * micro:(m + 1) * micro], 1.0 / GA)
        ev[1].record()
        if sync is not None:
            sync.finish_and_step(tc.grad_clip, tr._gnorm)
        else:
            from distributed_llm_trainer_b200.ops import kernels as K

            tr._sumsq.zero_()
            K.sumsq(eng.store.grad, 0, eng.store.total, tr._sumsq)
            K.clip_coef(tr._sumsq, float(tc.grad_clip), 1.0, tr.optimizer.grad_scale, tr._gnorm)
            tr.optimizer.step()
            eng.store.zero_grad()
        ev[2].record()
        torch.cuda.synchronize()
        t_micro.append(ev[0].elapsed_time(ev[1]))
        t_step.append(ev[1].elapsed_time(ev[2]))
        for (_, a), (nm, b) in zip(marks, marks[1:]):
            phases.setdefault(nm, []).append(a.elapsed_time(b))
    nv1 = nvlink_kib(local)
    t_micro.sort()
    t_step.sort()
    med_m = max_over_ranks(t_micro[len(t_micro) // 2], dev)
    med_s = max_over_ranks(t_step[len(t_step) // 2], dev)
    if rank == 0:
        print(json.dumps({"phases_ms_rank0_median": {k: sorted(v)[len(v) // 2] for k, v in phases.items()}}))
        if nv0 and nv1:
            owned = sum(b - a for a, b in sync.ranges) if sync is not None else 0
            total = eng.store.total
            print(json.dumps({"nvlink_MB_per_step_rank0": {"tx": (nv1[0] - nv0[0]) * 1024 / 1e6 / args.iters,
                                                           "rx": (nv1[1] - nv0[1]) * 1024 / 1e6 / args.iters},
                              "algorithmic_MB_per_step": {
                                  "gradient_out_fp32": 4 * (total - owned) / 1e6, "gradient_in_fp32": 4 * owned * (world - 1) / 1e6,
                                  "params_in_bf16": 2 * (total - owned) / 1e6,
                                  "params_out_bf16_multicast": 2 * owned / 1e6 if world > 1 else 0}}))
        print(json.dumps({"n_gpus": world, "micro_steps_ms": med_m, "end_of_step_ms": med_s, "total_ms": med_m + med_s,
                          "owned_params": sum(b - a for a, b in sync.ranges) if sync is not None else eng.store.total}))
    import torch.distributed as dist

    if dist.is_initialized():
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
