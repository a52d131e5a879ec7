"""Fully-sharded trainer.

Parity target: reference ``src/training/fsdp_trainer.py`` -- ``FSDPTrainer`` (:96-505) with
``train_step``, ``get_lr`` (clamped cosine), ``save_checkpoint`` / ``load_checkpoint`` (FULL_STATE_DICT
layout with FQN-keyed optimizer state + ``fsdp_config``), ``get_memory_stats``; ``main()`` (:530-621)
with ``--model_size --batch_size --max_steps --sharding --cpu_offload --no_activation_checkpointing``
and the ``Tokens/s | Mem`` console line.  Semantics kept: one AdamW group with weight
decay on *every* tensor (norms included), reduce-scatter every micro-step (no ``no_sync``), activation
checkpointing on by default, LR written after the step.

Differences, on purpose: the data flags of the DDP trainer are available here too, a
single-process / CPU path exists (the reference hard-requires CUDA + torchrun), model init is seeded
and identical on all ranks, ``--config`` / ``--resume_from`` work.
"""
from __future__ import annotations

import contextlib
import os
import time
from typing import Any, Dict, Optional

import torch
import torch.distributed as dist

from ..models import GPT, GPTConfig, TransformerBlock, count_parameters  # noqa: F401  (TransformerBlock: wrap-policy parity)
from ..utils.checkpoint import atomic_save, config_to_plain, find_latest_checkpoint, load_checkpoint_file
from ..utils.dist import init_distributed
from ..utils.logging import format_fsdp_line
from .config import SHARDING_STRATEGIES, FSDPConfig, TrainingConfig
from .ddp_trainer import build_dataloader, build_parser, resolve_configs, run_training_loop
from .lr import cosine_lr


class FSDPTrainer:
    def __init__(self, model_config: GPTConfig, training_config: TrainingConfig, fsdp_config: FSDPConfig,
                 model: Optional[GPT] = None):
        self.model_config = model_config
        self.training_config = training_config
        self.fsdp_config = fsdp_config
        if fsdp_config.sharding_strategy not in SHARDING_STRATEGIES:
            raise KeyError(fsdp_config.sharding_strategy)
        self._setup_distributed()
        self._setup_device()
        self._setup_model(model)
        self._setup_optimizer()
        self.global_step = 0
        self.tokens_seen = 0
        if training_config.resume_from:
            path = training_config.resume_from
            if path == "latest":
                path = find_latest_checkpoint(training_config.checkpoint_dir)
            if path and os.path.exists(path):
                self.load_checkpoint(path)

    # ------------------------------------------------------------------------------------------ setup
    def _setup_distributed(self) -> None:
        info = init_distributed(self.training_config.backend)
        self.distributed, self.rank, self.world_size, self.local_rank = (
            info.distributed, info.rank, info.world_size, info.local_rank)
        self.is_main_process = self.rank == 0
        if self.is_main_process:
            print(f"FSDP Training on {self.world_size} GPUs")

    def _setup_device(self) -> None:
        mp = self.fsdp_config.mixed_precision
        if torch.cuda.is_available():
            self.device = torch.device(f"cuda:{self.local_rank}")
            torch.cuda.set_device(self.device)
            self.compute_dtype = {"bf16": torch.bfloat16, "fp16": torch.float16}.get(mp, torch.float32)
        else:
            self.device = torch.device("cpu")
            self.compute_dtype = torch.float32
        self.path = "b200" if (self.device.type == "cuda" and self.compute_dtype == torch.bfloat16) else "torch"
        if self.is_main_process and self.device.type == "cuda" and self.path == "torch":
            print(f"Note: the hand-written sm_100a engine is bf16-only; mixed_precision={mp} runs the PyTorch path "
                  f"(autocast + cuBLAS + torch.distributed collectives)")

    def _setup_model(self, model: Optional[GPT]) -> None:
        cfg, fc = self.training_config, self.fsdp_config
        if self.is_main_process:
            print("Building model...")
        if cfg.seed is not None:
            torch.manual_seed(cfg.seed)
        self.model_config.gradient_checkpointing = bool(fc.activation_checkpointing)
        self.model = model if model is not None else GPT(self.model_config)
        self.model.gradient_checkpointing = bool(fc.activation_checkpointing)
        self.model = self.model.to(self.device)
        if self.is_main_process:
            print(f"Model parameters: {count_parameters(self.model):,}")
        strategy = fc.sharding_strategy
        self.sharded = None      # ShardedGPT (b200, sharded strategies)
        self.engine = None
        self.grad_sync = None    # NVLinkGradSync (b200, NO_SHARD with >1 rank)
        self.tsm = None          # TorchShardedModel (torch path)
        seed = (cfg.seed or 0) * 7919 + self.rank
        if self.path == "b200":
            if self.world_size > 1 and strategy in ("FULL_SHARD", "SHARD_GRAD_OP"):
                from ..parallel.fsdp import ShardedGPT

                self.sharded = ShardedGPT(self.model, strategy, self.rank, self.world_size, seed=seed,
                                          cpu_offload=fc.cpu_offload, limit_all_gathers=fc.limit_all_gathers,
                                          backward_prefetch=fc.backward_prefetch == "BACKWARD_PRE")
                self.engine = self.sharded.engine
            else:
                from ..parallel.ddp import build_engine_for_ddp

                self.engine, self.grad_sync = build_engine_for_ddp(
                    self.model, distributed=self.distributed, rank=self.rank, world_size=self.world_size, seed=seed)
        else:
            from ..parallel.fsdp import TorchShardedModel

            self.model.backend = "torch"
            self.tsm = TorchShardedModel(self.model, strategy, self.rank, self.world_size)
        if fc.activation_checkpointing and self.is_main_process:
            print("Activation checkpointing enabled")
        if self.is_main_process:
            print(f"FSDP sharding strategy: {strategy}")

    def _setup_optimizer(self) -> None:
        from ..ops.optim import FlatAdamW

        cfg = self.training_config
        kw = dict(lr=cfg.learning_rate, betas=(cfg.beta1, cfg.beta2), eps=cfg.adam_eps, weight_decay=cfg.weight_decay,
                  decay_norms=True, layout="fsdp")
        if self.sharded is not None:
            sh = self.sharded
            # formatting / param_groups facade (no state of its own) + the sharded state the kernels update
            self.optimizer = FlatAdamW(sh.store, ranges=[(0, 0)], **kw)
            dev = self.device
            self.opt_state = {
                # --cpu_offload keeps the moments (and the fp32 master shard) in pinned host memory
                "exp_avg": (torch.zeros(max(sh.own_total, 4), dtype=torch.float32, pin_memory=True) if sh.cpu_offload
                            else torch.zeros(max(sh.own_total, 4), dtype=torch.float32, device=dev)),
                "exp_avg_sq": (torch.zeros(max(sh.own_total, 4), dtype=torch.float32, pin_memory=True) if sh.cpu_offload
                               else torch.zeros(max(sh.own_total, 4), dtype=torch.float32, device=dev)),
                "step": 0, "grad_scale": self.optimizer.grad_scale, "decay": self.optimizer._decay,
            }
        elif self.engine is not None:
            ranges = self.grad_sync.owned_ranges() if self.grad_sync is not None else None
            self.optimizer = FlatAdamW(self.engine.store, ranges=ranges, **kw)
            if self.grad_sync is not None:
                self.grad_sync.attach_optimizer(self.optimizer)
        else:
            self.optimizer = FlatAdamW(self.tsm.store, ranges=self.tsm.owned_ranges(), **kw)

    # ------------------------------------------------------------------------------------------ schedule
    def get_lr(self, step: int) -> float:
        cfg = self.training_config
        return cosine_lr(step, cfg.learning_rate, cfg.warmup_steps, cfg.max_steps, clamp=True)

    # ------------------------------------------------------------------------------------------ step
    def train_step(self, batch: Dict[str, torch.Tensor]) -> Dict[str, Any]:
        cfg = self.training_config
        ga = cfg.gradient_accumulation_steps
        self.model.train()
        input_ids = batch["input_ids"].to(self.device, non_blocking=True)
        micro = input_ids.shape[0] // ga
        if self.path == "b200":
            if not hasattr(self, "_loss_acc"):
                self._loss_acc = torch.zeros(1, dtype=torch.float32, device=self.device)
                self._sumsq = torch.zeros(1, dtype=torch.float32, device=self.device)
                self._gnorm = torch.zeros(1, dtype=torch.float32, device=self.device)
            self._loss_acc.zero_()
            runner = self.sharded if self.sharded is not None else self.engine
            for m in range(ga):
                ids = input_ids[m * micro:(m + 1) * micro]
                if self.sharded is not None:
                    self._loss_acc.add_(runner.forward_backward(ids, ids, 1.0 / ga, last=(m == ga - 1), first=(m == 0)))
                else:
                    if self.grad_sync is not None:
                        self.grad_sync.begin_micro_step(m, ga)
                    self._loss_acc.add_(runner.forward_backward(ids, ids, 1.0 / ga))
            g0 = self.optimizer.param_groups[0]
            if self.sharded is not None:
                self.sharded.step(self.opt_state, g0["lr"], g0["betas"], g0["eps"], g0["weight_decay"], cfg.grad_clip,
                                  self._gnorm)
                self.optimizer.step_count = self.opt_state["step"]
            elif self.grad_sync is not None:
                self.grad_sync.finish_and_step(cfg.grad_clip, self._gnorm)
            else:
                from ..ops import kernels as K

                st = self.engine.store
                self._sumsq.zero_()
                K.sumsq(st.grad, 0, st.total, self._sumsq)
                K.clip_coef(self._sumsq, float(cfg.grad_clip), 1.0, self.optimizer.grad_scale, self._gnorm)
                self.optimizer.step()
                st.zero_grad()
            total = None
        else:
            total = 0.0
            tsm = self.tsm
            # reference fsdp_trainer.py:378: explicit autocast to the compute dtype (fp16 runs without a
            # GradScaler there too, C26); CPU and fp32 run un-cast
            amp = (torch.autocast("cuda", dtype=self.compute_dtype)
                   if self.device.type == "cuda" and self.compute_dtype != torch.float32 else contextlib.nullcontext())
            for m in range(ga):
                ids = input_ids[m * micro:(m + 1) * micro]
                with amp:
                    _, loss = self.model(ids, labels=ids.clone())
                loss = loss / ga
                loss.backward()
                total += loss.item()
            tsm.reduce_gradients()
            if cfg.grad_clip > 0:
                norm = float(tsm.grad_sumsq().sqrt().item())
                self.optimizer.grad_scale.fill_(min(1.0, cfg.grad_clip / (norm + 1e-6)))
            else:
                self.optimizer.grad_scale.fill_(1.0)
            self.optimizer.step()
            tsm.gather_params()
            tsm.store.zero_grad()
        lr = self.get_lr(self.global_step)
        for g in self.optimizer.param_groups:
            g["lr"] = lr
        self.global_step += 1
        self.tokens_seen += input_ids.numel() * self.world_size
        if total is None:
            total = float(self._loss_acc.item())
        return {"loss": total, "lr": lr, "tokens": self.tokens_seen}

    # ------------------------------------------------------------------------------------------ checkpoints
    def _full_state(self) -> Dict[str, Any]:
        """Collective: assemble the full (unsharded) model and FQN-keyed optimizer state."""
        opt = self.optimizer
        if self.sharded is not None:
            sh = self.sharded
            sh.gather_full_master()
            fm = sh.gather_flat(self.opt_state["exp_avg"].to(self.device)).cpu()
            fv = sh.gather_flat(self.opt_state["exp_avg_sq"].to(self.device)).cpu()
            opt.step_count = self.opt_state["step"]
            return {"model": self.model.state_dict(), "optimizer": opt.state_dict(fm, fv)}
        if self.grad_sync is not None:
            self.grad_sync.gather_master_and_state()
            return {"model": self.model.state_dict(), "optimizer": self.grad_sync.full_optimizer_state_dict()}
        if self.tsm is not None:
            fm, fv = self.tsm.full_optimizer_state(opt)
            return {"model": self.model.state_dict(), "optimizer": opt.state_dict(fm, fv)}
        return {"model": self.model.state_dict(), "optimizer": opt.state_dict()}

    def save_checkpoint(self, path: str) -> None:
        """FULL_STATE_DICT save on rank 0 + barrier (reference fsdp_trainer.py:444-465)."""
        state = self._full_state()
        if self.is_main_process:
            extra = {}
            if self.engine is not None:
                extra["rng_state"] = self.engine.rng_state.cpu()
            atomic_save({
                "model": {k: v.detach().cpu() for k, v in state["model"].items()}, "optimizer": state["optimizer"],
                "global_step": self.global_step, "tokens_seen": self.tokens_seen,
                "model_config": config_to_plain(self.model_config),
                "training_config": config_to_plain(self.training_config),
                "fsdp_config": config_to_plain(self.fsdp_config), "extra": extra,
            }, path)
            print(f"Saved checkpoint to {path} (step {self.global_step})")
        del state
        if self.sharded is not None:
            self.sharded.release_full_master()      # back to 1/N of the fp32 parameters per rank
        if self.distributed:
            dist.barrier()

    def load_checkpoint(self, path: str) -> None:
        """Every rank reads the file and keeps its own slices (the reference pickles the whole checkpoint
        through ``broadcast_object_list``, fsdp_trainer.py:467-494 / SURVEY K13)."""
        ckpt = load_checkpoint_file(path, map_location="cpu")
        if self.sharded is not None:
            self.sharded.materialize_full_master()  # module parameters need storage to load into
        self.model.load_state_dict(ckpt["model"])
        opt = self.optimizer
        if self.sharded is not None:
            sh = self.sharded
            sh._refresh_shard_from_master()
            sh._root_stale = True
            full = type(opt)(sh.store, lr=0.0, layout="fsdp", decay_norms=True)   # temporary full-size holder
            full.load_state_dict(ckpt["optimizer"])
            sh.scatter_flat(full.exp_avg, self.opt_state["exp_avg"])
            sh.scatter_flat(full.exp_avg_sq, self.opt_state["exp_avg_sq"])
            if hasattr(sh, "_host"):
                del sh._host                      # host master shard is rebuilt from the loaded weights
            self.opt_state["step"] = full.step_count
            opt.step_count = full.step_count
            # hyper-parameters travel with the checkpoint too (lr is written AFTER each step, so the first step
            # after a resume must run at the saved value, not at the peak rate the facade was built with)
            for g, src in zip(opt.param_groups, full.param_groups):
                for k in ("lr", "betas", "eps", "weight_decay"):
                    g[k] = src[k]
            del full
            sh.release_full_master()
        else:
            opt.load_state_dict(ckpt["optimizer"])
            if self.engine is not None:
                if self.grad_sync is not None:
                    self.grad_sync.shadow_buf.local[: self.engine.store.total].copy_(self.engine.store.master)
                else:
                    self.engine.store.refresh_shadow()
        if self.engine is not None:
            rs = (ckpt.get("extra") or {}).get("rng_state")
            if rs is not None:
                self.engine.rng_state.copy_(rs)
        self.global_step = ckpt["global_step"]
        self.tokens_seen = ckpt["tokens_seen"]
        if self.is_main_process:
            print(f"Loaded checkpoint from {path} (step {self.global_step})")

    def get_memory_stats(self) -> Dict[str, float]:
        """Reference fsdp_trainer.py:496-505 (GB = 1e9 bytes)."""
        if not torch.cuda.is_available():
            return {}
        return {"allocated_gb": torch.cuda.memory_allocated() / 1e9, "reserved_gb": torch.cuda.memory_reserved() / 1e9,
                "max_allocated_gb": torch.cuda.max_memory_allocated() / 1e9}


# ================================================================================================ CLI
def main(argv=None) -> None:
    p = build_parser()
    p.description = "FSDP trainer (reference flags + the DDP trainer's data flags, --config, --resume_from)"
    p.add_argument("--sharding", default=None, choices=list(SHARDING_STRATEGIES))
    p.add_argument("--cpu_offload", action="store_true")
    p.add_argument("--no_activation_checkpointing", action="store_true")
    args = p.parse_args(argv)
    model_cfg, train_cfg, fsdp_cfg, extras = resolve_configs(args, fsdp=True)
    if args.sharding:
        fsdp_cfg.sharding_strategy = args.sharding
    if args.cpu_offload:
        fsdp_cfg.cpu_offload = True
    if args.no_activation_checkpointing:
        fsdp_cfg.activation_checkpointing = False
    if args.mixed_precision:
        fsdp_cfg.mixed_precision = args.mixed_precision
    trainer = FSDPTrainer(model_cfg, train_cfg, fsdp_cfg)
    loader = build_dataloader(args, extras, model_cfg, train_cfg, trainer)
    if trainer.is_main_process and torch.cuda.is_available():
        print(f"Initial memory: {trainer.get_memory_stats()['allocated_gb']:.2f} GB allocated")

    def line(step, out, tps):
        mem = trainer.get_memory_stats().get("allocated_gb", 0.0)
        return format_fsdp_line(step, out["loss"], out["lr"], tps, mem)

    start = run_training_loop(trainer, loader, train_cfg, line, "Starting FSDP training...")
    if trainer.is_main_process:
        print(f"\nTraining complete! Total time: {time.time() - start:.2f}s")
        print(f"Tokens processed: {trainer.tokens_seen:,}")
        if torch.cuda.is_available():
            print(f"Peak memory: {trainer.get_memory_stats()['max_allocated_gb']:.2f} GB")
    if trainer.distributed:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
