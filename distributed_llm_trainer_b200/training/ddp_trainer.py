"""Data-parallel trainer.

Parity target: reference ``src/training/ddp_trainer.py`` -- ``DistributedTrainer`` (:66-456) with
``train_step(batch) -> {"loss", "lr", "tokens"}``, ``get_lr``, ``save_checkpoint``,
``load_checkpoint``; ``main()`` (:490-625) with the same flags and the same console line.  Semantics
kept on purpose: two AdamW param groups (names containing ``bias``/``norm`` are not
decayed, the embedding is), un-clamped cosine schedule, the LR computed for ``global_step`` is written
*after* the optimizer step (so step k runs at ``get_lr(k-1)``), ``--batch_size`` is the micro-batch and
the loader yields ``batch_size * gradient_accumulation_steps`` sequences per rank, tokens counted as
``numel * world_size``.

Execution paths
* **b200** (CUDA + bf16, the benchmark path): ``FusedGPTEngine`` forward/backward into flat fp32
  gradients, ``NVLinkGradSync`` (sharded reduce + fused AdamW over peer memory) when distributed,
  one device-side clip coefficient, one host sync per step (the reference syncs every micro-step).
* **torch** (CPU, or fp32/fp16 on CUDA): plain PyTorch model under autocast/GradScaler exactly like
  the reference, wrapped in ``BucketedDDP`` (our own bucketed all-reduce over ``torch.distributed``)
  when distributed -- this is what CPU/gloo multi-process tests exercise.

Additions over the reference: ``--config`` YAML, ``--resume_from`` (its ``resume_from``
field and ``load_checkpoint`` are dead code), ``--seed``, periodic eval, JSONL metrics, atomic
checkpoints with plain-dict configs (D9).
"""
from __future__ import annotations

import argparse
import os
import time
from contextlib import nullcontext
from typing import Any, Dict, Optional

import torch
import torch.distributed as dist

from ..data import create_dummy_dataloader, create_openwebtext_dataloader, create_tinystories_dataloader
from ..models import GPT, GPTConfig, count_parameters
from ..utils.checkpoint import atomic_save, config_to_plain, find_latest_checkpoint, load_checkpoint_file
from ..utils.dist import init_distributed
from ..utils.logging import MetricsLogger, format_ddp_line
from ..utils.tracing import StepTracer
from .config import TrainingConfig, load_yaml_config
from .lr import cosine_lr


class DistributedTrainer:
    """DDP trainer (reference ddp_trainer.py:66-456)."""

    def __init__(self, model_config: GPTConfig, training_config: TrainingConfig, model: Optional[GPT] = None):
        self.model_config = model_config
        self.training_config = training_config
        self._setup_distributed()
        self._setup_device()
        self._setup_model(model)
        self._setup_optimizer()
        self.global_step = 0
        self.tokens_seen = 0
        self.last_grad_norm: Optional[float] = None
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
            print(f"Distributed training: {self.distributed}")
            print(f"World size: {self.world_size}")

    def _setup_device(self) -> None:
        mp = self.training_config.mixed_precision
        self.scaler = None
        if torch.cuda.is_available():
            self.device = torch.device(f"cuda:{self.local_rank}")
            torch.cuda.set_device(self.device)
            if mp == "bf16" and torch.cuda.is_bf16_supported():
                self.dtype = torch.bfloat16
                self.autocast_ctx = torch.autocast(device_type="cuda", dtype=torch.bfloat16)
            elif mp == "fp16":
                self.dtype = torch.float16
                self.autocast_ctx = torch.autocast(device_type="cuda", dtype=torch.float16)
                self.scaler = torch.amp.GradScaler()
            else:
                self.dtype = torch.float32
                self.autocast_ctx = nullcontext()
        elif getattr(torch.backends, "mps", None) is not None and torch.backends.mps.is_available():
            # reference ddp_trainer.py:120-146: Apple-silicon path (fp16 autocast or fp32), kept for parity
            self.device = torch.device("mps")
            if mp == "fp16":
                self.dtype = torch.float16
                self.autocast_ctx = torch.autocast(device_type="mps", dtype=torch.float16)
            else:
                self.dtype = torch.float32
                self.autocast_ctx = nullcontext()
        else:
            self.device = torch.device("cpu")
            self.dtype = torch.float32
            self.autocast_ctx = nullcontext()
        # the hand-written engine is the bf16 CUDA path; everything else is the PyTorch path
        self.path = "b200" if (self.device.type == "cuda" and self.dtype == torch.bfloat16) else "torch"
        if self.is_main_process:
            print(f"Device: {self.device}")
            print(f"Mixed precision: {mp} (dtype: {self.dtype})")
            if self.device.type == "cuda" and self.path == "torch":
                print(f"Note: the hand-written sm_100a engine is bf16-only; {mp} runs the PyTorch path "
                      f"(autocast + cuBLAS + NCCL, as the reference)")

    def _setup_model(self, model: Optional[GPT]) -> None:
        cfg = self.training_config
        if cfg.seed is not None:
            torch.manual_seed(cfg.seed)  # rank-independent init: every rank builds identical weights
        self.model = (model if model is not None else GPT(self.model_config)).to(self.device)
        self.raw_model = self.model
        if self.is_main_process:
            print(f"Model parameters: {count_parameters(self.model):,}")
        self.engine = None
        self.grad_sync = None
        if self.path == "b200":
            from ..parallel.ddp import NVLinkGradSync, build_engine_for_ddp

            self.engine, self.grad_sync = build_engine_for_ddp(
                self.model, distributed=self.distributed, rank=self.rank, world_size=self.world_size,
                seed=(cfg.seed or 0) * 7919 + self.rank)
        else:
            self.model.backend = "torch"
            if self.distributed:
                from ..parallel.ddp import BucketedDDP

                self.model = BucketedDDP(self.model)

    def _setup_optimizer(self) -> None:
        cfg = self.training_config
        if self.path == "b200":
            from ..ops.optim import FlatAdamW

            ranges = self.grad_sync.owned_ranges() if self.grad_sync is not None else None
            self.optimizer = FlatAdamW(self.engine.store, lr=cfg.learning_rate, betas=(cfg.beta1, cfg.beta2),
                                       eps=cfg.adam_eps, weight_decay=cfg.weight_decay, ranges=ranges, layout="ddp")
            if self.grad_sync is not None:
                self.grad_sync.attach_optimizer(self.optimizer)
            return
        decay, no_decay = [], []
        for name, p in self.raw_model.named_parameters():
            if not p.requires_grad:
                continue
            (no_decay if ("bias" in name or "norm" in name) else decay).append(p)
        groups = [{"params": decay, "weight_decay": cfg.weight_decay}, {"params": no_decay, "weight_decay": 0.0}]
        self.optimizer = torch.optim.AdamW(groups, lr=cfg.learning_rate, betas=(cfg.beta1, cfg.beta2),
                                           eps=cfg.adam_eps, fused=torch.cuda.is_available())

    # ------------------------------------------------------------------------------------------ schedule
    def get_lr(self, step: int) -> float:
        cfg = self.training_config
        return cosine_lr(step, cfg.learning_rate, cfg.warmup_steps, cfg.max_steps, clamp=False)

    # ------------------------------------------------------------------------------------------ step
    def train_step(self, batch: Dict[str, torch.Tensor], sync: bool = True) -> Dict[str, Any]:
        """One optimizer step.  ``sync=False`` (b200 path) returns the loss as a device tensor instead of a
        float, so a caller that logs every N steps pays no host sync in between."""
        if self.path == "b200":
            return self._train_step_b200(batch, sync)
        return self._train_step_torch(batch)

    def _finish_step(self, input_ids: torch.Tensor) -> float:
        lr = self.get_lr(self.global_step)
        for g in self.optimizer.param_groups:
            g["lr"] = lr
        self.global_step += 1
        self.tokens_seen += input_ids.numel() * self.world_size
        return lr

    def _train_step_b200(self, batch, sync: bool = True):
        cfg = self.training_config
        ga = cfg.gradient_accumulation_steps
        self.model.train()
        input_ids = batch["input_ids"].to(self.device, non_blocking=True)
        micro = input_ids.shape[0] // ga
        eng = self.engine
        if not hasattr(self, "_loss_acc"):
            self._loss_acc = torch.zeros(1, dtype=torch.float32, device=self.device)
            self._sumsq = torch.zeros(1, dtype=torch.float32, device=self.device)
            self._gnorm = torch.zeros(1, dtype=torch.float32, device=self.device)
        self._loss_acc.zero_()
        for m in range(ga):
            ids = input_ids[m * micro:(m + 1) * micro]
            if self.grad_sync is not None:
                self.grad_sync.begin_micro_step(m, ga)
            loss = eng.forward_backward(ids, ids, 1.0 / ga)
            self._loss_acc.add_(loss)
        if self.grad_sync is not None:
            # sharded reduce (overlapped with the last backward) -> global norm -> fused AdamW + broadcast
            self.grad_sync.finish_and_step(cfg.grad_clip, self._gnorm)
        else:
            from ..ops import kernels as K

            self._sumsq.zero_()
            K.sumsq(eng.store.grad, 0, eng.store.total, self._sumsq)
            K.clip_coef(self._sumsq, float(cfg.grad_clip), 1.0, self.optimizer.grad_scale, self._gnorm)
            self.optimizer.step()
            eng.store.zero_grad()
        lr = self._finish_step(input_ids)
        total_loss = float(self._loss_acc.item()) if sync else self._loss_acc.clone()  # the one host sync of the step
        return {"loss": total_loss, "lr": lr, "tokens": self.tokens_seen}

    def _train_step_torch(self, batch):
        cfg = self.training_config
        ga = cfg.gradient_accumulation_steps
        self.model.train()
        self.optimizer.zero_grad(set_to_none=True)
        input_ids = batch["input_ids"].to(self.device)
        labels = input_ids.clone()
        micro = input_ids.shape[0] // ga
        total = 0.0
        for m in range(ga):
            ids, lab = input_ids[m * micro:(m + 1) * micro], labels[m * micro:(m + 1) * micro]
            sync = self.model.no_sync() if (self.distributed and m < ga - 1) else nullcontext()
            with sync:
                with self.autocast_ctx:
                    _, loss = self.model(ids, labels=lab)
                    loss = loss / ga
                if self.scaler is not None:
                    self.scaler.scale(loss).backward()
                else:
                    loss.backward()
            total += loss.item()
        if self.distributed:
            self.model.finish_grad_sync()
        if cfg.grad_clip > 0:
            if self.scaler is not None:
                self.scaler.unscale_(self.optimizer)
            gn = torch.nn.utils.clip_grad_norm_(self.raw_model.parameters(), cfg.grad_clip)
            self.last_grad_norm = float(gn)
        if self.scaler is not None:
            self.scaler.step(self.optimizer)
            self.scaler.update()
        else:
            self.optimizer.step()
        self.optimizer.zero_grad(set_to_none=True)
        lr = self._finish_step(input_ids)
        return {"loss": total, "lr": lr, "tokens": self.tokens_seen}

    @torch.no_grad()
    def evaluate(self, batches) -> float:
        """Mean loss over an iterable of ``{"input_ids"}`` batches, dropout off (the reference's
        ``eval_interval`` is a dead field; this makes it live)."""
        self.model.eval()
        tot, n = 0.0, 0
        for b in batches:
            ids = b["input_ids"].to(self.device)
            if self.path == "b200":
                tot += float(self.engine.forward_loss(ids, ids).item())
            else:
                with self.autocast_ctx:
                    _, loss = self.raw_model(ids, labels=ids)
                tot += float(loss.item())
            n += 1
        self.model.train()
        return tot / max(n, 1)

    # ------------------------------------------------------------------------------------------ checkpoints
    def _full_state(self) -> Dict[str, Any]:
        if self.grad_sync is not None:
            self.grad_sync.gather_master_and_state()     # sharded optimizer -> every rank holds the full state
        model_sd = self.raw_model.state_dict()
        opt_sd = self.grad_sync.full_optimizer_state_dict() if self.grad_sync is not None else self.optimizer.state_dict()
        return {"model": model_sd, "optimizer": opt_sd}

    def save_checkpoint(self, path: str) -> None:
        """Rank 0 writes ``{model, optimizer, global_step, tokens_seen, model_config, training_config}``
        (reference ddp_trainer.py:401-417); collective when the optimizer is sharded."""
        state = self._full_state() if (self.grad_sync is not None or self.is_main_process) else None
        if not self.is_main_process:
            return
        extra = {}
        if self.engine is not None:
            extra["rng_state"] = self.engine.rng_state.cpu()
        ckpt = {
            "model": state["model"], "optimizer": state["optimizer"], "global_step": self.global_step,
            "tokens_seen": self.tokens_seen, "model_config": config_to_plain(self.model_config),
            "training_config": config_to_plain(self.training_config), "extra": extra,
        }
        atomic_save(ckpt, path)

    def load_checkpoint(self, path: str) -> None:
        """Reference ddp_trainer.py:447-456 (which nothing calls); every rank reads the file."""
        ckpt = load_checkpoint_file(path, map_location="cpu")
        self.raw_model.load_state_dict(ckpt["model"])
        if self.engine is not None:
            self.engine.store.refresh_shadow()
            rs = (ckpt.get("extra") or {}).get("rng_state")
            if rs is not None:
                self.engine.rng_state.copy_(rs)
            if self.grad_sync is not None:
                self.grad_sync.load_full_optimizer_state_dict(ckpt["optimizer"])
            else:
                self.optimizer.load_state_dict(ckpt["optimizer"])
        else:
            self.optimizer.load_state_dict(ckpt["optimizer"])
        self.global_step = ckpt["global_step"]
        self.tokens_seen = ckpt["tokens_seen"]
        if self.is_main_process:
            print(f"Loaded Checkpoint from {path} (step {self.global_step})")


# ================================================================================================ CLI
def build_parser() -> argparse.ArgumentParser:
    p = argparse.ArgumentParser(description="DDP trainer (reference flags + --config/--resume_from/--seed)")
    p.add_argument("--config", type=str, default=None, help="YAML file (configs/small_model.yaml schema)")
    p.add_argument("--model_size", type=str, default=None, choices=["small", "medium", "large", "xl"])
    p.add_argument("--batch_size", type=int, default=None)
    p.add_argument("--max_steps", type=int, default=None)
    p.add_argument("--mixed_precision", type=str, default=None, choices=["fp32", "fp16", "bf16"])
    p.add_argument("--gradient_checkpointing", action="store_true")
    p.add_argument("--dataset", type=str, default=None, choices=["dummy", "tinystories", "openwebtext"])
    p.add_argument("--data_path", type=str, default=None)
    p.add_argument("--max_tokens", type=int, default=None)
    p.add_argument("--streaming", action="store_true")
    p.add_argument("--cache_max_tokens", type=int, default=None)
    # additions
    p.add_argument("--resume_from", type=str, default=None, help="checkpoint path, or 'latest'")
    p.add_argument("--seed", type=int, default=None)
    p.add_argument("--gradient_accumulation_steps", type=int, default=None)
    p.add_argument("--warmup_steps", type=int, default=None)
    p.add_argument("--learning_rate", type=float, default=None)
    p.add_argument("--log_interval", type=int, default=None)
    p.add_argument("--save_interval", type=int, default=None)
    p.add_argument("--eval_interval", type=int, default=None)
    p.add_argument("--checkpoint_dir", type=str, default=None)
    p.add_argument("--metrics_path", type=str, default=None)
    p.add_argument("--backend", type=str, default=None, choices=["nccl", "gloo"])
    p.add_argument("--profile_steps", type=int, default=None, help="torch.profiler over this many steps")
    p.add_argument("--profile_start", type=int, default=None, help="first profiled step (default 5)")
    p.add_argument("--profile_dir", type=str, default=None, help="where trace_rank*.json / kernels_rank*.txt go")
    p.add_argument("--nvtx", action="store_true", help="NVTX range per optimizer step")
    # tiny-model overrides so the CPU plumbing run finishes in seconds
    for name in ("vocab_size", "hidden_size", "num_layers", "num_heads", "max_seq_len"):
        p.add_argument(f"--{name}", type=int, default=None)
    p.add_argument("--dropout", type=float, default=None)
    return p


def resolve_configs(args, fsdp: bool = False):
    """YAML (if any) -> dataclasses, then explicit CLI flags override."""
    from .config import FSDPConfig

    extras: Dict[str, Any] = {"data": {}}
    if args.config:
        model_cfg, train_cfg, fsdp_cfg, extras = load_yaml_config(args.config, fsdp=fsdp)
        if args.model_size:
            keep = {k: getattr(model_cfg, k) for k in ("dropout", "attention_dropout", "max_seq_len", "vocab_size",
                                                      "use_flash_attention", "gradient_checkpointing")}
            model_cfg = GPTConfig.preset(args.model_size, **keep)
    else:
        model_cfg = GPTConfig.preset(args.model_size or ("medium" if fsdp else "small"))
        # reference CLI defaults: --batch_size 8 / --max_steps 1000 (DDP), 4 / 1000 (FSDP)
        train_cfg = TrainingConfig.fsdp_defaults(max_steps=1000) if fsdp else TrainingConfig(max_steps=1000)
        fsdp_cfg = FSDPConfig()
    overrides = {}
    for k in ("vocab_size", "hidden_size", "num_layers", "num_heads", "max_seq_len"):
        v = getattr(args, k, None)
        if v is not None:
            overrides[k] = v
    if overrides or getattr(args, "dropout", None) is not None:
        d = model_cfg.to_dict()
        d.update(overrides)
        if "hidden_size" in overrides:
            d["intermediate_size"] = None
        if args.dropout is not None:
            d["dropout"] = d["attention_dropout"] = args.dropout
        model_cfg = GPTConfig.from_dict(d)
    if getattr(args, "gradient_checkpointing", False):
        model_cfg.gradient_checkpointing = True
    for k in ("batch_size", "max_steps", "mixed_precision", "resume_from", "seed", "gradient_accumulation_steps",
              "warmup_steps", "learning_rate", "log_interval", "save_interval", "eval_interval", "checkpoint_dir",
              "metrics_path", "backend", "profile_steps", "profile_start", "profile_dir"):
        v = getattr(args, k, None)
        if v is not None:
            setattr(train_cfg, k, v)
    if getattr(args, "nvtx", False):
        train_cfg.nvtx = True
    train_cfg.max_seq_len = model_cfg.max_seq_len
    return model_cfg, train_cfg, fsdp_cfg, extras


def build_dataloader(args, extras, model_cfg, train_cfg, trainer, always_distributed: bool = False):
    dataset = args.dataset or "dummy"
    bs = train_cfg.batch_size * train_cfg.gradient_accumulation_steps
    common = dict(batch_size=bs, seq_len=model_cfg.max_seq_len, distributed=trainer.distributed or always_distributed,
                  rank=trainer.rank, world_size=trainer.world_size)
    if dataset in ("tinystories", "openwebtext"):
        if args.data_path is None:
            raise ValueError(f"{dataset} dataset requires --data_path")
        fn = create_tinystories_dataloader if dataset == "tinystories" else create_openwebtext_dataloader
        return fn(path=args.data_path, max_tokens=args.max_tokens, streaming=args.streaming,
                  cache_max_tokens=args.cache_max_tokens, num_workers=0 if args.streaming else 2, **common)
    return create_dummy_dataloader(vocab_size=model_cfg.vocab_size, seed=train_cfg.seed,
                                   num_batches=min(1000, max(8, train_cfg.max_steps)), **common)


def run_training_loop(trainer, dataloader, train_cfg, line_fn, banner: str) -> None:
    """Shared by both CLIs; mirrors the reference loop (ddp_trainer.py:574-625)."""
    if trainer.is_main_process:
        print("\n" + "=" * 60)
        print(banner)
        print("=" * 60 + "\n")
    metrics = MetricsLogger(train_cfg.metrics_path, enabled=trainer.is_main_process)
    tracer = StepTracer(train_cfg, rank=getattr(trainer, "rank", 0), device=getattr(trainer, "device", None))
    it = iter(dataloader)
    epoch = 0
    start = time.time()
    start_tokens = trainer.tokens_seen
    for step in range(trainer.global_step, train_cfg.max_steps):
        try:
            batch = next(it)
        except StopIteration:
            epoch += 1
            smp = getattr(dataloader, "sampler", None)
            if hasattr(smp, "set_epoch"):
                smp.set_epoch(epoch)
            it = iter(dataloader)
            batch = next(it)
        ids = batch[0] if isinstance(batch, (list, tuple)) else batch
        tracer.begin(step)
        out = trainer.train_step({"input_ids": ids})
        tracer.end(step)
        if step % train_cfg.log_interval == 0 and trainer.is_main_process:
            elapsed = max(time.time() - start, 1e-9)
            tps = (out["tokens"] - start_tokens) / elapsed
            print(line_fn(step, out, tps))
            rec = {"step": step, "loss": out["loss"], "lr": out["lr"], "tokens": out["tokens"], "tokens_per_sec": tps}
            step_ms = tracer.read_step_ms()     # device time of this step (CUDA events), not in the console line
            if step_ms is not None:
                rec["step_ms"] = step_ms
            metrics.log(rec)
        if train_cfg.eval_interval and step > 0 and step % train_cfg.eval_interval == 0 and hasattr(trainer, "evaluate"):
            ev = []
            ev_it = iter(dataloader)
            for _ in range(train_cfg.eval_batches):
                try:
                    b = next(ev_it)
                except StopIteration:
                    break
                b = b[0] if isinstance(b, (list, tuple)) else b
                ev.append({"input_ids": b[: train_cfg.batch_size]})
            val = trainer.evaluate(ev)
            if trainer.is_main_process:
                print(f"Eval step {step:6d} | Loss: {val:.4f}")
                metrics.log({"step": step, "eval_loss": val})
        if step > 0 and step % train_cfg.save_interval == 0:
            trainer.save_checkpoint(f"{train_cfg.checkpoint_dir}/step_{step}.pt")
    tracer.close()
    trainer.save_checkpoint(f"{train_cfg.checkpoint_dir}/final.pt")
    metrics.close()
    return start


def main(argv=None) -> None:
    args = build_parser().parse_args(argv)
    model_cfg, train_cfg, _, extras = resolve_configs(args, fsdp=False)
    if args.dataset is None and extras.get("data", {}).get("dataset") in ("tinystories", "openwebtext") and args.data_path:
        args.dataset = extras["data"]["dataset"]
    trainer = DistributedTrainer(model_cfg, train_cfg)
    loader = build_dataloader(args, extras, model_cfg, train_cfg, trainer)
    start = run_training_loop(
        trainer, loader, train_cfg,
        lambda step, out, tps: format_ddp_line(step, out["loss"], out["lr"], tps), "Starting training...")
    if trainer.distributed:
        dist.barrier()
        dist.destroy_process_group()
    if trainer.is_main_process:
        print(f"\nTraining complete! Total time: {time.time() - start:.2f}s")
        print(f"Total tokens processed: {trainer.tokens_seen:,}")


if __name__ == "__main__":
    main()
