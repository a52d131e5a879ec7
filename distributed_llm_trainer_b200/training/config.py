"""Training / FSDP configuration dataclasses and the YAML -> dataclass loader.

Parity targets
* DDP ``TrainingConfig``  -- reference ``ddp_trainer.py:34-63`` (same fields, same defaults).
* FSDP ``TrainingConfig`` -- reference ``fsdp_trainer.py:78-93`` (different defaults: batch 4,
  lr 3e-4, GA 8, log 10, dir ``checkpoints_fsdp``).
* ``FSDPConfig``          -- reference ``fsdp_trainer.py:53-75``.
* YAML schema             -- reference ``configs/small_model.yaml`` / ``medium_model.yaml``.  The
  reference never parses those files (neither trainer defines ``--config``); here
  ``load_yaml_config`` turns them into the dataclasses and explicit CLI flags override the file.

New fields (all optional, defaults keep reference behaviour): ``seed``, ``metrics_path``,
``resume_from``/``eval_interval`` are live instead of dead, ``backend``, the tracing switches
``profile_steps`` / ``profile_start`` / ``profile_dir`` / ``nvtx``.
"""
from __future__ import annotations

from dataclasses import asdict, dataclass, fields
from typing import Any, Dict, Optional, Tuple

from ..models.config import GPTConfig


@dataclass
class TrainingConfig:
    """Defaults are the reference DDP trainer's (ddp_trainer.py:34-63)."""
    batch_size: int = 8
    max_seq_len: int = 1024
    learning_rate: float = 6e-4
    weight_decay: float = 0.1
    beta1: float = 0.9
    beta2: float = 0.95
    grad_clip: float = 1.0
    max_steps: int = 10000
    warmup_steps: int = 1000
    log_interval: int = 1
    eval_interval: int = 500
    save_interval: int = 1000
    mixed_precision: str = "bf16"  # "fp32" | "fp16" | "bf16"
    gradient_accumulation_steps: int = 4
    checkpoint_dir: str = "checkpoints"
    resume_from: Optional[str] = None
    # -- additions ---------------------------------------------------------
    seed: Optional[int] = None
    metrics_path: Optional[str] = None   # JSONL sink
    eval_batches: int = 4
    backend: str = "nccl"                # "gloo" for CPU multi-process tests
    adam_eps: float = 1e-8
    # tracing (utils/tracing.py): torch.profiler window, NVTX ranges
    profile_steps: int = 0
    profile_start: int = 5
    profile_dir: str = "profiles_run"
    nvtx: bool = False

    def to_dict(self) -> Dict[str, Any]:
        return asdict(self)

    @classmethod
    def from_dict(cls, d: Dict[str, Any]) -> "TrainingConfig":
        known = {f.name for f in fields(cls)}
        return cls(**{k: v for k, v in d.items() if k in known})

    @classmethod
    def fsdp_defaults(cls, **kw: Any) -> "TrainingConfig":
        """Reference FSDP trainer defaults (fsdp_trainer.py:78-93)."""
        base = dict(batch_size=4, learning_rate=3e-4, log_interval=10, gradient_accumulation_steps=8,
                    checkpoint_dir="checkpoints_fsdp")
        base.update(kw)
        return cls(**base)


@dataclass
class FSDPConfig:
    """Reference fsdp_trainer.py:53-75."""
    sharding_strategy: str = "FULL_SHARD"       # FULL_SHARD | SHARD_GRAD_OP | NO_SHARD
    cpu_offload: bool = False
    mixed_precision: str = "bf16"
    backward_prefetch: str = "BACKWARD_PRE"
    activation_checkpointing: bool = True
    limit_all_gathers: bool = True

    def to_dict(self) -> Dict[str, Any]:
        return asdict(self)

    @classmethod
    def from_dict(cls, d: Dict[str, Any]) -> "FSDPConfig":
        known = {f.name for f in fields(cls)}
        return cls(**{k: v for k, v in d.items() if k in known})


SHARDING_STRATEGIES = ("FULL_SHARD", "SHARD_GRAD_OP", "NO_SHARD")


# ----------------------------------------------------------------------------- YAML
def _num(v: Any) -> Any:
    """PyYAML reads ``6e-4`` (no dot) as a string; coerce numeric-looking strings."""
    if isinstance(v, str):
        try:
            return float(v) if any(c in v for c in ".eE") else int(v)
        except ValueError:
            return v
    return v


def load_yaml_config(path: str, fsdp: bool = False) -> Tuple[GPTConfig, TrainingConfig, FSDPConfig, Dict[str, Any]]:
    """Parse a reference-schema YAML file.

    Returns ``(model_config, training_config, fsdp_config, extras)`` where ``extras`` carries the
    ``data`` section (``dataset``, ``num_workers``, ``data_path`` ...) for the CLI.
    """
    import yaml

    with open(path, "r", encoding="utf-8") as fh:
        doc = yaml.safe_load(fh) or {}
    m = {k: _num(v) for k, v in (doc.get("model") or {}).items()}
    m.pop("name", None)
    model_cfg = GPTConfig.from_dict(m)

    t = {k: _num(v) for k, v in (doc.get("training") or {}).items()}
    dist_sec = doc.get("distributed") or {}
    ck = doc.get("checkpoint") or {}
    if "mixed_precision" in dist_sec:
        t["mixed_precision"] = dist_sec["mixed_precision"]
    if "backend" in dist_sec:
        t["backend"] = dist_sec["backend"]
    if ck.get("dir"):
        t["checkpoint_dir"] = ck["dir"]
    if ck.get("resume_from"):
        t["resume_from"] = ck["resume_from"]
    t.setdefault("max_seq_len", model_cfg.max_seq_len)
    base = TrainingConfig.fsdp_defaults() if fsdp else TrainingConfig()
    merged = base.to_dict()
    merged.update({k: v for k, v in t.items() if k in merged})
    train_cfg = TrainingConfig.from_dict(merged)

    f = dict(doc.get("fsdp") or {})
    if "mixed_precision" in dist_sec:
        f.setdefault("mixed_precision", dist_sec["mixed_precision"])
    # model.gradient_checkpointing in the YAML is the FSDP trainer's activation checkpointing knob too
    fsdp_cfg = FSDPConfig.from_dict(f)
    extras = {"data": dict(doc.get("data") or {}), "raw": doc}
    return model_cfg, train_cfg, fsdp_cfg, extras
