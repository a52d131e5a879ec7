"""Trainer semantics on the PyTorch path (CPU): parity with the reference trainers' ``train_step``,
checkpoint layout / cross-loading, resume, CLI and YAML handling, inference script."""
import os
import subprocess
import sys

import pytest
import torch

from distributed_llm_trainer_b200.models import GPT, GPTConfig
from distributed_llm_trainer_b200.training.config import FSDPConfig, TrainingConfig
from distributed_llm_trainer_b200.training.ddp_trainer import DistributedTrainer, build_parser, resolve_configs
from distributed_llm_trainer_b200.training.fsdp_trainer import FSDPTrainer
from distributed_llm_trainer_b200.utils.checkpoint import find_latest_checkpoint, load_checkpoint_file

from ._ref import import_ref

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
TINY = dict(vocab_size=97, hidden_size=32, num_layers=2, num_heads=1, max_seq_len=16, dropout=0.0, attention_dropout=0.0)


def _batches(n, bs=8):
    g = torch.Generator().manual_seed(5)
    return [torch.randint(0, 97, (bs, 16), generator=g) for _ in range(n)]


def test_ddp_train_step_matches_reference_trainer(tmp_path):
    rd = import_ref("training.ddp_trainer")
    rcfg = import_ref("models.config")
    torch.manual_seed(0)
    ref = rd.DistributedTrainer(rcfg.GPTConfig(**TINY),
                                rd.TrainingConfig(batch_size=2, max_steps=20, warmup_steps=3, mixed_precision="fp32"))
    ours = DistributedTrainer(GPTConfig(**TINY), TrainingConfig(batch_size=2, max_steps=20, warmup_steps=3,
                                                                mixed_precision="fp32"))
    ours.raw_model.load_state_dict(ref.model.state_dict())
    for b in _batches(5):
        r = ref.train_step({"input_ids": b})
        o = ours.train_step({"input_ids": b})
        assert abs(r["loss"] - o["loss"]) < 1e-5 and r["lr"] == o["lr"] and r["tokens"] == o["tokens"]
    for (n, p), (_, q) in zip(ref.model.named_parameters(), ours.raw_model.named_parameters()):
        assert torch.allclose(p, q, atol=1e-6), n
    # LR quirk kept: the value returned at step k is get_lr(k), applied from step k+1 on
    assert ours.optimizer.param_groups[0]["lr"] == ours.get_lr(4)
    # checkpoints: same top-level keys, ours loads into the reference model strictly and vice versa
    ref.save_checkpoint(str(tmp_path / "ref" / "c.pt"))
    ours.save_checkpoint(str(tmp_path / "ours" / "c.pt"))
    a = load_checkpoint_file(str(tmp_path / "ref" / "c.pt"))
    b = load_checkpoint_file(str(tmp_path / "ours" / "c.pt"))
    assert set(a) <= set(b) and list(a["model"]) == list(b["model"])
    assert isinstance(b["model_config"], dict) and b["global_step"] == 5
    assert a["optimizer"]["param_groups"][0]["params"] == b["optimizer"]["param_groups"][0]["params"]
    rgpt = import_ref("models.gpt")
    rgpt.GPT(rcfg.GPTConfig(**TINY)).load_state_dict(b["model"])
    fresh = DistributedTrainer(GPTConfig(**TINY), TrainingConfig(batch_size=2, mixed_precision="fp32"))
    fresh.load_checkpoint(str(tmp_path / "ref" / "c.pt"))         # a *reference-written* checkpoint
    assert fresh.global_step == 5
    nb = _batches(6)[-1]
    assert abs(fresh.train_step({"input_ids": nb})["loss"] - ours.train_step({"input_ids": nb})["loss"]) < 1e-5


def test_fsdp_trainer_single_process_semantics(tmp_path):
    """FSDP trainer differences kept from the reference: clamped LR, one param group decaying norms too,
    FQN-keyed optimizer state, fsdp_config in the checkpoint."""
    torch.manual_seed(0)
    model = GPT(GPTConfig(**TINY))
    ref_model = GPT(GPTConfig(**TINY))
    ref_model.load_state_dict(model.state_dict())
    tr = FSDPTrainer(GPTConfig(**TINY), TrainingConfig.fsdp_defaults(batch_size=2, gradient_accumulation_steps=2,
                                                                    max_steps=10, warmup_steps=2),
                     FSDPConfig(mixed_precision="fp32", activation_checkpointing=True), model=model)
    topt = torch.optim.AdamW(ref_model.parameters(), lr=3e-4, betas=(0.9, 0.95), weight_decay=0.1)
    for step, b in enumerate(_batches(3, bs=4)):
        out = tr.train_step({"input_ids": b})
        tot = 0.0
        for m in range(2):
            _, loss = ref_model(b[m * 2:(m + 1) * 2], labels=b[m * 2:(m + 1) * 2])
            (loss / 2).backward()
            tot += loss.item() / 2
        torch.nn.utils.clip_grad_norm_(ref_model.parameters(), 1.0)
        topt.step()
        topt.zero_grad()
        for g in topt.param_groups:
            g["lr"] = tr.get_lr(step)
        assert abs(out["loss"] - tot) < 1e-5
    for (n, p), (_, q) in zip(tr.model.named_parameters(), ref_model.named_parameters()):
        assert torch.allclose(p, q, atol=1e-6), n
    assert tr.get_lr(1000) == tr.get_lr(10)
    path = str(tmp_path / "f" / "final.pt")
    tr.save_checkpoint(path)
    ck = load_checkpoint_file(path)
    assert "fsdp_config" in ck and set(ck["optimizer"]["state"]) == {n for n, _ in ref_model.named_parameters()}
    assert ck["optimizer"]["param_groups"][0]["params"][0] == "embed_tokens.weight"
    tr2 = FSDPTrainer(GPTConfig(**TINY), TrainingConfig.fsdp_defaults(batch_size=2, gradient_accumulation_steps=2,
                                                                     max_steps=10, warmup_steps=2, resume_from=path),
                      FSDPConfig(mixed_precision="fp32"))
    assert tr2.global_step == 3 and tr2.optimizer.step_count == 3
    b = _batches(4, bs=4)[-1]
    assert abs(tr2.train_step({"input_ids": b})["loss"] - tr.train_step({"input_ids": b})["loss"]) < 1e-5


def test_cli_config_precedence():
    args = build_parser().parse_args(["--config", os.path.join(ROOT, "configs", "small_model.yaml"), "--batch_size", "2",
                                      "--max_steps", "7", "--hidden_size", "64", "--num_heads", "1"])
    m, t, f, _ = resolve_configs(args)
    assert (t.batch_size, t.max_steps, t.learning_rate, t.warmup_steps) == (2, 7, 6e-4, 1000)
    assert m.hidden_size == 64 and m.intermediate_size == 256 and m.num_layers == 12
    args = build_parser().parse_args([])
    m, t, f, _ = resolve_configs(args)
    assert (m.hidden_size, t.batch_size, t.max_steps, t.mixed_precision) == (768, 8, 1000, "bf16")   # reference CLI defaults
    m, t, f, _ = resolve_configs(args, fsdp=True)
    assert (m.hidden_size, t.batch_size, t.gradient_accumulation_steps, t.checkpoint_dir) == (1024, 4, 8, "checkpoints_fsdp")


def _run(cmd, **kw):
    env = dict(os.environ, PYTHONPATH=ROOT)
    return subprocess.run([sys.executable] + cmd, cwd=ROOT, env=env, capture_output=True, text=True, timeout=600, **kw)


def test_cli_end_to_end_train_resume_infer(tmp_path):
    tiny = ["--hidden_size", "32", "--num_layers", "2", "--num_heads", "1", "--vocab_size", "300", "--max_seq_len", "32",
            "--batch_size", "2", "--gradient_accumulation_steps", "2", "--warmup_steps", "2", "--seed", "1"]
    ck = str(tmp_path / "ck")
    r = _run(["src/training/ddp_trainer.py", "--max_steps", "4", "--save_interval", "2", "--checkpoint_dir", ck,
              "--metrics_path", os.path.join(ck, "m.jsonl")] + tiny)
    assert r.returncode == 0, r.stderr[-2000:]
    assert "Step      0 | Loss: " in r.stdout and "| LR: 0.00e+00 | Tokens/sec: " in r.stdout
    assert os.path.exists(os.path.join(ck, "final.pt")) and os.path.exists(os.path.join(ck, "step_2.pt"))
    assert find_latest_checkpoint(ck).endswith("final.pt")
    r2 = _run(["src/training/ddp_trainer.py", "--max_steps", "6", "--checkpoint_dir", ck, "--resume_from", "latest"] + tiny)
    assert r2.returncode == 0 and "Loaded Checkpoint" in r2.stdout and "Step      4 |" in r2.stdout
    assert "Step      0 |" not in r2.stdout
    inf = _run(["src/eval/infer.py", "--checkpoint", os.path.join(ck, "final.pt"), "--prompt_ids", "1,2,3",
                "--max_new_tokens", "5", "--greedy", "--device", "cpu"])
    assert inf.returncode == 0, inf.stderr[-2000:]
    assert "token_ids: [1, 2, 3," in inf.stdout
    f = _run(["src/training/fsdp_trainer.py", "--max_steps", "3", "--checkpoint_dir", str(tmp_path / "fk"),
              "--sharding", "SHARD_GRAD_OP", "--log_interval", "1", "--mixed_precision", "fp32"] + tiny)
    assert f.returncode == 0, f.stderr[-2000:]
    assert "Tokens/s: " in f.stdout and "FSDP sharding strategy: SHARD_GRAD_OP" in f.stdout
    inf2 = _run(["src/eval/infer.py", "--checkpoint", str(tmp_path / "fk" / "final.pt"), "--prompt_ids", "4,5",
                 "--max_new_tokens", "3", "--device", "cpu", "--seed", "0"])
    assert inf2.returncode == 0, inf2.stderr[-2000:]
