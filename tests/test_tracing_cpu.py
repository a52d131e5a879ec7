"""Tracing hooks of the training loop (utils/tracing.py): profiler window, metrics with device step time."""
import json
import os

from distributed_llm_trainer_b200.training import ddp_trainer


def test_profile_window_writes_trace_and_kernel_table(tmp_path, monkeypatch):
    monkeypatch.chdir(tmp_path)
    for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK"):
        monkeypatch.delenv(k, raising=False)
    ddp_trainer.main(["--hidden_size", "64", "--num_layers", "1", "--num_heads", "1", "--vocab_size", "256",
                      "--max_seq_len", "32", "--batch_size", "2", "--gradient_accumulation_steps", "2",
                      "--max_steps", "6", "--mixed_precision", "fp32", "--profile_steps", "2", "--profile_start", "2",
                      "--profile_dir", "prof", "--metrics_path", "m.jsonl", "--checkpoint_dir", "ck"])
    assert os.path.getsize("prof/trace_rank0.json") > 1000
    table = open("prof/kernels_rank0.txt").read()
    assert "Name" in table and "aten::" in table
    recs = [json.loads(line) for line in open("m.jsonl")]
    assert [r["step"] for r in recs if "loss" in r] == list(range(6))
    assert os.path.exists("ck/final.pt")


def test_tracer_is_inert_by_default():
    import torch

    from distributed_llm_trainer_b200.training.config import TrainingConfig
    from distributed_llm_trainer_b200.utils.tracing import StepTracer

    tr = StepTracer(TrainingConfig(), rank=0, device=torch.device("cpu"))
    for s in range(3):
        tr.begin(s)
        tr.end(s)
    assert tr.read_step_ms() is None and tr._prof is None
    tr.close()
