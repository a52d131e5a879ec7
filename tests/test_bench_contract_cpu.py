"""bench.py contract pieces that can be checked without a GPU: defaults, the config dict shared by both arms, the
headline shape named by BASELINE.json, and the reference arm's "unavailable" answer."""
import importlib
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _bench():
    if ROOT not in sys.path:
        sys.path.insert(0, ROOT)
    return importlib.import_module("bench")


def test_defaults_and_flags(monkeypatch):
    bench = _bench()
    monkeypatch.setattr(sys, "argv", ["bench.py"])
    a = bench.parse()
    assert a.gpus == 1 and a.impl == "b200" and a.warmup >= 3 and a.steps >= 1
    monkeypatch.setattr(sys, "argv", ["bench.py", "--impl", "reference", "--gpus", "8", "--steps", "7", "--warmup", "4"])
    a = bench.parse()
    assert (a.impl, a.gpus, a.steps, a.warmup) == ("reference", 8, 7, 4)


def test_config_dict_is_the_headline_shape_and_arm_independent():
    bench = _bench()
    cfg = bench.ddp_config("small", 151862784, 0.1, 8)
    # BASELINE.json: GPT-2 "124M" preset, batch 8 x seq 1024, grad-accum 4, one rank per GPU
    assert cfg["micro_batch"] == 8 and cfg["seq_len"] == 1024 and cfg["grad_accum"] == 4
    assert cfg["global_batch"] == 8 * 4 * 8 and cfg["parallelism"] == "dp8" and cfg["model"] == "gpt2-small"
    # both arms call the same function with the same arguments: nothing arm-specific may leak into it
    assert "impl" not in cfg and "path" not in cfg and "cuda_graphs" not in cfg
    f = bench.fsdp_config_dict("medium", 454166528, 0.1, 8)
    assert f["micro_batch"] == 4 and f["grad_accum"] == 8 and f["activation_checkpointing"] is True
    assert f["parallelism"] == "fsdp8:FULL_SHARD" and f["global_batch"] == 4 * 8 * 8
    json.dumps(cfg), json.dumps(f)   # serialisable


def test_reference_arm_reports_unavailable_without_the_install(tmp_path):
    """A checkout without baseline/_ref answers with one JSON line and exit code 0 (nothing of this repo's engine runs)."""
    work = tmp_path / "repo"
    work.mkdir()
    for name in ("bench.py",):
        (work / name).write_text(open(os.path.join(ROOT, name)).read())
    os.symlink(os.path.join(ROOT, "distributed_llm_trainer_b200"), work / "distributed_llm_trainer_b200")
    env = {k: v for k, v in os.environ.items() if k not in ("RANK", "WORLD_SIZE", "LOCAL_RANK")}
    r = subprocess.run([sys.executable, "bench.py", "--impl", "reference"], cwd=work, env=env, capture_output=True,
                       text=True, timeout=300)
    assert r.returncode == 0, r.stderr[-2000:]
    line = json.loads(r.stdout.strip().splitlines()[-1])
    assert line["impl"] == "reference" and "unavailable" in line and "baseline/_ref" in line["unavailable"]


def test_multi_gpu_request_without_torchrun_is_refused(monkeypatch, capsys):
    bench = _bench()
    monkeypatch.setattr(sys, "argv", ["bench.py", "--gpus", "4"])
    for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK"):
        monkeypatch.delenv(k, raising=False)
    try:
        bench.main()
    except SystemExit as e:
        assert e.code == 2
    else:  # pragma: no cover
        raise AssertionError("bench.py --gpus 4 without torchrun should exit")
    assert "needs torchrun" in capsys.readouterr().out
