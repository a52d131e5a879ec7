"""Host-side logic of the parallel engines under real multi-process collectives: 2 ranks, gloo, CPU.
Each mode checks N-rank training against the single-process run on the same global batch, the collective
checkpoint save (reference layouts) and resume."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _torchrun(mode, tmp, nproc=2, port=29611):
    env = dict(os.environ, PYTHONPATH=ROOT, OMP_NUM_THREADS="2", CUDA_VISIBLE_DEVICES="")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={nproc}",
           "--master-addr", "127.0.0.1", "--master-port", str(port), os.path.join(ROOT, "tests", "dist_worker.py"),
           mode, str(tmp)]
    return subprocess.run(cmd, cwd=ROOT, env=env, capture_output=True, text=True, timeout=900)


@pytest.mark.parametrize("mode,port", [("ddp", 29611), ("fsdp:FULL_SHARD", 29612), ("fsdp:NO_SHARD", 29613)])
def test_two_rank_gloo(mode, port, tmp_path):
    r = _torchrun(mode, tmp_path, port=port)
    assert r.returncode == 0, (r.stdout[-1500:], r.stderr[-3000:])
    assert r.stdout.count("OK") == 2
