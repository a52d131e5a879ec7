"""NVLink comm kernels and the sharded DDP / FSDP engines on 2+ real GPUs (one process per GPU)."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _ngpu():
    return torch.cuda.device_count() if torch.cuda.is_available() else 0


def _torchrun(mode, tmp, nproc, port):
    env = dict(os.environ, PYTHONPATH=ROOT)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={nproc}",
           "--master-addr", "127.0.0.1", "--master-port", str(port), os.path.join(ROOT, "tests", "dist_worker.py"),
           mode, str(tmp)]
    return subprocess.run(cmd, cwd=ROOT, env=env, capture_output=True, text=True, timeout=600)


@pytest.mark.parametrize("mode,port", [("comm", 29711), ("ddp", 29712), ("fsdp:FULL_SHARD", 29713),
                                       ("fsdp:SHARD_GRAD_OP", 29714), ("fsdp:NO_SHARD", 29715),
                                       ("fsdp:FULL_SHARD:offload", 29716), ("rsgemm", 29717), ("aggemm", 29718),
                                       ("fsdp:FULL_SHARD:fuseag", 29719), ("fsdp:FULL_SHARD:smgather", 29720),
                                       ("fsdp:FULL_SHARD:everymicro", 29721), ("fsdp:FULL_SHARD:nographs", 29722), ("nvls", 29723), ("ddp:ga1", 29724), ("ddp:ga3", 29725)])
def test_multi_gpu(mode, port, tmp_path):
    n = _ngpu()
    if n < 2:
        pytest.skip("needs at least 2 GPUs")
    nproc = 2 if n < 4 else (4 if n < 8 else 8)
    r = _torchrun(mode, tmp_path, nproc, port)
    assert r.returncode == 0, (r.stdout[-1500:], r.stderr[-4000:])
    assert r.stdout.count("OK") == nproc
