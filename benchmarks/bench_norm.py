"""RMSNorm backward (+ dropout-masked bf16 dY of the next branch) at the benchmark shape, vs its HBM traffic."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributed_llm_trainer_b200.ops import kernels as k  # noqa: E402
from benchmarks.bench_gemm import timeit  # noqa: E402

rows, h = 8192, 768
dy = torch.randn(rows, h, device="cuda").to(torch.bfloat16)
x = torch.randn(rows, h, device="cuda")
w = torch.ones(h, device="cuda", dtype=torch.bfloat16)
rstd = torch.rand(rows, device="cuda") + 0.5
dx = torch.randn(rows, h, device="cuda")
dw = torch.zeros(h, device="cuda")
dyn = torch.empty(rows, h, device="cuda", dtype=torch.bfloat16)
rng = torch.tensor([42, 3], dtype=torch.int64, device="cuda")
f = lambda: k.rmsnorm_bwd(dy, x, w, rstd, dx, dw, dy_next=dyn, p_next=0.1, rng_state=rng, rng_stream=5)
f()
torch.cuda.synchronize()
t = timeit(f)
mb = rows * h * (2 + 4 + 4 + 4 + 2) / 1e6
print(f"rmsnorm_bwd {t * 1e3:.1f} us incl. ~6 us of launch + event overhead  ({mb / t:.0f} GB/s)")
