"""Experiment for docs/ROADMAP.md section 0: does a wgrad GEMM (one 225 KB-smem CTA per SM, ~1 kW) overlap with the
memory-bound RMSNorm backward (3 KB smem, ~600 W) when the two are launched on different streams?

Measures, with CUDA events around the pair (L2 flushed): back to back on one stream, and on two streams in both
launch orders."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributed_llm_trainer_b200.ops import kernels as k  # noqa: E402

M, h, I = 8192, 768, 3072
dgu = (torch.randn(M, 2 * I, device="cuda") * 0.1).to(torch.bfloat16)
xn = torch.randn(M, h, device="cuda").to(torch.bfloat16)
gw = torch.zeros(2 * I, h, device="cuda")
dy = torch.randn(M, h, device="cuda").to(torch.bfloat16)
x = torch.randn(M, h, device="cuda")
w = torch.ones(h, device="cuda", dtype=torch.bfloat16)
rstd = torch.rand(M, device="cuda") + 0.5
dx = torch.randn(M, h, device="cuda")
dw = torch.zeros(h, device="cuda")
dyn = torch.empty(M, h, device="cuda", dtype=torch.bfloat16)
rng = torch.tensor([42, 3], dtype=torch.int64, device="cuda")
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")

gemm = lambda: k.gemm(dgu, xn, gw, a_mn=True, b_mn=True, epilogue=k.EPI_F32_ATOMIC)       # gate|up wgrad
norm = lambda: k.rmsnorm_bwd(dy, x, w, rstd, dx, dw, dy_next=dyn, p_next=0.1, rng_state=rng, rng_stream=5)
side = torch.cuda.Stream()


def timed(fn, iters=20):
    ts = []
    for _ in range(iters):
        flush.zero_()
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b) * 1e3)
    ts.sort()
    return ts[len(ts) // 2]


def two_streams(first, second):
    main = torch.cuda.current_stream()
    side.wait_stream(main)
    first()
    with torch.cuda.stream(side):
        second()
    main.wait_stream(side)


def reps(fn, n=4):
    return lambda: [fn() for _ in range(n)]


for f in (gemm, norm):
    f()
torch.cuda.synchronize()
n = 4
print(f"4 x gemm alone            {timed(reps(gemm, n)):.1f} us")
print(f"4 x norm alone            {timed(reps(norm, n)):.1f} us")
print(f"4 x (gemm, norm) 1 stream {timed(reps(lambda: (gemm(), norm()), n)):.1f} us")
print(f"4 x gemm | norm 2 streams {timed(reps(lambda: two_streams(gemm, norm), n)):.1f} us")
print(f"4 x norm | gemm 2 streams {timed(reps(lambda: two_streams(norm, gemm), n)):.1f} us")
