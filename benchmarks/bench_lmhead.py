"""The LM-head stage of one micro-step chunk, kernel by kernel (small preset: 4224 rows x 768 x V=50257):
logits GEMM, softmax-CE, dX GEMM (split-K), dW GEMM -- pair kernel with both tile orders vs cuBLAS (torch.matmul).
CUDA events, L2 flushed between iterations, median of 20."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from benchmarks.bench_gemm import timeit  # noqa: E402
from distributed_llm_trainer_b200.ops import kernels as k  # noqa: E402

rows, h, V = int(os.environ.get("LM_ROWS", 4224)), 768, 50257
Vp = (V + 127) // 128 * 128
torch.manual_seed(0)
x = (torch.randn(rows, h, device="cuda") * 0.5).to(torch.bfloat16)
w = (torch.randn(Vp, h, device="cuda") * 0.02).to(torch.bfloat16)
logits = torch.empty(rows, Vp, device="cuda", dtype=torch.bfloat16)
labels = torch.randint(0, V, (rows,), device="cuda")
loss = torch.zeros(1, device="cuda")
dx32 = torch.zeros(rows, h, device="cuda")
gw = torch.zeros(V, h, device="cuda")
fl = 2.0 * rows * Vp * h


def report(name, t, flops=fl, bytes_=None):
    extra = f"  {bytes_ / t / 1e6:.0f} GB/s" if bytes_ else ""
    print(f"{name:46s} {t * 1e3:7.1f} us  {flops / t / 1e9:6.0f} TFLOP/s{extra}" if flops else f"{name:46s} {t * 1e3:7.1f} us{extra}")


for r in (1, 2):
    report(f"logits GEMM raster={r}", timeit(lambda: k.gemm(x, w, logits, tile_n=256, N=Vp, raster=r)))
report("logits cuBLAS", timeit(lambda: torch.matmul(x, w.t(), out=logits)))
k.gemm(x, w, logits, tile_n=256, N=Vp)
keep = logits.clone()
report("softmax-CE (in place)", timeit(lambda: k.ce_fwd_bwd(logits, labels, loss, V, 1e-4, 1e-4)), flops=0,
       bytes_=2.0 * rows * Vp * 2)
logits.copy_(keep)
k.ce_fwd_bwd(logits, labels, loss, V, 1e-4, 1e-4)
for r in (1, 2):
    for split in (1, 3, 4, 7, 10):
        report(f"dX GEMM split={split} raster={r}",
               timeit(lambda: k.gemm(logits, w, dx32, b_mn=True, epilogue=k.EPI_F32_ATOMIC, split_k=split, tile_n=256, K=V,
                                     raster=r)))
dxb = torch.empty(rows, h, device="cuda", dtype=torch.bfloat16)
report("dX cuBLAS", timeit(lambda: torch.matmul(logits[:, :V], w[:V], out=dxb)))
for r in (1, 2):
    report(f"dW GEMM raster={r}",
           timeit(lambda: k.gemm(logits, x, gw, a_mn=True, b_mn=True, epilogue=k.EPI_F32_ATOMIC, tile_n=256, M=V, K=rows,
                                 raster=r)))
gwb = torch.empty(V, h, device="cuda", dtype=torch.bfloat16)
report("dW cuBLAS", timeit(lambda: torch.matmul(logits[:, :V].t(), x, out=gwb)))
