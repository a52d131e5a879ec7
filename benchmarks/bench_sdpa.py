"""The library baselines the attention kernels must be judged against: torch
``F.scaled_dot_product_attention`` with the flash, cuDNN and memory-efficient backends, forward and forward+backward,
at the benchmark shape (B=8, S=1024, H=12, D=64, causal, bf16) with dropout 0.1 and 0 -- same timing rules as
benchmarks/run_attn.py (CUDA events, L2 flushed between iterations, median of 20).  Also the reference's manual
attention path (gpt.py:230-234) for scale.  Prints a markdown table (causal FLOPs: fwd 12.9 G, bwd 32.2 G)."""
import os
import sys

import torch
import torch.nn.functional as F
from torch.nn.attention import SDPBackend, sdpa_kernel

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from benchmarks.bench_gemm import timeit  # noqa: E402
from distributed_llm_trainer_b200.ops import kernels as k  # noqa: E402

B, S, H, D = 8, 1024, 12, 64
h = H * D
FL = 4.0 * B * H * S * S * D / 2
torch.manual_seed(0)
rows = []


def add(name, p, tf, tfb):
    rows.append((name, p, tf, tfb))
    fb = "-" if tfb is None else f"{tfb * 1e3:7.1f} us ({3.5 * FL / tfb / 1e9:4.0f})"
    print(f"| {name:34s} | {p:3.1f} | {tf * 1e3:7.1f} us ({FL / tf / 1e9:4.0f}) | {fb} |", flush=True)


print("| implementation | dropout | forward (TFLOP/s causal) | forward + backward (TFLOP/s) |\n|---|---|---|---|")
for p in (0.1, 0.0):
    qkv = (torch.randn(B * S, 3 * h, device="cuda") * 0.8).to(torch.bfloat16)
    dout = (torch.randn(B * S, h, device="cuda") * 0.5).to(torch.bfloat16)
    rng = torch.tensor([42, 3], dtype=torch.int64, device="cuda")
    out = torch.zeros(B * S, h, device="cuda", dtype=torch.bfloat16)
    lse = torch.zeros(B, H, S, device="cuda")
    dqkv = torch.zeros(B * S, 3 * h, device="cuda", dtype=torch.bfloat16)
    delta = torch.zeros(B, H, S, device="cuda")
    dq_acc = torch.empty(B * S, h, device="cuda")
    fwd = lambda: k.attn_fwd(qkv, out, lse, B, S, H, p, rng, 9)
    both = lambda: (k.attn_fwd(qkv, out, lse, B, S, H, p, rng, 9),
                    k.attn_bwd(qkv, out, dout, lse, dqkv, delta, dq_acc, B, S, H, p, rng, 9))
    add("this repo (tcgen05 flash, fused)", p, timeit(fwd), timeit(both))
    q, kk, v = [t.reshape(B, S, H, D).transpose(1, 2).contiguous().requires_grad_(True)
                for t in qkv.view(B * S, 3, h).unbind(1)]
    g = dout.view(B, S, H, D).transpose(1, 2).contiguous()
    for name, be in (("torch SDPA flash", SDPBackend.FLASH_ATTENTION), ("torch SDPA cuDNN", SDPBackend.CUDNN_ATTENTION),
                     ("torch SDPA mem-efficient", SDPBackend.EFFICIENT_ATTENTION)):
        try:
            with sdpa_kernel(be):
                f = lambda: F.scaled_dot_product_attention(q, kk, v, dropout_p=p, is_causal=True)

                def fb():
                    o = F.scaled_dot_product_attention(q, kk, v, dropout_p=p, is_causal=True)
                    o.backward(g)

                f()
                fb()
                add(name, p, timeit(f), timeit(fb))
        except Exception as e:  # noqa: BLE001
            print(f"| {name} | {p} | unavailable: {type(e).__name__}: {str(e)[:80]} | |")

    def manual():          # the reference's as-run path: QK^T, triu mask, fp32 softmax, dropout, @V
        sc = (q @ kk.transpose(-2, -1)) * (D ** -0.5)
        sc = sc.masked_fill(torch.ones(S, S, dtype=torch.bool, device="cuda").triu(1), float("-inf"))
        pr = F.dropout(F.softmax(sc.float(), dim=-1).to(q.dtype), p)
        return pr @ v

    add("reference manual path (torch ops)", p, timeit(manual, iters=5), timeit(lambda: manual().backward(g), iters=5))
