"""Split-K sweep for the reduce-add GEMMs whose tile count does not fill the 74 CTA pairs (CUDA events, median)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributed_llm_trainer_b200.ops import kernels as k  # noqa: E402
from benchmarks.bench_gemm import timeit  # noqa: E402


def bf(*shape, s=0.1):
    return (torch.randn(*shape, device="cuda") * s).to(torch.bfloat16)


def main():
    T, h, I, V = 8192, 768, 3072, 50257
    Vp = (V + 127) // 128 * 128
    rows = 4096
    logits, emb = bf(rows, Vp), bf(Vp, h)
    dx32 = torch.zeros(rows, h, device="cuda")
    qkv, x = bf(T, 3 * h), bf(T, h)
    gw_qkv = torch.zeros(3 * h, h, device="cuda")
    dy, act = bf(T, h), bf(T, I)
    gw_d, gw_o = torch.zeros(h, I, device="cuda"), torch.zeros(h, h, device="cuda")
    cases = [
        ("lm head dX chunk (4096x768, K=50257)", 2.0 * rows * h * V,
         lambda s: k.gemm(logits, emb, dx32, b_mn=True, epilogue=k.EPI_F32_ATOMIC, split_k=s, tile_n=256, K=V)),
        ("qkv wgrad (2304x768, K=8192)", 2.0 * 3 * h * h * T,
         lambda s: k.gemm(qkv, x, gw_qkv, a_mn=True, b_mn=True, epilogue=k.EPI_F32_ATOMIC, split_k=s)),
        ("down wgrad (768x3072, K=8192)", 2.0 * h * I * T,
         lambda s: k.gemm(dy, act, gw_d, a_mn=True, b_mn=True, epilogue=k.EPI_F32_ATOMIC, split_k=s)),
        ("o_proj wgrad (768x768, K=8192)", 2.0 * h * h * T,
         lambda s: k.gemm(dy, x, gw_o, a_mn=True, b_mn=True, epilogue=k.EPI_F32_ATOMIC, split_k=s)),
    ]
    splits = (1, 2, 3, 4, 5, 6, 8, 12, 16)
    print("| GEMM | " + " | ".join(f"split {s}" for s in splits) + " |")
    print("|---|" + "---|" * len(splits))
    for name, fl, fn in cases:
        cells = []
        for s in splits:
            t = timeit(lambda: fn(s))
            cells.append(f"{t * 1e3:.1f} us ({fl / t / 1e9:.0f})")
        print(f"| {name} | " + " | ".join(cells) + " |")


if __name__ == "__main__":
    main()
