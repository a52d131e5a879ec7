"""256-wide vs 384-wide CTA-pair tile on the N = 768 / 1024 / 1536 GEMMs of the model (CUDA events, median)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributed_llm_trainer_b200.ops import kernels as k  # noqa: E402
from benchmarks.bench_gemm import timeit  # noqa: E402


def mk(*shape):
    return (torch.randn(*shape, device="cuda") * 0.1).to(torch.bfloat16)


def main():
    M = 8192
    rows = []
    for name, N, K, b_mn in (("o_proj fwd", 768, 768, False), ("down fwd", 768, 3072, False),
                             ("qkv dgrad", 768, 2304, True), ("o_proj dgrad", 768, 768, True),
                             ("gate_up dgrad", 768, 6144, True), ("N=1536 K=1536", 1536, 1536, False)):
        a = mk(M, K)
        b = mk(K, N) if b_mn else mk(N, K)
        c = torch.zeros(M, N, device="cuda", dtype=torch.bfloat16)
        t = {tn: timeit(lambda: k.gemm(a, b, c, b_mn=b_mn, pair=True, tile_n=tn)) for tn in (256, 384, 0)}
        fl = 2.0 * M * N * K
        rows.append(f"| {name} | {M}x{N}x{K} | {t[256]*1e3:.1f} us ({fl/t[256]/1e9:.0f}) | {t[384]*1e3:.1f} us ({fl/t[384]/1e9:.0f}) | {t[0]*1e3:.1f} us |")
    print("| GEMM | MxNxK | tile 256 (TFLOP/s) | tile 384 (TFLOP/s) | launcher's choice |")
    print("|---|---|---|---|---|")
    print("\n".join(rows))


if __name__ == "__main__":
    main()
