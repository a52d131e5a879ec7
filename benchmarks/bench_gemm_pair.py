"""1-CTA vs CTA-pair tcgen05 GEMM vs cuBLAS on the model's shapes (CUDA events, L2 flushed, median)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributed_llm_trainer_b200.ops import kernels as k  # noqa: E402
from benchmarks.bench_gemm import timeit  # noqa: E402


def case(name, M, N, K, a_mn, b_mn, epi=k.EPI_BF16, splits=(1,)):
    a = torch.randn((K, M) if a_mn else (M, K), device="cuda").to(torch.bfloat16)
    b = torch.randn((K, N) if b_mn else (N, K), device="cuda").to(torch.bfloat16)
    f32 = epi != k.EPI_BF16
    c = torch.zeros(M, N, device="cuda", dtype=torch.float32 if f32 else torch.bfloat16)
    A = a.t() if a_mn else a
    B = b if b_mn else b.t()
    ref = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
    t_ref = timeit(lambda: torch.matmul(A, B, out=ref))
    best1 = min(timeit(lambda: k.gemm(a, b, c, a_mn=a_mn, b_mn=b_mn, epilogue=epi, split_k=s, tile_n=tn, pair=False))
                for s in splits for tn in (128, 256))
    res2 = {s: timeit(lambda: k.gemm(a, b, c, a_mn=a_mn, b_mn=b_mn, epilogue=epi, split_k=s, pair=True)) for s in splits}
    s2 = min(res2, key=res2.get)
    fl = 2.0 * M * N * K
    print(f"| {name} | {M}x{N}x{K} | {'MN' if a_mn else 'K'}/{'MN' if b_mn else 'K'} | {best1*1e3:.1f} | "
          f"{res2[s2]*1e3:.1f} (split {s2}) | {t_ref*1e3:.1f} | {fl/res2[s2]/1e9:.0f} | {fl/t_ref/1e9:.0f} |")


def main():
    print("| case | MxNxK | layout | 1-CTA us | pair us | cuBLAS us | pair TF | cuBLAS TF |")
    print("|---|---|---|---|---|---|---|---|")
    for h, I, T, tag in ((768, 3072, 8192, "small"), (1024, 4096, 4096, "medium")):
        case(f"{tag} qkv fwd", T, 3 * h, h, False, False)
        case(f"{tag} o fwd", T, h, h, False, False)
        case(f"{tag} gate|up fwd", T, 2 * I, h, False, False)
        case(f"{tag} down fwd", T, h, I, False, False)
        case(f"{tag} qkv dgrad", T, h, 3 * h, False, True)
        case(f"{tag} o dgrad", T, h, h, False, True)
        case(f"{tag} down dgrad", T, I, h, False, True)
        case(f"{tag} gate|up dgrad", T, h, 2 * I, False, True)
        sp = (1, 2, 3, 4, 6, 8)
        case(f"{tag} qkv wgrad", 3 * h, h, T, True, True, k.EPI_F32_ATOMIC, sp)
        case(f"{tag} o wgrad", h, h, T, True, True, k.EPI_F32_ATOMIC, sp)
        case(f"{tag} gate|up wgrad", 2 * I, h, T, True, True, k.EPI_F32_ATOMIC, sp)
        case(f"{tag} down wgrad", h, I, T, True, True, k.EPI_F32_ATOMIC, sp)
    case("lm head fwd (4224-row chunk)", 4224, 50304, 768, False, False)
    case("lm head dX", 4224, 768, 50304, False, True)
    case("lm head dW", 50304, 768, 4224, True, True, k.EPI_F32_ATOMIC, (1,))
    case("square 8192", 8192, 8192, 8192, False, False)


if __name__ == "__main__":
    main()
