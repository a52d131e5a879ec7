"""tcgen05 GEMM vs cuBLAS (torch.matmul) on the shapes the GPT small/medium step issues.
CUDA-event timed, L2 flushed between iterations, median of N.  Prints a markdown table."""
import json
import sys
import os

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributed_llm_trainer_b200.ops import kernels as k  # noqa: E402

PEAK = 1670.1e12
flush = None


def timeit(fn, iters=20, warm=3):
    global flush
    if flush is None:
        flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device="cuda")
    for _ in range(warm):
        fn()
    ts = []
    for _ in range(iters):
        flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    ts.sort()
    return ts[len(ts) // 2]


def case(name, M, N, K, a_mn, b_mn, epi=k.EPI_BF16, split_k=1, tile_n=0):
    a = torch.randn((K, M) if a_mn else (M, K), device="cuda").to(torch.bfloat16)
    b = torch.randn((K, N) if b_mn else (N, K), device="cuda").to(torch.bfloat16)
    f32 = epi in (k.EPI_F32, k.EPI_F32_ATOMIC)
    c = torch.zeros(M, N, device="cuda", dtype=torch.float32 if f32 else torch.bfloat16)
    A = a.t() if a_mn else a
    B = b if b_mn else b.t()
    ref = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
    t_ref = timeit(lambda: torch.matmul(A, B, out=ref))
    t_our = timeit(lambda: k.gemm(a, b, c, a_mn=a_mn, b_mn=b_mn, epilogue=epi, split_k=split_k, tile_n=tile_n))
    fl = 2.0 * M * N * K
    return dict(name=name, M=M, N=N, K=K, layout=f"{'MN' if a_mn else 'K'}/{'MN' if b_mn else 'K'}", tile_n=tile_n,
                split_k=split_k, ours_us=t_our * 1e3, cublas_us=t_ref * 1e3, ours_tflops=fl / t_our / 1e9,
                cublas_tflops=fl / t_ref / 1e9, frac_peak=fl / (t_our * 1e-3) / PEAK)


def main():
    rows = []
    T = 8192
    for h, I, tag in ((768, 3072, "small"), (1024, 4096, "medium")):
        Tm = T if tag == "small" else 4096
        for tn in (128, 256):
            rows.append(case(f"{tag} qkv fwd", Tm, 3 * h, h, False, False, tile_n=tn))
            rows.append(case(f"{tag} o fwd", Tm, h, h, False, False, tile_n=tn))
            rows.append(case(f"{tag} gate|up fwd", Tm, 2 * I, h, False, False, tile_n=tn))
            rows.append(case(f"{tag} down fwd", Tm, h, I, False, False, tile_n=tn))
            rows.append(case(f"{tag} qkv dgrad", Tm, h, 3 * h, False, True, tile_n=tn))
            rows.append(case(f"{tag} down dgrad", Tm, I, h, False, True, tile_n=tn))
            rows.append(case(f"{tag} gate|up dgrad", Tm, h, 2 * I, False, True, tile_n=tn))
        for sk in (1, 2, 4, 8):
            rows.append(case(f"{tag} qkv wgrad", 3 * h, h, Tm, True, True, epi=k.EPI_F32_ATOMIC, split_k=sk, tile_n=256))
            rows.append(case(f"{tag} o wgrad", h, h, Tm, True, True, epi=k.EPI_F32_ATOMIC, split_k=sk, tile_n=128))
            rows.append(case(f"{tag} gate|up wgrad", 2 * I, h, Tm, True, True, epi=k.EPI_F32_ATOMIC, split_k=sk, tile_n=256))
            rows.append(case(f"{tag} down wgrad", h, I, Tm, True, True, epi=k.EPI_F32_ATOMIC, split_k=sk, tile_n=256))
    rows.append(case("lm head fwd chunk", 2048, 50304, 768, False, False, tile_n=256))
    rows.append(case("lm head dX chunk", 2048, 768, 50304, False, True, tile_n=128))
    rows.append(case("lm head dW chunk", 50304, 768, 2048, True, True, epi=k.EPI_F32_ATOMIC, tile_n=256))
    rows.append(case("square 8192", 8192, 8192, 8192, False, False, tile_n=256))
    print("| case | M | N | K | layout | BN | splitK | ours us | cuBLAS us | ours TF | cuBLAS TF | frac of measured peak |")
    print("|---|---|---|---|---|---|---|---|---|---|---|---|")
    for r in rows:
        print(f"| {r['name']} | {r['M']} | {r['N']} | {r['K']} | {r['layout']} | {r['tile_n']} | {r['split_k']} | "
              f"{r['ours_us']:.1f} | {r['cublas_us']:.1f} | {r['ours_tflops']:.0f} | {r['cublas_tflops']:.0f} | {r['frac_peak']:.2f} |")
    os.makedirs("gpurun_out", exist_ok=True)
    json.dump(rows, open("gpurun_out/bench_gemm.json", "w"), indent=1)


if __name__ == "__main__":
    main()
