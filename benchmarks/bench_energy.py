"""Energy per launch of the step's hot kernels (NVML total-energy counter around a >= 0.6 s loop).

The 124M benchmark step runs AT the 1000 W board power cap (nvidia-smi reports sw_power_cap and ~990 W
while bench.py runs), so the step time is set by joules per step as much as by kernel time.  This script
gives the joule side: mJ per launch, time per launch under sustained load (clocks as the governor leaves
them), average power, and each kernel's share of a step's energy for the launch counts of the benchmark
(12 layers x 4 micro-batches of 8 x 1024 tokens).
"""
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributed_llm_trainer_b200.ops import kernels as k  # noqa: E402

import pynvml  # noqa: E402

pynvml.nvmlInit()
H = pynvml.nvmlDeviceGetHandleByIndex(int(os.environ.get("LOCAL_RANK", 0)))


def energy_mj():
    return pynvml.nvmlDeviceGetTotalEnergyConsumption(H)


def measure(fn, min_s=0.6):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    # calibrate the loop length
    t0 = time.perf_counter()
    for _ in range(10):
        fn()
    torch.cuda.synchronize()
    per = (time.perf_counter() - t0) / 10
    n = max(20, int(min_s / max(per, 1e-6)))
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0 = energy_mj()
    a.record()
    for _ in range(n):
        fn()
    b.record()
    torch.cuda.synchronize()
    e1 = energy_mj()
    ms = a.elapsed_time(b)
    return (e1 - e0) / n, ms * 1e3 / n, (e1 - e0) / ms   # mJ/launch, us/launch, W


def bf(*shape, s=0.1):
    return (torch.randn(*shape, device="cuda") * s).to(torch.bfloat16)


def main():
    T, h, I, V, B, S, Hh = 8192, 768, 3072, 50257, 8, 1024, 12
    L, MB = 12, 4
    rng = torch.tensor([42, 3], dtype=torch.int64, device="cuda")
    cases = []   # (name, launches per step, fn)

    x, wqkv, qkv = bf(T, h), bf(3 * h, h), bf(T, 3 * h)
    cos, sin = torch.randn(S, 32, device="cuda"), torch.randn(S, 32, device="cuda")
    cases.append(("qkv fwd (rope)", L * MB, lambda: k.gemm(x, wqkv, qkv, epilogue=k.EPI_ROPE_BF16, rope_cos=cos, rope_sin=sin,
                                                           seq_len=S, rope_cols=2 * h)))
    res, xo, wo = torch.randn(T, h, device="cuda"), torch.zeros(T, h, device="cuda"), bf(h, h)
    cases.append(("o_proj fwd (resid+dropout)", L * MB, lambda: k.gemm(x, wo, xo, epilogue=k.EPI_RESID_DROPOUT, residual=res,
                                                                       dropout_p=0.1, rng_state=rng, rng_stream=1)))
    wgu, gu, act = bf(2 * I, h), bf(T, 2 * I), bf(T, I)
    cases.append(("gate|up fwd (swiglu)", L * MB, lambda: k.gemm(x, wgu, gu, epilogue=k.EPI_SWIGLU, aux=act, half_n=I, tile_n=256)))
    wd = bf(h, I)
    cases.append(("down fwd (resid+dropout)", L * MB, lambda: k.gemm(act, wd, xo, epilogue=k.EPI_RESID_DROPOUT, residual=res,
                                                                     dropout_p=0.1, rng_state=rng, rng_stream=2)))
    dy, dgu, dxn = bf(T, h), bf(T, 2 * I), bf(T, h)
    cases.append(("down dgrad (swiglu')", L * MB, lambda: k.gemm(dy, wd, dgu, b_mn=True, epilogue=k.EPI_SWIGLU_BWD, aux=gu, half_n=I, N=I)))
    cases.append(("gate|up dgrad", L * MB, lambda: k.gemm(dgu, wgu, dxn, b_mn=True)))
    cases.append(("o_proj dgrad", L * MB, lambda: k.gemm(dy, wo, dxn, b_mn=True)))
    cases.append(("qkv dgrad", L * MB, lambda: k.gemm(qkv, wqkv, dxn, b_mn=True)))
    gw_gu, gw_d = torch.zeros(2 * I, h, device="cuda"), torch.zeros(h, I, device="cuda")
    gw_qkv, gw_o = torch.zeros(3 * h, h, device="cuda"), torch.zeros(h, h, device="cuda")
    cases.append(("gate|up wgrad", L * MB, lambda: k.gemm(dgu, x, gw_gu, a_mn=True, b_mn=True, epilogue=k.EPI_F32_ATOMIC)))
    cases.append(("down wgrad", L * MB, lambda: k.gemm(dy, act, gw_d, a_mn=True, b_mn=True, epilogue=k.EPI_F32_ATOMIC, split_k=2)))
    cases.append(("qkv wgrad", L * MB, lambda: k.gemm(qkv, x, gw_qkv, a_mn=True, b_mn=True, epilogue=k.EPI_F32_ATOMIC, split_k=2)))
    cases.append(("o_proj wgrad", L * MB, lambda: k.gemm(dy, x, gw_o, a_mn=True, b_mn=True, epilogue=k.EPI_F32_ATOMIC, split_k=8)))

    # cuBLAS on the plain-epilogue shapes, for the joules-per-FLOP comparison
    ref_out = torch.empty(T, h, device="cuda", dtype=torch.bfloat16)
    cases.append(("cuBLAS gate|up dgrad shape", 0, lambda: torch.matmul(dgu, wgu, out=ref_out)))
    cases.append(("cuBLAS qkv dgrad shape", 0, lambda: torch.matmul(qkv, wqkv, out=ref_out)))
    ref_gu = torch.empty(T, 2 * I, device="cuda", dtype=torch.bfloat16)
    cases.append(("cuBLAS gate|up fwd shape (no swiglu)", 0, lambda: torch.matmul(x, wgu.t(), out=ref_gu)))
    big_a, big_b, big_c = bf(8192, 8192), bf(8192, 8192), torch.empty(8192, 8192, device="cuda", dtype=torch.bfloat16)
    cases.append(("cuBLAS 8192^3", 0, lambda: torch.matmul(big_a, big_b.t(), out=big_c)))
    cases.append(("ours   8192^3", 0, lambda: k.gemm(big_a, big_b, big_c)))

    out, lse = torch.zeros(T, h, device="cuda", dtype=torch.bfloat16), torch.zeros(B, Hh, S, device="cuda")
    dqkv, delta, dq = torch.zeros(T, 3 * h, device="cuda", dtype=torch.bfloat16), torch.zeros(B, Hh, S, device="cuda"), torch.empty(T, h, device="cuda")
    qkv_s = bf(T, 3 * h, s=0.8)
    cases.append(("attention fwd", L * MB, lambda: k.attn_fwd(qkv_s, out, lse, B, S, Hh, 0.1, rng, 9)))
    k.attn_fwd(qkv_s, out, lse, B, S, Hh, 0.1, rng, 9)
    cases.append(("attention bwd (+delta, dq cast)", L * MB, lambda: k.attn_bwd(qkv_s, out, dy, lse, dqkv, delta, dq, B, S, Hh, 0.1, rng, 9)))

    Vp = (V + 7) // 8 * 8
    rows = 4224
    emb, logits = bf(Vp, h), torch.zeros(rows, Vp, device="cuda", dtype=torch.bfloat16)
    xs = bf(rows, h)
    cases.append(("lm head fwd (chunk)", 2 * MB, lambda: k.gemm(xs, emb, logits, tile_n=256, N=Vp)))
    labels = torch.randint(0, V, (rows,), device="cuda")
    loss = torch.zeros(1, device="cuda")
    cases.append(("cross entropy fwd+bwd (chunk)", 2 * MB, lambda: k.ce_fwd_bwd(logits, labels, loss, V, 1.0, 1.0)))
    dxb, gemb = torch.zeros(rows, h, device="cuda", dtype=torch.bfloat16), torch.zeros(Vp, h, device="cuda")
    cases.append(("lm head dX (chunk)", 2 * MB, lambda: k.gemm(logits, emb, dxb, b_mn=True, tile_n=256, K=V)))
    cases.append(("lm head dW (chunk)", 2 * MB, lambda: k.gemm(logits, xs, gemb, a_mn=True, b_mn=True, epilogue=k.EPI_F32_ATOMIC, tile_n=256, M=V, K=rows)))

    xf, wln, y, rstd = torch.randn(T, h, device="cuda"), bf(h), torch.zeros(T, h, device="cuda", dtype=torch.bfloat16), torch.zeros(T, device="cuda")
    cases.append(("rmsnorm fwd", (2 * L + 1) * MB, lambda: k.rmsnorm_fwd(xf, wln, y, rstd, 1e-6)))
    k.rmsnorm_fwd(xf, wln, y, rstd, 1e-6)
    dxa, dwa = torch.zeros(T, h, device="cuda"), torch.zeros(h, device="cuda")
    cases.append(("rmsnorm bwd", (2 * L + 1) * MB, lambda: k.rmsnorm_bwd(y, xf, wln, rstd, dxa, dwa)))

    print("| kernel | launches/step | us/launch (loaded) | mJ/launch | W | J/step |")
    print("|---|---|---|---|---|---|")
    tot_j = tot_ms = 0.0
    for name, n, fn in cases:
        mj, us, w = measure(fn)
        tot_j += mj * n / 1e3
        tot_ms += us * n / 1e3
        print(f"| {name} | {n} | {us:.1f} | {mj:.1f} | {w:.0f} | {mj * n / 1e3:.2f} |")
    print(f"\nsum over listed kernels: {tot_j:.1f} J/step, {tot_ms:.1f} ms/step of back-to-back launches")
    torch.cuda.synchronize()
    e0 = energy_mj()
    time.sleep(1.0)
    print(f"idle power (context alive, no kernels): {(energy_mj() - e0) / 1000.0:.0f} W")


if __name__ == "__main__":
    main()
