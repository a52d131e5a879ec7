"""Fused-epilogue GEMMs: single-CTA kernel vs CTA-pair kernel (both tile widths) vs unfused alternative."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributed_llm_trainer_b200.ops import kernels as k  # noqa: E402
from distributed_llm_trainer_b200.ops.reference import rope_tables  # noqa: E402
from benchmarks.bench_gemm import timeit  # noqa: E402


def mk(*shape, s=1.0):
    return (torch.randn(*shape, device="cuda") * s).to(torch.bfloat16)


def row(name, fns):
    print(f"| {name} | " + " | ".join(f"{lbl} {timeit(fn)*1e3:.1f}" for lbl, fn in fns) + " |")


def main():
    for h, I, T, S, tag in ((768, 3072, 8192, 1024, "small"), (1024, 4096, 4096, 1024, "medium")):
        x, wqkv = mk(T, h), mk(3 * h, h, s=0.05)
        qkv = torch.empty(T, 3 * h, device="cuda", dtype=torch.bfloat16)
        cos, sin = rope_tables(S, 64, device="cuda")
        kw = dict(epilogue=k.EPI_ROPE_BF16, rope_cos=cos, rope_sin=sin, seq_len=S, rope_cols=2 * h)
        row(f"{tag} qkv+rope", [("1cta", lambda: k.gemm(x, wqkv, qkv, pair=False, **kw)),
                                ("pair256", lambda: k.gemm(x, wqkv, qkv, pair=True, tile_n=256, **kw)),
                                ("pair128", lambda: k.gemm(x, wqkv, qkv, pair=True, tile_n=128, **kw)),
                                ("plain+rope", lambda: (k.gemm(x, wqkv, qkv, pair=True), k.rope_inplace(qkv, cos, sin, 2 * h, S)))])
        attn, wo = mk(T, h), mk(h, h, s=0.05)
        res, out = torch.randn(T, h, device="cuda"), torch.empty(T, h, device="cuda")
        rng = torch.tensor([1, 2], dtype=torch.int64, device="cuda")
        kw = dict(epilogue=k.EPI_RESID_DROPOUT, residual=res, dropout_p=0.1, rng_state=rng, rng_stream=3)
        y = torch.empty(T, h, device="cuda", dtype=torch.bfloat16)
        row(f"{tag} o_proj+resid", [("1cta", lambda: k.gemm(attn, wo, out, pair=False, **kw)),
                                    ("pair256", lambda: k.gemm(attn, wo, out, pair=True, tile_n=256, **kw)),
                                    ("pair128", lambda: k.gemm(attn, wo, out, pair=True, tile_n=128, **kw)),
                                    ("plain+ew", lambda: (k.gemm(attn, wo, y, pair=True), k.resid_dropout_fwd(y, res, out, 0.1, rng, 3)))])
        act, wd = mk(T, I), mk(h, I, s=0.05)
        row(f"{tag} down+resid", [("1cta", lambda: k.gemm(act, wd, out, pair=False, **kw)),
                                  ("pair256", lambda: k.gemm(act, wd, out, pair=True, tile_n=256, **kw)),
                                  ("pair128", lambda: k.gemm(act, wd, out, pair=True, tile_n=128, **kw)),
                                  ("plain+ew", lambda: (k.gemm(act, wd, y, pair=True), k.resid_dropout_fwd(y, res, out, 0.1, rng, 3)))])
        wgu = mk(2 * I, h, s=0.05)
        gu = torch.empty(T, 2 * I, device="cuda", dtype=torch.bfloat16)
        a2 = torch.empty(T, I, device="cuda", dtype=torch.bfloat16)
        kw = dict(epilogue=k.EPI_SWIGLU, aux=a2, half_n=I)
        row(f"{tag} gate|up+swiglu", [("1cta", lambda: k.gemm(x, wgu, gu, pair=False, tile_n=256, **kw)),
                                      ("pair", lambda: k.gemm(x, wgu, gu, pair=True, **kw)),
                                      ("plain+ew", lambda: (k.gemm(x, wgu, gu, pair=True), k.swiglu_fwd(gu, a2)))])
        dy = mk(T, h)
        dgu = torch.empty(T, 2 * I, device="cuda", dtype=torch.bfloat16)
        dact = torch.empty(T, I, device="cuda", dtype=torch.bfloat16)
        kw = dict(b_mn=True, epilogue=k.EPI_SWIGLU_BWD, aux=gu, half_n=I, N=I)
        row(f"{tag} down dgrad+swiglu'", [("1cta", lambda: k.gemm(dy, wd, dgu, pair=False, **kw)),
                                          ("pair256", lambda: k.gemm(dy, wd, dgu, pair=True, tile_n=256, **kw)),
                                          ("pair128", lambda: k.gemm(dy, wd, dgu, pair=True, tile_n=128, **kw)),
                                          ("plain+ew", lambda: (k.gemm(dy, wd, dact, b_mn=True, pair=True), k.swiglu_bwd(dact, gu, dgu)))])
        # plain narrow-N GEMMs: which pair tile?
        dgu_in, dxn = mk(T, 2 * I), torch.empty(T, h, device="cuda", dtype=torch.bfloat16)
        row(f"{tag} gate|up dgrad", [("pair256", lambda: k.gemm(dgu_in, wgu, dxn, b_mn=True, pair=True, tile_n=256)),
                                     ("pair128", lambda: k.gemm(dgu_in, wgu, dxn, b_mn=True, pair=True, tile_n=128)),
                                     ("auto", lambda: k.gemm(dgu_in, wgu, dxn, b_mn=True, pair=True))])
        row(f"{tag} down fwd plain", [("pair256", lambda: k.gemm(act, wd, y, pair=True, tile_n=256)),
                                      ("pair128", lambda: k.gemm(act, wd, y, pair=True, tile_n=128)),
                                      ("auto", lambda: k.gemm(act, wd, y, pair=True))])


if __name__ == "__main__":
    main()
