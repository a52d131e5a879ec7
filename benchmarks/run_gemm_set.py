"""One launch each of the step's main GEMM configurations (target for `ncu --set full -k regex:gemm2 -c 6`)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributed_llm_trainer_b200.ops import kernels as k  # noqa: E402


def bf(*shape, s=0.1):
    return (torch.randn(*shape, device="cuda") * s).to(torch.bfloat16)


T, h, I, V = 8192, 768, 3072, 50257
x, wgu, gu, act = bf(T, h), bf(2 * I, h), bf(T, 2 * I), bf(T, I)
dgu, dxn = bf(T, 2 * I), bf(T, h)
gw = torch.zeros(2 * I, h, device="cuda")
wd, res, xo = bf(h, I), torch.randn(T, h, device="cuda"), torch.zeros(T, h, device="cuda")
rng = torch.tensor([1, 2], dtype=torch.int64, device="cuda")
Vp = (V + 127) // 128 * 128
emb, logits, xs = bf(Vp, h), torch.zeros(4096, Vp, device="cuda", dtype=torch.bfloat16), bf(4096, h)
torch.cuda.synchronize()
# 1: gate|up forward + SwiGLU epilogue (256-wide tile)
k.gemm(x, wgu, gu, epilogue=k.EPI_SWIGLU, aux=act, half_n=I, tile_n=256)
# 2: gate|up dgrad, N = 768 (384-wide tile, MN-major B)
k.gemm(dgu, wgu, dxn, b_mn=True)
# 3: gate|up wgrad (MN-major A and B, TMA reduce-add epilogue)
k.gemm(dgu, x, gw, a_mn=True, b_mn=True, epilogue=k.EPI_F32_ATOMIC)
# 4: down forward + residual + dropout epilogue, N = 768 (384-wide tile)
k.gemm(act, wd, xo, epilogue=k.EPI_RESID_DROPOUT, residual=res, dropout_p=0.1, rng_state=rng, rng_stream=2)
# 5: down dgrad + SwiGLU' epilogue
k.gemm(dxn, wd, dgu, b_mn=True, epilogue=k.EPI_SWIGLU_BWD, aux=gu, half_n=I, N=I)
# 6: LM head forward chunk
k.gemm(xs, emb, logits, tile_n=256, N=Vp)
torch.cuda.synchronize()
print("ok")
