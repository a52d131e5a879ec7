"""Launch one GEMM configuration a few times (target for `ncu -k regex:... -c 1`)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributed_llm_trainer_b200.ops import kernels as k  # noqa: E402
from distributed_llm_trainer_b200.ops.reference import rope_tables  # noqa: E402

which = sys.argv[1] if len(sys.argv) > 1 else "swiglu"
T, h, I, S = 8192, 768, 3072, 1024
x = (torch.randn(T, h, device="cuda")).to(torch.bfloat16)
if which == "swiglu":
    w = (torch.randn(2 * I, h, device="cuda") * 0.05).to(torch.bfloat16)
    gu = torch.empty(T, 2 * I, device="cuda", dtype=torch.bfloat16)
    act = torch.empty(T, I, device="cuda", dtype=torch.bfloat16)
    fn = lambda: k.gemm(x, w, gu, epilogue=k.EPI_SWIGLU, aux=act, half_n=I, pair=True)
elif which == "plain":
    w = (torch.randn(2 * I, h, device="cuda") * 0.05).to(torch.bfloat16)
    gu = torch.empty(T, 2 * I, device="cuda", dtype=torch.bfloat16)
    fn = lambda: k.gemm(x, w, gu, pair=True)
elif which == "rope":
    w = (torch.randn(3 * h, h, device="cuda") * 0.05).to(torch.bfloat16)
    qkv = torch.empty(T, 3 * h, device="cuda", dtype=torch.bfloat16)
    cos, sin = rope_tables(S, 64, device="cuda")
    fn = lambda: k.gemm(x, w, qkv, epilogue=k.EPI_ROPE_BF16, rope_cos=cos, rope_sin=sin, seq_len=S, rope_cols=2 * h, pair=True)
for _ in range(5):
    fn()
torch.cuda.synchronize()
print("ok", which)
