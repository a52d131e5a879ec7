"""One launch each of the step's memory-bound kernels (target for `ncu --set full -k regex:<names> -c N`)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributed_llm_trainer_b200.ops import kernels as k  # noqa: E402

T, h, V = 8192, 768, 50257
Vp = (V + 127) // 128 * 128
rng = torch.tensor([1, 2], dtype=torch.int64, device="cuda")
x = torch.randn(T, h, device="cuda")
w = (torch.rand(h, device="cuda") + 0.5).to(torch.bfloat16)
y = torch.empty(T, h, device="cuda", dtype=torch.bfloat16)
rstd = torch.empty(T, device="cuda")
dy = torch.randn(T, h, device="cuda").to(torch.bfloat16)
dx, dw = torch.zeros(T, h, device="cuda"), torch.zeros(h, device="cuda")
dyn = torch.empty(T, h, device="cuda", dtype=torch.bfloat16)
logits = (torch.randn(4096, Vp, device="cuda") * 2).to(torch.bfloat16)
labels = torch.randint(0, V, (4096,), device="cuda")
loss = torch.zeros(1, device="cuda")
n = 151_862_784 // 128 * 128
master, grad = torch.randn(n, device="cuda") * 0.02, torch.randn(n, device="cuda") * 1e-3
m, v = torch.zeros(n, device="cuda"), torch.zeros(n, device="cuda")
shadow = torch.empty(n, device="cuda", dtype=torch.bfloat16)
decay = torch.ones(n // 128, dtype=torch.uint8, device="cuda")
scale = torch.ones(1, device="cuda")
torch.cuda.synchronize()
k.rmsnorm_fwd(x, w, y, rstd, 1e-6)
k.rmsnorm_bwd(dy, x, w, rstd, dx, dw, dy_next=dyn, p_next=0.1, rng_state=rng, rng_stream=3)
k.ce_fwd_bwd(logits, labels, loss, V, 1.0, 1.0)
k.adamw_step(master, grad, m, v, shadow, decay, 0, n, 1, 6e-4, 0.9, 0.95, 1e-8, 0.1, scale)
torch.cuda.synchronize()
print("ok")
