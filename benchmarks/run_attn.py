"""Attention fwd+bwd at the benchmark shape: timing (default) or a target for `ncu -k regex:attn_`."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributed_llm_trainer_b200.ops import kernels as k  # noqa: E402
from benchmarks.bench_gemm import timeit  # noqa: E402

B, S, H = 8, 1024, 12
p = float(sys.argv[1]) if len(sys.argv) > 1 else 0.1
h = H * 64
qkv = (torch.randn(B * S, 3 * h, device="cuda") * 0.8).to(torch.bfloat16)
dout = (torch.randn(B * S, h, device="cuda") * 0.5).to(torch.bfloat16)
rng = torch.tensor([42, 3], dtype=torch.int64, device="cuda")
out = torch.zeros(B * S, h, device="cuda", dtype=torch.bfloat16)
lse = torch.zeros(B, H, S, device="cuda")
dqkv = torch.zeros(B * S, 3 * h, device="cuda", dtype=torch.bfloat16)
delta = torch.zeros(B, H, S, device="cuda")
dq_acc = torch.empty(B * S, h, device="cuda")
fwd = lambda: k.attn_fwd(qkv, out, lse, B, S, H, p, rng, 9)
bwd = lambda: k.attn_bwd(qkv, out, dout, lse, dqkv, delta, dq_acc, B, S, H, p, rng, 9)
ang = torch.rand(S, 32, device="cuda") * 6.28
cos, sin = ang.cos().contiguous(), ang.sin().contiguous()
# what the engine runs: inverse RoPE of dQ|dK fused into the final pass; compare with bwd + rope_inplace
bwd_rope = lambda: k.attn_bwd(qkv, out, dout, lse, dqkv, delta, dq_acc, B, S, H, p, rng, 9, rope_cos=cos, rope_sin=sin)
bwd_then_rope = lambda: (bwd(), k.rope_inplace(dqkv, cos, sin, 2 * h, S, inverse=True))
fwd(); bwd(); torch.cuda.synchronize()
if os.environ.get("ATTN_TIME", "1") == "1":
    fl = 4.0 * B * H * S * S * 64 / 2
    tf, tb = timeit(fwd), timeit(bwd)
    print(f"p={p}: fwd {tf*1e3:.1f} us ({fl/tf/1e9:.0f} TFLOP/s causal)  bwd {tb*1e3:.1f} us ({2.5*fl/tb/1e9:.0f} TFLOP/s)")
    tr, tu = timeit(bwd_rope), timeit(bwd_then_rope)
    print(f"p={p}: bwd + inverse RoPE fused {tr*1e3:.1f} us   unfused (bwd, then rope_inplace) {tu*1e3:.1f} us")
