"""ncu target for the fused compute+collective kernels on ONE GPU (self-peer: the "owners" / "ranks" are separate local
buffers, so the kernels run exactly the instructions they run across NVLink, against local addresses):

    wgrad GEMM -> reduce-scatter (gate|up shape, world 8, with and without the carried accumulator), the same GEMM
    accumulating locally, embed_bwd_rs, sumsq_chunks, adamw_chunks (per-rank stores), pull_reduce_chunks.

ATTN_TIME-style switch: FUSED_TIME=1 prints CUDA-event timings instead (L2 flushed, median of 20)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from benchmarks.bench_gemm import timeit  # noqa: E402
from distributed_llm_trainer_b200.ops import kernels as K  # noqa: E402
from distributed_llm_trainer_b200.parallel.ownership import SlabGroup  # noqa: E402

DEV = "cuda"
world = 8
T, n_out, k_in = 8192, 6144, 768
torch.manual_seed(0)
dy = (torch.randn(T, n_out, device=DEV) * 0.1).to(torch.bfloat16)
x = (torch.randn(T, k_in, device=DEV) * 0.1).to(torch.bfloat16)
g = SlabGroup("gu", 0, n_out, k_in, 3, ("gu",))
bufs, bases = [], []
for o in range(world):
    a, b = g.row_range(o, world)
    t = torch.zeros(max(b - a, 1), k_in, device=DEV)
    bufs.append(t)
    bases.append(t.data_ptr() if b > a else None)
prev = torch.randn(n_out, k_in, device=DEV)
local = torch.zeros(n_out, k_in, device=DEV)
rs = K.make_rs_args(bases, g.rot)
rs_prev = K.make_rs_args(bases, g.rot, prev_ptr=prev.data_ptr())
gemm_rs = lambda: K.gemm(dy, x, local, a_mn=True, b_mn=True, epilogue=K.EPI_F32_ATOMIC, tile_n=256, M=n_out, N=k_in, K=T, rs=rs)
gemm_rs_prev = lambda: K.gemm(dy, x, local, a_mn=True, b_mn=True, epilogue=K.EPI_F32_ATOMIC, tile_n=256, M=n_out, N=k_in, K=T,
                              rs=rs_prev)
gemm_local = lambda: K.gemm(dy, x, local, a_mn=True, b_mn=True, epilogue=K.EPI_F32_ATOMIC, tile_n=256, M=n_out, N=k_in, K=T)

V, h = 50257, 768
ge = SlabGroup("emb", 0, V, h, 0, ("emb",))
ebufs, ebases = [], []
for o in range(world):
    a, b = ge.row_range(o, world)
    t = torch.zeros(b - a, h, device=DEV)
    ebufs.append(t)
    ebases.append(t.data_ptr())
ids = torch.randint(0, V, (T,), device=DEV)
dx = torch.randn(T, h, device=DEV)
ers = K.make_rs_args(ebases, 0)
embed = lambda: K.embed_bwd_rs(ids, dx, ers, V)

n = 20_253_696 // 128 * 128          # parameters one rank owns at 8 GPUs (small preset)
master, grad = torch.randn(n, device=DEV), torch.randn(n, device=DEV)
m, v = torch.zeros(n, device=DEV), torch.zeros(n, device=DEV)
shadows = [torch.zeros(n, dtype=torch.bfloat16, device=DEV) for _ in range(world)]
ptrs = K.ptr_array([t.data_ptr() for t in shadows])
decay = torch.ones(n // 128, dtype=torch.uint8, device=DEV)
scale = torch.ones(1, device=DEV)
table, nch = K.make_chunk_table([(0, 0, 0, 0, n, -1)], torch.device(DEV))
out = torch.zeros(1, device=DEV)
sumsq = lambda: K.sumsq_chunks(grad, table, nch, out, 1.0)
adamw = lambda: K.adamw_chunks(master, grad, m, v, ptrs, 0, world, decay, table, nch, 1, 1e-3, 0.9, 0.95, 1e-8, 0.1, scale,
                               zero_grad=True)
small = 128 * 150
srcs = [torch.randn(small, device=DEV) for _ in range(world)]
ptab, pn = K.make_chunk_table([(0, 0, 0, 0, small, -1)], torch.device(DEV))
dst = torch.zeros(small, device=DEV)
pull = lambda: K.pull_reduce_chunks(K.ptr_array([s.data_ptr() for s in srcs]), dst, ptab, pn, 0, world, zero_remote=False)

fns = [("wgrad GEMM -> reduce-scatter (8 owners)", gemm_rs, 2.0 * T * n_out * k_in),
       ("wgrad GEMM -> reduce-scatter + carried accumulator", gemm_rs_prev, 2.0 * T * n_out * k_in),
       ("wgrad GEMM, local reduce-add", gemm_local, 2.0 * T * n_out * k_in), ("embed_bwd_rs", embed, 0),
       ("sumsq_chunks (20.3 M)", sumsq, 0), ("adamw_chunks (20.3 M, 8 stores)", adamw, 0), ("pull_reduce_chunks", pull, 0)]
for _, f, _ in fns:
    f()
torch.cuda.synchronize()
if os.environ.get("FUSED_TIME", "0") == "1":
    for name, f, fl in fns:
        t = timeit(f)
        print(f"{name:52s} {t * 1e3:8.1f} us" + (f"  {fl / t / 1e9:6.0f} TFLOP/s" if fl else ""))
print("ok")
