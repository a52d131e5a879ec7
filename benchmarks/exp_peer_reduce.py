"""Feasibility experiment for the GEMM -> reduce-scatter fusion: a wgrad GEMM whose TMA reduce-add epilogue
targets the OWNER's gradient buffer over NVLink (a peer-mapped address) instead of local memory.

    torchrun --nproc-per-node 2 --master-addr 127.0.0.1 benchmarks/exp_peer_reduce.py

Every rank computes dW_r = dY_r^T X_r for the gate|up weight shape and reduce-adds the top half of the rows into
rank 0's buffer and the bottom half into rank 1's (row-aligned ownership), so after a barrier each owner holds the
cross-rank sum of its rows.  Checks the result against torch and times local vs remote accumulation."""
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributed_llm_trainer_b200.ops import kernels as K  # noqa: E402
from distributed_llm_trainer_b200.parallel.comm import SignalPads, SymmetricBuffer, wrap_pointer  # noqa: E402


def timed(fn, pads, iters=20):
    ts = []
    for _ in range(iters):
        pads.barrier()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    ts.sort()
    return ts[len(ts) // 2]


def main():
    local = int(os.environ.get("LOCAL_RANK", 0))
    dev = torch.device(f"cuda:{local}")
    torch.cuda.set_device(dev)
    dist.init_process_group("nccl", device_id=dev)
    rank, world = dist.get_rank(), dist.get_world_size()
    assert world == 2, "experiment is written for two ranks"
    pads = SignalPads(dev)
    T, n_out, k_in = 8192, 6144, 768
    torch.manual_seed(10 + rank)
    dy = (torch.randn(T, n_out, device=dev) * 0.1).to(torch.bfloat16)
    x = (torch.randn(T, k_in, device=dev) * 0.1).to(torch.bfloat16)
    half = n_out // world
    gbuf = SymmetricBuffer(half * k_in, torch.float32, dev)      # this rank's owned rows of the gradient
    views = [wrap_pointer(p, half * k_in, torch.float32, dev).view(half, k_in) for p in gbuf.raw_ptrs]
    local_full = torch.zeros(n_out, k_in, device=dev)

    def remote():   # rows [r*half, (r+1)*half) of dW go to rank r's buffer (TMA reduce-add over NVLink for r != rank)
        for r in range(world):
            K.gemm(dy[:, r * half:(r + 1) * half], x, views[r], a_mn=True, b_mn=True, epilogue=K.EPI_F32_ATOMIC,
                   M=half, N=k_in, K=T)

    def local_only():
        K.gemm(dy, x, local_full, a_mn=True, b_mn=True, epilogue=K.EPI_F32_ATOMIC, M=n_out, N=k_in, K=T)

    def local_halves():   # same two launches as remote(), both into local memory: isolates the NVLink cost
        for r in range(world):
            K.gemm(dy[:, r * half:(r + 1) * half], x, local_full[r * half:(r + 1) * half], a_mn=True, b_mn=True,
                   epilogue=K.EPI_F32_ATOMIC, M=half, N=k_in, K=T)

    # correctness
    gbuf.local.zero_()
    pads.barrier()
    remote()
    pads.barrier()
    torch.cuda.synchronize()
    total = dy.float().t() @ x.float()          # this rank's contribution to every row ...
    dist.all_reduce(total)                        # ... summed over ranks
    both = total[rank * half:(rank + 1) * half]
    err = (views[rank] - both).abs().max().item()
    scale = both.abs().max().item()
    ok = err < 2e-2 * scale
    t_remote, t_local, t_halves = timed(remote, pads), timed(local_only, pads), timed(local_halves, pads)
    out = torch.tensor([t_remote, t_local, t_halves], device=dev)
    dist.all_reduce(out, op=dist.ReduceOp.MAX)
    if rank == 0:
        mb = half * k_in * 4 / 1e6
        print(f"peer TMA reduce-add: max err {err:.3e} (scale {scale:.3e}) -> {'OK' if ok else 'MISMATCH'}")
        print(f"wgrad {n_out}x{k_in} K={T}: one local launch {out[1].item() * 1e3:.1f} us; two half launches, both local "
              f"{out[2].item() * 1e3:.1f} us; two half launches, owner-directed ({mb:.1f} MB reduce-added into the peer "
              f"over NVLink) {out[0].item() * 1e3:.1f} us")
    assert ok
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
