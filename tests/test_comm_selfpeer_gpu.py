"""comm.cu on ONE GPU: the "peers" of every entry point are separate local buffers (or, for the signal pads, a pad
whose peer slots are pre-filled as if the peers had already arrived), so the driver's single-GPU `pytest -m gpu`
covers the kernels that otherwise only run in the multi-GPU tests.  Oracles are plain fp32 torch."""
import pytest
import torch

from distributed_llm_trainer_b200.ops import _lib
from distributed_llm_trainer_b200.ops import kernels as K

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _ptrs(ts):
    return K.ptr_array([t.data_ptr() for t in ts])


@pytest.mark.parametrize("world", [2, 4, 8, 3])
def test_reduce_scatter_f32_and_sumsq(world):
    torch.manual_seed(world)
    n = 128 * 96
    srcs = [torch.randn(n, device=DEV) for _ in range(world)]
    want = torch.stack(srcs).sum(0) / world
    for rank in range(world):
        lo, hi = rank * (n // world) // 4 * 4, (rank + 1) * (n // world) // 4 * 4
        dst = torch.zeros(n, device=DEV)
        sumsq = torch.zeros(1, device=DEV)
        K.reduce_scatter_f32(_ptrs(srcs), dst, lo, hi, 1.0 / world, sumsq, rank, world)
        torch.cuda.synchronize()
        assert torch.allclose(dst[lo:hi], want[lo:hi], rtol=1e-6, atol=1e-6)
        assert torch.all(dst[:lo] == 0) and torch.all(dst[hi:] == 0)
        assert abs(sumsq.item() - want[lo:hi].double().pow(2).sum().item()) < 1e-4 * sumsq.item()


@pytest.mark.parametrize("world", [2, 8])
def test_accumulating_reduce_scatters(world):
    torch.manual_seed(1)
    n = 4096
    full = [torch.randn(4 * n, device=DEV) for _ in range(world)]
    acc = torch.ones(n, device=DEV)
    K.reduce_scatter_f32_accum(_ptrs(full), acc, 2 * n, n, 0.5, 1, world)
    torch.cuda.synchronize()
    want = 1 + 0.5 * torch.stack(full).sum(0)[2 * n: 3 * n]
    assert torch.allclose(acc, want, rtol=1e-5, atol=1e-5)
    fullb = [t.to(torch.bfloat16) for t in full]
    acc = torch.ones(n, device=DEV)
    K.reduce_scatter_bf16_accum(_ptrs(fullb), acc, 2 * n, n, 0.5, 0, world)
    torch.cuda.synchronize()
    want = 1 + 0.5 * torch.stack([t.float() for t in fullb]).sum(0)[2 * n: 3 * n]
    assert torch.allclose(acc, want, rtol=1e-5, atol=1e-5)


@pytest.mark.parametrize("world", [2, 4, 8])
def test_all_gather_pull_with_exclusion(world):
    shard = 2048
    shards = [(torch.arange(shard, device=DEV) % 251 + 7 * r).to(torch.bfloat16) for r in range(world)]
    dst = torch.full((shard * world,), -1.0, dtype=torch.bfloat16, device=DEV)
    K.all_gather_bf16(_ptrs(shards), dst, shard, 1, world, exclude=(shard + 64, shard + 512))
    torch.cuda.synchronize()
    want = torch.cat(shards)
    keep = torch.ones_like(want, dtype=torch.bool)
    keep[shard + 64: shard + 512] = False
    assert torch.equal(dst[keep], want[keep]) and torch.all(dst[~keep] == -1.0)


@pytest.mark.parametrize("world", [2, 8])
def test_adamw_bcast_matches_torch_adamw(world):
    torch.manual_seed(2)
    n, lo, hi, rank = 128 * 64, 128 * 8, 128 * 40, 1
    master = torch.randn(n, device=DEV)
    init = master.clone()
    grad = torch.randn(n, device=DEV)
    m = torch.zeros(hi - lo, device=DEV)
    v = torch.zeros(hi - lo, device=DEV)
    shadows = [torch.zeros(n, dtype=torch.bfloat16, device=DEV) for _ in range(world)]
    decay = torch.ones(n // 128, dtype=torch.uint8, device=DEV)
    scale = torch.tensor([0.5], device=DEV)
    p = init[lo:hi].clone().requires_grad_(True)
    opt = torch.optim.AdamW([p], lr=1e-2, betas=(0.9, 0.95), eps=1e-8, weight_decay=0.1)
    for step in (1, 2, 3):
        K.adamw_bcast(master, grad, m, v, _ptrs(shadows), rank, world, decay, lo, hi, step, 1e-2, 0.9, 0.95, 1e-8, 0.1, scale)
        p.grad = grad[lo:hi] * 0.5
        opt.step()
    torch.cuda.synchronize()
    assert torch.allclose(master[lo:hi], p.detach(), rtol=2e-5, atol=2e-6)
    assert torch.equal(master[:lo], init[:lo]) and torch.equal(master[hi:], init[hi:])
    for t in shadows:
        assert torch.equal(t[lo:hi], master[lo:hi].to(torch.bfloat16)) and torch.all(t[:lo] == 0) and torch.all(t[hi:] == 0)


def _pad(world, epoch_before, arrived):
    """A signal pad for rank 0 of `world` whose peers have already published epoch `arrived`."""
    n = _lib.lib().b200_comm_pad_bytes() // 4
    pad = torch.zeros(n, dtype=torch.int32, device=DEV)
    pad[1:world] = arrived          # row 0: slot [t] = epoch published by rank t
    others = [torch.zeros(n, dtype=torch.int32, device=DEV) for _ in range(world - 1)]
    counter = torch.tensor([epoch_before], dtype=torch.int32, device=DEV)
    return pad, others, counter


@pytest.mark.parametrize("world", [2, 8])
@pytest.mark.parametrize("epoch_before", [0, 41, 2 ** 31 - 2, -2])   # -2 = 0xFFFFFFFE: the next epochs wrap through 0
def test_barrier_epochs_and_wraparound(world, epoch_before):
    """The barrier of rank 0 publishes epoch+1 into every peer's pad and returns once all peers show an epoch that
    is not behind (signed distance): checked for small epochs, around 2^31 and across the 2^32 wrap."""
    e = epoch_before
    nxt = ((e + 1 + 2 ** 31) % 2 ** 32) - 2 ** 31          # as int32
    pad, others, counter = _pad(world, e, nxt)
    K.comm_barrier(_ptrs([pad] + others), 0, world, counter)
    torch.cuda.synchronize()                                # returns (peers "arrived"); a bug would trap after 20 s
    assert counter.item() == nxt
    assert pad[0].item() == nxt and all(o[0].item() == nxt for o in others)
    # peers AHEAD of this rank (they already entered the next barrier) must not block it either
    ahead = ((e + 3 + 2 ** 31) % 2 ** 32) - 2 ** 31
    pad[1:world] = ahead
    K.comm_barrier(_ptrs([pad] + others), 0, world, counter)
    torch.cuda.synchronize()
    assert counter.item() == ((e + 2 + 2 ** 31) % 2 ** 32) - 2 ** 31


@pytest.mark.parametrize("world", [2, 8])
def test_allreduce_scalars_self_peer(world):
    n = _lib.lib().b200_comm_pad_bytes() // 4
    pads = [torch.zeros(n, dtype=torch.int32, device=DEV) for _ in range(world)]
    counter = torch.zeros(1, dtype=torch.int32, device=DEV)
    # epoch 1: parity 1 half of the exchange area; peers' contributions pre-written into rank 0's pad
    xchg = pads[0].view(torch.float32)[64:]                  # after kMaxRanks*4 barrier slots (16*4 uint32)
    for r in range(1, world):
        xchg[16 * 8 + r * 8: 16 * 8 + r * 8 + 2] = torch.tensor([float(r + 1), 2.0], device=DEV)
    pads[0][1:world] = 1
    inp = torch.tensor([1.0, 2.0, 0, 0, 0, 0, 0, 0], device=DEV)
    out = torch.zeros(8, device=DEV)
    K.allreduce_scalars(_ptrs(pads), 0, world, counter, inp, out, 2)
    torch.cuda.synchronize()
    assert out[0].item() == world * (world + 1) / 2 and out[1].item() == 2.0 * world
    for r in range(1, world):                                # rank 0's values landed in every peer's exchange slot 0
        assert pads[r].view(torch.float32)[64 + 16 * 8: 64 + 16 * 8 + 2].tolist() == [1.0, 2.0]
