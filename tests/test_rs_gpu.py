"""GEMM -> reduce-scatter and the chunk-table optimizer kernels on ONE GPU: the "ranks" are separate local
buffers, so the owner-directed epilogue, the embedding scatter and the optimizer side are checked against
plain fp32 PyTorch without needing a second device (the multi-GPU runs in test_multi_gpu.py exercise the same
entry points over NVLink)."""
import pytest
import torch

from distributed_llm_trainer_b200.ops import kernels as K
from distributed_llm_trainer_b200.parallel.ownership import SLAB, SlabGroup

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _owner_buffers(rows, cols, world, rot):
    g = SlabGroup("t", 0, rows, cols, rot % world, ("t",))
    bufs, bases = [], []
    for o in range(world):
        a, b = g.row_range(o, world)
        # one guard row of NaN-free sentinel on both sides: a slab written through the wrong map shows up
        t = torch.full((b - a + 2, cols), 7.0, dtype=torch.float32, device=DEV)
        t[1:-1].zero_()
        bufs.append(t)
        bases.append(t[1:].data_ptr() if b > a else None)
    return g, bufs, bases


def _assemble(g, bufs, world):
    full = torch.zeros(g.rows, g.cols, device=DEV)
    for o in range(world):
        a, b = g.row_range(o, world)
        if b > a:
            full[a:b] = bufs[o][1:-1]
        assert torch.all(bufs[o][0] == 7.0) and torch.all(bufs[o][-1] == 7.0), f"owner {o}: guard rows overwritten"
    return full


@pytest.mark.parametrize("world", [1, 2, 4, 8])
@pytest.mark.parametrize("shape", [(768, 768, 1024), (2304, 768, 512), (1000, 256, 384), (128, 128, 256),
                                   (640, 1600, 256)])
def test_gemm_rs_matches_fp32(world, shape):
    n_out, k_in, T = shape
    torch.manual_seed(world * 131 + n_out)
    dy = (torch.randn(T, n_out, device=DEV) * 0.5).to(torch.bfloat16)
    x = (torch.randn(T, k_in, device=DEV) * 0.5).to(torch.bfloat16)
    ref = dy.float().t() @ x.float()
    for rot in (0, 3):
        g, bufs, bases = _owner_buffers(n_out, k_in, world, rot)
        rs = K.make_rs_args(bases, g.rot)
        layout = torch.full((n_out, k_in), 3.0, dtype=torch.float32, device=DEV)   # describes ld only: never written
        bn = 256 if k_in % 256 == 0 else 128
        for rep in (1, 2):          # second launch accumulates on top of the first
            K.gemm(dy, x, layout, a_mn=True, b_mn=True, epilogue=K.EPI_F32_ATOMIC, split_k=2 if rep == 2 else 1,
                   tile_n=bn, M=n_out, N=k_in, K=T, rs=rs)
            torch.cuda.synchronize()
            got = _assemble(g, bufs, world)
            err = (got - rep * ref).norm() / (rep * ref).norm()
            assert err < 2e-3, f"world {world} rot {rot} rep {rep}: rel L2 {err:.3e}"
        assert torch.all(layout == 3.0)


@pytest.mark.parametrize("world", [1, 4])
@pytest.mark.parametrize("shape,split", [((768, 768, 1024), 1), ((768, 768, 2048), 4), ((1000, 256, 384), 2),
                                         ((2304, 768, 512), 1)])
def test_gemm_rs_carries_local_accumulation(world, shape, split):
    """owner += A*B + prev: the locally accumulated gradient rides in the same reduce-add (first K split only)."""
    n_out, k_in, T = shape
    torch.manual_seed(7)
    dy = (torch.randn(T, n_out, device=DEV) * 0.5).to(torch.bfloat16)
    x = (torch.randn(T, k_in, device=DEV) * 0.5).to(torch.bfloat16)
    prev = torch.randn(n_out, k_in, device=DEV) * 30.0
    keep = prev.clone()
    ref = dy.float().t() @ x.float() + prev
    g, bufs, bases = _owner_buffers(n_out, k_in, world, 2)
    for o in range(world):          # owners already hold other ranks' contributions
        bufs[o][1:-1].fill_(1.5)
    rs = K.make_rs_args(bases, g.rot, prev_ptr=prev.data_ptr())
    K.gemm(dy, x, prev, a_mn=True, b_mn=True, epilogue=K.EPI_F32_ATOMIC, split_k=split,
           tile_n=256 if k_in % 256 == 0 else 128, M=n_out, N=k_in, K=T, rs=rs)
    torch.cuda.synchronize()
    got = _assemble(g, bufs, world)
    err = (got - (ref + 1.5)).norm() / ref.norm()
    assert err < 2e-3, f"rel L2 {err:.3e}"
    assert torch.equal(prev, keep), "the local accumulator is read, never written"


def test_slab_rule_matches_native():
    import ctypes as C

    from distributed_llm_trainer_b200.ops import _lib

    lib = _lib.lib()
    for rows in (128, 640, 768, 1000, 2304, 50257):
        for world in (1, 2, 3, 4, 8):
            for rot in (0, 1, 5):
                g = SlabGroup("t", 0, rows, 64, rot % world, ("t",))
                for s in range(g.slabs):
                    assert lib.b200_slab_owner(s, g.slabs, world, rot % world) == g.owner_of_slab(s, world)
                seen = []
                for o in range(world):
                    a, b = C.c_int(), C.c_int()
                    assert lib.b200_owner_slab_range(o, g.slabs, world, rot % world, C.byref(a), C.byref(b)) == 0
                    assert (a.value, b.value) == g.slab_range(o, world)
                    seen += list(range(a.value, b.value))
                    assert all(g.owner_of_slab(s, world) == o for s in range(a.value, b.value))
                assert sorted(seen) == list(range(g.slabs))


@pytest.mark.parametrize("world", [1, 2, 8])
def test_embed_bwd_rs(world):
    vocab, hidden, rows = 1000, 256, 4096
    torch.manual_seed(3)
    ids = torch.randint(0, vocab, (rows,), device=DEV)
    ids[:7] = torch.tensor([-100, vocab - 1, 0, 127, 128, 999, 511], device=DEV)
    dx = torch.randn(rows, hidden, device=DEV)
    ref = torch.zeros(vocab, hidden, device=DEV)
    ok = ids >= 0
    ref.index_add_(0, ids[ok], dx[ok])
    g, bufs, bases = _owner_buffers(vocab, hidden, world, 1)
    K.embed_bwd_rs(ids, dx, K.make_rs_args(bases, g.rot), vocab)
    torch.cuda.synchronize()
    got = _assemble(g, bufs, world)
    assert (got - ref).abs().max() < 1e-4 * ref.abs().max()


def test_chunk_kernels_vs_torch():
    """sumsq / AdamW / pull-reduce over a ragged list of runs, against torch.optim.AdamW."""
    torch.manual_seed(5)
    world, rank = 4, 1
    total = 128 * 400
    runs_flat = [(0, 4096 + 128), (128 * 50, 128 * 50 + 9000 // 4 * 4), (128 * 300, 128 * 400)]
    master = torch.randn(total, device=DEV)
    grad = torch.randn(total, device=DEV)
    decay = (torch.arange(total // 128, device=DEV) % 3 != 0).to(torch.uint8)
    n_state = sum(b - a for a, b in runs_flat)
    m = torch.zeros(n_state, device=DEV)
    v = torch.zeros(n_state, device=DEV)
    shadows = [torch.zeros(total, dtype=torch.bfloat16, device=DEV) for _ in range(world)]
    ptrs = K.ptr_array([t.data_ptr() for t in shadows])
    runs, off = [], 0
    for a, b in runs_flat:
        runs.append((a, a, off, a, b - a, -1))
        off += b - a
    table, n = K.make_chunk_table(runs, torch.device(DEV))
    scale = torch.tensor([0.37], device=DEV)
    out = torch.zeros(1, device=DEV)
    K.sumsq_chunks(grad, table, n, out, 0.25)
    want = sum(grad[a:b].double().pow(2).sum() for a, b in runs_flat) * 0.25
    assert abs(out.item() - want.item()) < 1e-4 * want.item()
    g0 = grad.clone()
    for step in (1, 2):
        grad.copy_(g0)
        K.adamw_chunks(master, grad, m, v, ptrs, rank, world, decay, table, n, step, 1e-2, 0.9, 0.95, 1e-8, 0.1, scale,
                       zero_grad=True)
        torch.cuda.synchronize()
    # torch reference: decayed and undecayed elements as two optimizers over masks
    mask_owned = torch.zeros(total, dtype=torch.bool, device=DEV)
    for a, b in runs_flat:
        mask_owned[a:b] = True
    dec_el = decay.bool().repeat_interleave(128)
    # (master was updated in place: rebuild the expectation from the initial values)
    torch.manual_seed(5)
    init = torch.randn(total, device=DEV)
    for wd, sel in ((0.1, mask_owned & dec_el), (0.0, mask_owned & ~dec_el)):
        p = init[sel].clone().requires_grad_(True)
        o = torch.optim.AdamW([p], lr=1e-2, betas=(0.9, 0.95), eps=1e-8, weight_decay=wd)
        for _ in (1, 2):
            p.grad = g0[sel] * 0.37
            o.step()
        assert torch.allclose(master[sel], p.detach(), rtol=2e-5, atol=2e-6)
    assert torch.equal(master[~mask_owned], init[~mask_owned]), "elements outside the runs were touched"
    assert torch.all(grad[mask_owned] == 0) and torch.equal(grad[~mask_owned], g0[~mask_owned])
    for t in shadows:
        assert torch.equal(t[mask_owned], master[mask_owned].to(torch.bfloat16))
        assert torch.all(t[~mask_owned] == 0)
    # dst >= 0: only that rank's buffer, at the shard offset
    for t in shadows:
        t.zero_()
    runs2 = [(a, a, o_, 64 + o_, n_, 2) for (a, _, o_, _, n_, _) in runs]
    table2, n2 = K.make_chunk_table(runs2, torch.device(DEV))
    grad.copy_(g0)
    K.adamw_chunks(master, grad, m, v, ptrs, rank, world, decay, table2, n2, 3, 1e-2, 0.9, 0.95, 1e-8, 0.1, scale)
    torch.cuda.synchronize()
    cat = torch.cat([master[a:b] for a, b in runs_flat]).to(torch.bfloat16)
    assert torch.equal(shadows[2][64: 64 + n_state], cat)
    assert all(torch.all(shadows[r] == 0) for r in (0, 1, 3))
    # pull-reduce: owner sums the peers' copies and clears them
    srcs = [torch.randn(total, device=DEV) for _ in range(world)]
    keep = [s.clone() for s in srcs]
    K.pull_reduce_chunks(K.ptr_array([s.data_ptr() for s in srcs]), srcs[rank], table, n, rank, world, zero_remote=True)
    torch.cuda.synchronize()
    want = sum(keep)
    assert torch.allclose(srcs[rank][mask_owned], want[mask_owned], rtol=1e-5, atol=1e-5)
    assert torch.equal(srcs[rank][~mask_owned], keep[rank][~mask_owned])
    for r in range(world):
        if r != rank:
            assert torch.all(srcs[r][mask_owned] == 0) and torch.equal(srcs[r][~mask_owned], keep[r][~mask_owned])
