"""Bandwidth-bound kernels vs PyTorch fp32 oracles of the same op."""
import math

import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("rows,hidden", [(64, 768), (1000, 1024), (333, 1600), (130, 256), (77, 512), (4100, 768)])
def test_rmsnorm_fwd_bwd(rows, hidden):
    from distributed_llm_trainer_b200.ops import kernels as k
    from distributed_llm_trainer_b200.ops import reference as R

    torch.manual_seed(0)
    x = torch.randn(rows, hidden, device="cuda") * 2
    w = (1 + 0.1 * torch.randn(hidden, device="cuda")).to(torch.bfloat16)
    y = torch.empty(rows, hidden, device="cuda", dtype=torch.bfloat16)
    rstd = torch.empty(rows, device="cuda")
    k.rmsnorm_fwd(x, w, y, rstd, 1e-6)
    ref, r = R.rmsnorm(x, w)
    torch.cuda.synchronize()
    assert (y.float() - ref).abs().max().item() < 0.02 * ref.abs().max().item()
    assert torch.allclose(rstd, r, rtol=1e-4, atol=1e-6)
    # backward
    dy = torch.randn(rows, hidden, device="cuda").to(torch.bfloat16)
    xx = x.clone().requires_grad_(True)
    ww = w.float().clone().requires_grad_(True)
    (R.rmsnorm(xx, ww)[0] * dy.float()).sum().backward()
    dx = torch.ones(rows, hidden, device="cuda")          # accumulates on top of existing content
    dw = torch.zeros(hidden, device="cuda")
    k.rmsnorm_bwd(dy, x, w, rstd, dx, dw)
    torch.cuda.synchronize()
    assert (dx - 1 - xx.grad).abs().max().item() < 2e-3 * max(1.0, xx.grad.abs().max().item())
    assert (dw - ww.grad).abs().max().item() < 2e-3 * max(1.0, ww.grad.abs().max().item())
    # fused variant: also emits dropout'(updated dx) in bf16, identical to the two-pass dropout_bwd_cast
    rng = torch.tensor([11, 4], dtype=torch.int64, device="cuda")
    for p in (0.0, 0.1):
        dx2 = torch.ones(rows, hidden, device="cuda")
        dw2 = torch.zeros(hidden, device="cuda")
        dyn = torch.full((rows, hidden), float("nan"), device="cuda", dtype=torch.bfloat16)
        k.rmsnorm_bwd(dy, x, w, rstd, dx2, dw2, dy_next=dyn, p_next=p, rng_state=rng, rng_stream=6)
        two_pass = torch.empty_like(dyn)
        k.dropout_bwd_cast(dx2, two_pass, p, rng, 6)
        torch.cuda.synchronize()
        assert torch.equal(dx2, dx) and torch.allclose(dw2, dw, rtol=1e-4, atol=1e-4)
        assert torch.equal(dyn, two_pass), f"fused dropout-cast differs from dropout_bwd_cast at p={p}"


def test_embedding_fwd_bwd():
    from distributed_llm_trainer_b200.ops import kernels as k

    torch.manual_seed(1)
    V, h, rows = 1000, 256, 4096
    w = torch.randn(V, h, device="cuda").to(torch.bfloat16)
    ids = torch.randint(0, V, (rows,), device="cuda")
    out = torch.empty(rows, h, device="cuda")
    k.embed_fwd(ids, w, out)
    torch.cuda.synchronize()
    assert torch.equal(out, w.float()[ids])
    dx = torch.randn(rows, h, device="cuda")
    dw = torch.zeros(V, h, device="cuda")
    k.embed_bwd(ids, dx, dw)
    ref = torch.zeros(V, h, device="cuda").index_add_(0, ids, dx)
    torch.cuda.synchronize()
    assert (dw - ref).abs().max().item() < 1e-3


@pytest.mark.parametrize("vocab,rows", [(50257, 96), (1000, 96), (50257, 1000), (1000, 777)])
def test_cross_entropy_fwd_bwd(vocab, rows):
    """rows < 2 x SMs: one CTA per row; more rows: the persistent kernel whose producer warp moves rows by bulk copy."""
    from distributed_llm_trainer_b200.ops import kernels as k

    torch.manual_seed(2)
    ld = (vocab + 127) // 128 * 128
    logits = torch.zeros(rows, ld, device="cuda", dtype=torch.bfloat16)
    logits[:, :vocab] = (torch.randn(rows, vocab, device="cuda") * 3).to(torch.bfloat16)
    logits[:, vocab:] = 77.0  # garbage in the padding must be ignored
    labels = torch.randint(0, vocab, (rows,), device="cuda")
    labels[5] = -100
    labels[rows - 1] = -100
    ref_in = logits[:, :vocab].float().clone().requires_grad_(True)
    ref_loss = torch.nn.functional.cross_entropy(ref_in, labels, ignore_index=-100, reduction="sum")
    ref_loss.backward()
    loss = torch.zeros(1, device="cuda")
    k.ce_fwd_bwd(logits, labels, loss, vocab, 1.0, 1.0)
    torch.cuda.synchronize()
    assert abs(loss.item() - ref_loss.item()) < 1e-3 * abs(ref_loss.item())
    assert (logits[:, :vocab].float() - ref_in.grad).abs().max().item() < 1e-2
    assert logits[:, vocab:].abs().max().item() == 0
    assert logits[5].abs().max().item() == 0 and logits[rows - 1].abs().max().item() == 0


def test_adamw_matches_torch():
    from distributed_llm_trainer_b200.ops import kernels as k

    torch.manual_seed(3)
    n = 128 * 37
    p0 = torch.randn(n, device="cuda")
    decay = (torch.arange(n // 128, device="cuda") % 2).to(torch.uint8)
    master, m, v = p0.clone(), torch.zeros(n, device="cuda"), torch.zeros(n, device="cuda")
    shadow = torch.zeros(n, device="cuda", dtype=torch.bfloat16)
    tp_d = torch.nn.Parameter(p0.clone())
    mask = decay.bool().repeat_interleave(128)
    opt_d = torch.optim.AdamW([tp_d], lr=1e-2, betas=(0.9, 0.95), eps=1e-8, weight_decay=0.1)
    tp_n = torch.nn.Parameter(p0.clone())
    opt_n = torch.optim.AdamW([tp_n], lr=1e-2, betas=(0.9, 0.95), eps=1e-8, weight_decay=0.0)
    scale = torch.tensor([0.5], device="cuda")
    for step in range(1, 4):
        g = torch.randn(n, device="cuda")
        tp_d.grad = g * 0.5
        tp_n.grad = g * 0.5
        opt_d.step()
        opt_n.step()
        k.adamw_step(master, g, m, v, shadow, decay, 0, n, step, 1e-2, 0.9, 0.95, 1e-8, 0.1, scale)
    torch.cuda.synchronize()
    ref = torch.where(mask, tp_d.data, tp_n.data)
    assert (master - ref).abs().max().item() < 1e-5
    assert (shadow.float() - ref).abs().max().item() < 0.01 * ref.abs().max().item()
    # sub-range update leaves the rest untouched
    before = master.clone()
    k.adamw_step(master, g, m[:256], v[:256], shadow, decay, 128, 384, 4, 1e-2, 0.9, 0.95, 1e-8, 0.1, scale)
    torch.cuda.synchronize()
    assert torch.equal(master[:128], before[:128]) and torch.equal(master[384:], before[384:])
    assert not torch.equal(master[128:384], before[128:384])


def test_sumsq_and_clip():
    from distributed_llm_trainer_b200.ops import kernels as k

    x = torch.randn(128 * 1000, device="cuda")
    out = torch.zeros(1, device="cuda")
    k.sumsq(x, 0, x.numel(), out)
    scale = torch.zeros(1, device="cuda")
    norm = torch.zeros(1, device="cuda")
    k.clip_coef(out, 1.0, 0.25, scale, norm)
    torch.cuda.synchronize()
    ref_norm = x.norm().item() * 0.25
    assert abs(norm.item() - ref_norm) < 1e-3 * ref_norm
    assert abs(scale.item() - 0.25 * min(1.0, 1.0 / (ref_norm + 1e-6))) < 1e-6
    rng = torch.tensor([5, 10], dtype=torch.int64, device="cuda")
    k.rng_advance(rng, 3)
    torch.cuda.synchronize()
    assert rng.tolist() == [5, 13]
