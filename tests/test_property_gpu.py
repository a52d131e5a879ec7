"""Property tests (hypothesis) of the sm_100a kernels over random shapes, against fp32 PyTorch expressions
of the same math.  Shapes are drawn around the tile boundaries (128 / 256 / 384 columns, 64-deep K blocks,
ragged M), which is where a tcgen05 / TMA kernel breaks if its clipping or descriptor arithmetic is off."""
import math

import pytest
import torch

hypothesis = pytest.importorskip("hypothesis")
from hypothesis import given, settings, HealthCheck  # noqa: E402
from hypothesis import strategies as st  # noqa: E402

pytestmark = pytest.mark.gpu

import os  # noqa: E402

# B200_PROP_SCALE=5 runs five times as many examples (bug hunts); the default keeps the suite at seconds
_SCALE = int(os.environ.get("B200_PROP_SCALE", "1"))
_SETTINGS = dict(max_examples=20 * _SCALE, deadline=None, derandomize=_SCALE == 1,
                 suppress_health_check=[HealthCheck.too_slow, HealthCheck.function_scoped_fixture])


def _bf(shape, scale=1.0):
    return (torch.randn(*shape, device="cuda") * scale).to(torch.bfloat16)


@settings(**_SETTINGS)
@given(m=st.integers(1, 1100), n8=st.integers(1, 150), k8=st.integers(1, 100), a_mn=st.booleans(), b_mn=st.booleans(),
       pair=st.booleans())
def test_gemm_any_shape(m, n8, k8, a_mn, b_mn, pair):
    from distributed_llm_trainer_b200.ops import kernels as k

    n, kk = 8 * n8, 8 * k8
    if a_mn:
        m = (m + 7) // 8 * 8          # MN-major A: its leading dimension is M
    torch.manual_seed(m * 7 + n + kk)
    a = _bf((kk, m) if a_mn else (m, kk))
    b = _bf((kk, n) if b_mn else (n, kk))
    c = torch.full((m, n), float("nan"), device="cuda", dtype=torch.bfloat16)
    k.gemm(a, b, c, a_mn=a_mn, b_mn=b_mn, pair=pair)
    torch.cuda.synchronize()
    A = a.float().t() if a_mn else a.float()
    B = b.float() if b_mn else b.float().t()
    ref = A @ B
    err = (c.float() - ref).abs().max().item()
    tol = 0.02 * math.sqrt(kk) + 0.05 * ref.abs().max().item() * 2 ** -7
    assert err <= tol, f"{m}x{n}x{kk} a_mn={a_mn} b_mn={b_mn} pair={pair}: {err} > {tol}"


@settings(**_SETTINGS)
@given(m=st.integers(1, 700), n384=st.integers(1, 4), k8=st.integers(1, 60), b_mn=st.booleans())
def test_gemm_tile384_any_shape(m, n384, k8, b_mn):
    from distributed_llm_trainer_b200.ops import kernels as k

    n, kk = 384 * n384, 8 * k8
    torch.manual_seed(m + n + kk)
    a = _bf((m, kk))
    b = _bf((kk, n) if b_mn else (n, kk))
    c = torch.full((m, n), float("nan"), device="cuda", dtype=torch.bfloat16)
    k.gemm(a, b, c, b_mn=b_mn, pair=True, tile_n=384)
    torch.cuda.synchronize()
    ref = a.float() @ (b.float() if b_mn else b.float().t())
    err = (c.float() - ref).abs().max().item()
    tol = 0.02 * math.sqrt(kk) + 0.05 * ref.abs().max().item() * 2 ** -7
    assert err <= tol, f"tile384 {m}x{n}x{kk} b_mn={b_mn}: {err} > {tol}"


@settings(**_SETTINGS)
@given(rows=st.integers(1, 600), h8=st.integers(1, 256), p=st.sampled_from([0.0, 0.1, 0.5]))
def test_rmsnorm_any_shape(rows, h8, p):
    from distributed_llm_trainer_b200.ops import kernels as k
    from distributed_llm_trainer_b200.ops import reference as R

    hidden = 8 * h8
    torch.manual_seed(rows + hidden)
    x = torch.randn(rows, hidden, device="cuda")
    w = (torch.rand(hidden, device="cuda") + 0.5).to(torch.bfloat16)
    y = torch.empty(rows, hidden, device="cuda", dtype=torch.bfloat16)
    rstd = torch.empty(rows, device="cuda")
    k.rmsnorm_fwd(x, w, y, rstd, 1e-6)
    ref, _ = R.rmsnorm(x, w)
    torch.cuda.synchronize()
    assert (y.float() - ref).abs().max().item() <= 0.02 * max(1.0, ref.abs().max().item())
    dy = torch.randn(rows, hidden, device="cuda").to(torch.bfloat16)
    xx = x.clone().requires_grad_(True)
    ww = w.float().clone().requires_grad_(True)
    (R.rmsnorm(xx, ww)[0] * dy.float()).sum().backward()
    dx, dw = torch.zeros(rows, hidden, device="cuda"), torch.zeros(hidden, device="cuda")
    dyn = torch.empty(rows, hidden, device="cuda", dtype=torch.bfloat16)
    rng = torch.tensor([rows, hidden], dtype=torch.int64, device="cuda")
    k.rmsnorm_bwd(dy, x, w, rstd, dx, dw, dy_next=dyn, p_next=p, rng_state=rng, rng_stream=3)
    two_pass = torch.empty_like(dyn)
    k.dropout_bwd_cast(dx, two_pass, p, rng, 3)
    torch.cuda.synchronize()
    assert (dx - xx.grad).abs().max().item() <= 3e-3 * max(1.0, xx.grad.abs().max().item())
    assert (dw - ww.grad).abs().max().item() <= 3e-3 * max(1.0, ww.grad.abs().max().item())
    assert torch.equal(dyn, two_pass)


@settings(**_SETTINGS)
@given(rows=st.integers(1, 300), vocab=st.integers(2, 9000), pad=st.integers(0, 3))
def test_cross_entropy_any_shape(rows, vocab, pad):
    from distributed_llm_trainer_b200.ops import kernels as k

    ld = (vocab + 7) // 8 * 8 + 8 * pad
    torch.manual_seed(rows * 3 + vocab)
    logits = torch.randn(rows, ld, device="cuda").mul_(3).to(torch.bfloat16)
    labels = torch.randint(0, vocab, (rows,), device="cuda")
    labels[::7] = -100                                    # ignored positions
    ref_in = logits[:, :vocab].float().clone().requires_grad_(True)
    ref_loss = torch.nn.functional.cross_entropy(ref_in, labels, ignore_index=-100, reduction="sum")
    ref_loss.backward()
    loss = torch.zeros(1, device="cuda")
    k.ce_fwd_bwd(logits, labels, loss, vocab, 1.0, 1.0)
    torch.cuda.synchronize()
    assert abs(loss.item() - ref_loss.item()) <= 2e-3 * max(1.0, abs(ref_loss.item()))
    assert (logits[:, :vocab].float() - ref_in.grad).abs().max().item() <= 1e-2
    assert logits[:, vocab:].float().abs().max().item() == 0.0 if ld > vocab else True


@settings(**dict(_SETTINGS, max_examples=10 * _SCALE))
@given(b=st.integers(1, 3), s=st.integers(1, 520), h=st.integers(1, 3), p=st.sampled_from([0.0, 0.1]))
def test_attention_any_shape(b, s, h, p):
    from distributed_llm_trainer_b200.ops import kernels as k
    from distributed_llm_trainer_b200.ops import reference as R
    from .test_attention_gpu import _attn_mask, _p_eff

    torch.manual_seed(b * 100 + s + h)
    hd = h * 64
    qkv = _bf((b * s, 3 * hd), 0.8)
    rng = torch.tensor([5, 9], dtype=torch.int64, device="cuda")
    out = torch.zeros(b * s, hd, device="cuda", dtype=torch.bfloat16)
    lse = torch.zeros(b, h, s, device="cuda")
    k.attn_fwd(qkv, out, lse, b, s, h, p, rng, 4)
    torch.cuda.synchronize()
    q, kk, v = [x.clone().requires_grad_(True) for x in qkv.float().reshape(b, s, 3, h, 64).permute(2, 0, 3, 1, 4)]
    mask = _attn_mask(b, h, s, p, rng, 4) if p > 0 else None
    ref = R.causal_attention(q, kk, v, mask, _p_eff(p)).permute(0, 2, 1, 3).reshape(b * s, hd)
    assert (out.float() - ref).abs().max().item() < 0.03
    dout = _bf((b * s, hd), 0.5)
    ref.backward(dout.float())
    dqkv = torch.zeros(b * s, 3 * hd, device="cuda", dtype=torch.bfloat16)
    delta, dq_acc = torch.zeros(b, h, s, device="cuda"), torch.empty(b * s, hd, device="cuda")
    k.attn_bwd(qkv, out, dout, lse, dqkv, delta, dq_acc, b, s, h, p, rng, 4)
    torch.cuda.synchronize()
    got = dqkv.float().reshape(b, s, 3, h, 64).permute(2, 0, 3, 1, 4)
    for name, g, r_ in (("dq", got[0], q.grad), ("dk", got[1], kk.grad), ("dv", got[2], v.grad)):
        assert (g - r_).abs().max().item() < 0.04 * max(1.0, r_.abs().max().item()), name


@settings(**dict(_SETTINGS, max_examples=8 * _SCALE))
@given(heads=st.integers(1, 4), layers=st.integers(1, 3), inter128=st.integers(1, 6), vocab=st.integers(50, 1500),
       batch=st.integers(1, 3), seq=st.integers(2, 300), ckpt=st.booleans())
def test_engine_matches_torch_on_ragged_configs(heads, layers, inter128, vocab, batch, seq, ckpt):
    """Whole fused step (every kernel chained, 384- and 256-wide tiles, ragged M / S / vocab) against the fp32
    PyTorch model with the same weights: loss and every parameter gradient."""
    import copy

    from distributed_llm_trainer_b200.models import GPT, GPTConfig

    cfg = GPTConfig(vocab_size=vocab, hidden_size=64 * heads, num_layers=layers, num_heads=heads,
                    intermediate_size=128 * inter128, max_seq_len=320, dropout=0.0, attention_dropout=0.0,
                    gradient_checkpointing=ckpt)
    torch.manual_seed(heads * 1000 + seq)
    model = GPT(cfg)
    ref = copy.deepcopy(model).cuda().float()
    ref.backend = "torch"
    model = model.cuda()
    eng = model.engine()
    eng.use_graphs = False
    ids = torch.randint(0, vocab, (batch, seq), device="cuda")
    ref.train()
    _, ref_loss = ref(ids, labels=ids)
    ref_loss.backward()
    model.train()
    eng.store.zero_grad()
    loss = eng.forward_backward(ids, ids, 1.0)
    torch.cuda.synchronize()
    assert abs(loss.item() - ref_loss.item()) < 0.02 * abs(ref_loss.item()) + 1e-3, (loss.item(), ref_loss.item())
    bad = []
    for (n, p), (_, q) in zip(model.named_parameters(), ref.named_parameters()):
        g, r = p.grad.float(), q.grad.float()
        err = (g - r).abs().max().item() / (r.abs().max().item() + 1e-6)
        if err > 0.1:
            bad.append((n, round(err, 3)))
    assert not bad, f"cfg h={64 * heads} L={layers} I={128 * inter128} V={vocab} B={batch} S={seq}: {bad}"


@settings(**_SETTINGS)
@given(m=st.integers(1, 900), n8=st.integers(1, 160), k8=st.integers(1, 80), p=st.sampled_from([0.0, 0.1]),
       pair=st.booleans(), tile=st.sampled_from([0, 256, 384]))
def test_gemm_resid_dropout_any_shape(m, n8, k8, p, pair, tile):
    from distributed_llm_trainer_b200.ops import kernels as k

    n, kk = 8 * n8, 8 * k8
    if tile == 384:
        n = max(384, n // 384 * 384)
        pair = True
    torch.manual_seed(m + n + kk)
    a, b = _bf((m, kk), 0.5), _bf((n, kk), 0.1)
    res = torch.randn(m, n, device="cuda")
    rng = torch.tensor([m, n], dtype=torch.int64, device="cuda")
    out = torch.full((m, n), float("nan"), device="cuda")
    k.gemm(a, b, out, epilogue=k.EPI_RESID_DROPOUT, residual=res, dropout_p=p, rng_state=rng, rng_stream=7, pair=pair,
           tile_n=tile if pair else 0)
    mask = torch.ones(m, n, dtype=torch.uint8, device="cuda")
    if p > 0:
        k.dropout_mask(mask, p, rng, 7)
    torch.cuda.synchronize()
    ref = res + (a.float() @ b.float().t()) * mask.float() / (1 - p)
    err = (out - ref).abs().max().item()
    tol = 0.02 * math.sqrt(kk) + 1e-3 * ref.abs().max().item()
    assert err <= tol, f"resid+dropout {m}x{n}x{kk} p={p} pair={pair} tile={tile}: {err} > {tol}"


@settings(**_SETTINGS)
@given(b=st.integers(1, 3), s=st.integers(1, 300), heads=st.integers(1, 4), pair=st.booleans())
def test_gemm_rope_any_shape(b, s, heads, pair):
    from distributed_llm_trainer_b200.ops import kernels as k
    from distributed_llm_trainer_b200.ops import reference as R

    h, m = 64 * heads, b * s
    torch.manual_seed(m + h)
    x, w = _bf((m, h)), _bf((3 * h, h), 0.05)
    cos, sin = R.rope_tables(s, 64, device="cuda")
    out = torch.full((m, 3 * h), float("nan"), device="cuda", dtype=torch.bfloat16)
    k.gemm(x, w, out, epilogue=k.EPI_ROPE_BF16, rope_cos=cos, rope_sin=sin, seq_len=s, rope_cols=2 * h, pair=pair)
    torch.cuda.synchronize()
    ref = R.rope_packed(x.float() @ w.float().t(), cos, sin, s, 2 * h)
    err = (out.float() - ref).abs().max().item()
    assert err <= 0.02 * math.sqrt(h) + 0.05 * ref.abs().max().item() * 2 ** -7, f"rope {b}x{s}x{h} pair={pair}: {err}"


@settings(**dict(_SETTINGS, max_examples=10 * _SCALE))
@given(b=st.integers(1, 3), s=st.integers(1, 400), h=st.integers(1, 3), p=st.sampled_from([0.0, 0.1]))
def test_attention_bwd_rope_any_shape(b, s, h, p):
    """The engine's entry point (delta + zero-fill fused, dQ cast + inverse RoPE fused) against the plain one."""
    from distributed_llm_trainer_b200.ops import kernels as k
    from distributed_llm_trainer_b200.ops import reference as R

    torch.manual_seed(b * 31 + s + h)
    hd = h * 64
    qkv, dout = _bf((b * s, 3 * hd), 0.8), _bf((b * s, hd), 0.5)
    rng = torch.tensor([3, 1], dtype=torch.int64, device="cuda")
    out = torch.zeros(b * s, hd, device="cuda", dtype=torch.bfloat16)
    lse = torch.zeros(b, h, s, device="cuda")
    k.attn_fwd(qkv, out, lse, b, s, h, p, rng, 2)
    cos, sin = R.rope_tables(s, 64, device="cuda")
    res = []
    for fused in (False, True):
        dqkv = torch.zeros(b * s, 3 * hd, device="cuda", dtype=torch.bfloat16)
        delta = torch.full((b, h, s), float("nan"), device="cuda")
        dq_acc = torch.full((b * s, hd), float("nan"), device="cuda")
        if fused:
            k.attn_bwd(qkv, out, dout, lse, dqkv, delta, dq_acc, b, s, h, p, rng, 2, rope_cos=cos, rope_sin=sin)
        else:
            k.attn_bwd(qkv, out, dout, lse, dqkv, delta, dq_acc, b, s, h, p, rng, 2)
            k.rope_inplace(dqkv, cos, sin, 2 * hd, s, inverse=True)
        torch.cuda.synchronize()
        res.append(dqkv.float())
    err = (res[0] - res[1]).abs().max().item()
    assert err <= 0.02 * max(1.0, res[0].abs().max().item()), f"{b}x{s}x{h} p={p}: {err}"
