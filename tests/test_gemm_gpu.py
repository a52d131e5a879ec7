"""tcgen05 GEMM vs a plain PyTorch fp32 reference of the same product (all operand layouts,
tile widths, ragged shapes, split-K accumulation and every fused epilogue)."""
import math

import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(autouse=True, params=["1cta", "pair"])
def _kernel_choice(request):
    """Every test runs against the single-CTA kernel and against the CTA-pair kernel (which takes over
    automatically for M >= 256 when enabled)."""
    from distributed_llm_trainer_b200.ops import kernels as k

    prev = k.use_pair_kernel
    k.use_pair_kernel = request.param == "pair"
    yield
    k.use_pair_kernel = prev


def _mk(shape, scale=1.0):
    return (torch.randn(*shape, device="cuda") * scale).to(torch.bfloat16)


def _ref(a, b, a_mn, b_mn):
    A = a.float().t() if a_mn else a.float()
    B = b.float() if b_mn else b.float().t()
    return A @ B


def _check(out, ref, K, what):
    err = (out.float() - ref).abs().max().item()
    tol = 0.02 * math.sqrt(K) + 0.05 * ref.abs().max().item() * 2 ** -7
    assert err <= tol, f"{what}: max err {err:.4f} > {tol:.4f}"


@pytest.mark.parametrize("a_mn,b_mn", [(False, False), (False, True), (True, True), (True, False)])
@pytest.mark.parametrize("tile_n", [128, 256])
@pytest.mark.parametrize("M,N,K", [(256, 256, 128), (384, 768, 320), (1024, 2304, 768)])
def test_gemm_layouts(a_mn, b_mn, tile_n, M, N, K):
    from distributed_llm_trainer_b200.ops import kernels as k

    torch.manual_seed(0)
    a = _mk((K, M) if a_mn else (M, K))
    b = _mk((K, N) if b_mn else (N, K))
    c = torch.zeros(M, N, device="cuda", dtype=torch.bfloat16)
    k.gemm(a, b, c, a_mn=a_mn, b_mn=b_mn, tile_n=tile_n)
    torch.cuda.synchronize()
    _check(c, _ref(a, b, a_mn, b_mn), K, f"layout a_mn={a_mn} b_mn={b_mn} bn={tile_n}")


def test_gemm_ragged_rows_and_vocab_like():
    """M not a tile multiple; N = padded vocab-like width; K tail handled by TMA zero fill."""
    from distributed_llm_trainer_b200.ops import kernels as k

    torch.manual_seed(1)
    M, N, K = 1000, 50304 // 8, 200
    a, b = _mk((M, K)), _mk((N, K))
    c = torch.zeros(M, N, device="cuda", dtype=torch.bfloat16)
    k.gemm(a, b, c)
    torch.cuda.synchronize()
    _check(c, _ref(a, b, False, False), K, "ragged")


def test_gemm_f32_and_splitk_accumulate():
    from distributed_llm_trainer_b200.ops import kernels as k

    torch.manual_seed(2)
    M, N, K = 768, 768, 4096
    a, b = _mk((K, M)), _mk((K, N))           # wgrad-like: both MN-major
    ref = _ref(a, b, True, True)
    c = torch.zeros(M, N, device="cuda")
    k.gemm(a, b, c, a_mn=True, b_mn=True, epilogue=k.EPI_F32)
    torch.cuda.synchronize()
    _check(c, ref, K, "f32")
    acc = torch.ones(M, N, device="cuda")
    k.gemm(a, b, acc, a_mn=True, b_mn=True, epilogue=k.EPI_F32_ATOMIC, split_k=8)
    k.gemm(a, b, acc, a_mn=True, b_mn=True, epilogue=k.EPI_F32_ATOMIC, split_k=1)
    torch.cuda.synchronize()
    _check(acc, 2 * ref + 1, 2 * K, "atomic split-k")


def test_gemm_rope_epilogue():
    from distributed_llm_trainer_b200.ops import kernels as k
    from distributed_llm_trainer_b200.ops import reference as R

    torch.manual_seed(3)
    B, S, h = 2, 128, 256
    M = B * S
    x, w = _mk((M, h)), _mk((3 * h, h), 0.05)
    cos, sin = R.rope_tables(S, 64, device="cuda")
    out = torch.zeros(M, 3 * h, device="cuda", dtype=torch.bfloat16)
    k.gemm(x, w, out, epilogue=k.EPI_ROPE_BF16, rope_cos=cos, rope_sin=sin, seq_len=S, rope_cols=2 * h)
    torch.cuda.synchronize()
    ref = R.rope_packed(x.float() @ w.float().t(), cos, sin, S, 2 * h)
    _check(out, ref, h, "rope epilogue")
    # unfused companion, forward then inverse is the identity
    plain = torch.zeros_like(out)
    k.gemm(x, w, plain)
    rot = plain.clone()
    k.rope_inplace(rot, cos, sin, 2 * h, S)
    torch.cuda.synchronize()
    _check(rot, ref, h, "rope kernel")
    k.rope_inplace(rot, cos, sin, 2 * h, S, inverse=True)
    torch.cuda.synchronize()
    assert (rot.float() - plain.float()).abs().max().item() < 0.05


def test_gemm_resid_dropout_epilogue():
    from distributed_llm_trainer_b200.ops import kernels as k

    torch.manual_seed(4)
    M, N, K, p = 512, 768, 256, 0.1
    a, b = _mk((M, K)), _mk((N, K), 0.05)
    res = torch.randn(M, N, device="cuda")
    rng = torch.tensor([1234, 7], dtype=torch.int64, device="cuda")
    out = torch.zeros(M, N, device="cuda")
    k.gemm(a, b, out, epilogue=k.EPI_RESID_DROPOUT, residual=res, dropout_p=p, rng_state=rng, rng_stream=5)
    mask = torch.zeros(M, N, dtype=torch.uint8, device="cuda")
    k.dropout_mask(mask, p, rng, 5)
    torch.cuda.synchronize()
    ref = res + (a.float() @ b.float().t()) * mask.float() / (1 - p)
    _check(out, ref, K, "resid+dropout epilogue")
    frac = 1.0 - mask.float().mean().item()
    assert abs(frac - p) < 0.01, f"drop fraction {frac}"
    # unfused companion agrees with the epilogue bit-for-bit in its mask
    y = torch.zeros(M, N, device="cuda", dtype=torch.bfloat16)
    k.gemm(a, b, y)
    out2 = torch.zeros(M, N, device="cuda")
    k.resid_dropout_fwd(y, res, out2, p, rng, 5)
    torch.cuda.synchronize()
    assert ((out2 - res) == 0).eq(mask == 0).float().mean().item() > 0.999
    # backward helper uses the same mask
    dy = torch.zeros(M, N, device="cuda", dtype=torch.bfloat16)
    g = torch.randn(M, N, device="cuda")
    k.dropout_bwd_cast(g, dy, p, rng, 5)
    torch.cuda.synchronize()
    assert (dy.float() - g * mask.float() / (1 - p)).abs().max().item() < 0.05


@pytest.mark.parametrize("M,h,I", [(512, 256, 1024), (256, 64, 384), (300, 128, 640), (1000, 192, 128)])
def test_gemm_swiglu_epilogues(M, h, I):
    """I = 384 / 640 / 128: intermediate sizes that are not a multiple of the 256-wide tile -- the last tile's
    columns beyond I must neither be read as gate nor written over the first dup columns."""
    from distributed_llm_trainer_b200.ops import kernels as k

    torch.manual_seed(5)
    x, w = _mk((M, h)), _mk((2 * I, h), 0.05)   # rows [0,I) gate, [I,2I) up
    gu = torch.zeros(M, 2 * I, device="cuda", dtype=torch.bfloat16)
    act = torch.zeros(M, I, device="cuda", dtype=torch.bfloat16)
    k.gemm(x, w, gu, epilogue=k.EPI_SWIGLU, aux=act, half_n=I)
    torch.cuda.synchronize()
    ref_gu = x.float() @ w.float().t()
    _check(gu, ref_gu, h, "swiglu gu")
    g, u = gu[:, :I].float(), gu[:, I:].float()
    ref_act = torch.nn.functional.silu(g) * u
    assert (act.float() - ref_act).abs().max().item() < 0.03 * max(1.0, ref_act.abs().max().item())
    act2 = torch.zeros_like(act)
    k.swiglu_fwd(gu, act2)
    torch.cuda.synchronize()
    assert (act2.float() - act.float()).abs().max().item() < 0.02 * max(1.0, ref_act.abs().max().item())
    # backward epilogue: d_act = dy @ Wd  (Wd [h, I] is the MN-major B operand)
    dy, wd = _mk((M, h)), _mk((h, I), 0.05)
    dgu = torch.zeros(M, 2 * I, device="cuda", dtype=torch.bfloat16)
    k.gemm(dy, wd, dgu, b_mn=True, epilogue=k.EPI_SWIGLU_BWD, aux=gu, half_n=I, N=I)
    dact = (dy.float() @ wd.float())
    gg = g.clone().requires_grad_(True)
    uu = u.clone().requires_grad_(True)
    (torch.nn.functional.silu(gg) * uu).backward(dact)
    torch.cuda.synchronize()
    ref = torch.cat([gg.grad, uu.grad], dim=1)
    err = (dgu.float() - ref).abs().max().item()
    assert err < 0.03 * max(1.0, ref.abs().max().item()), err
    dgu2 = torch.zeros_like(dgu)
    k.swiglu_bwd(dact.to(torch.bfloat16), gu, dgu2)
    torch.cuda.synchronize()
    assert (dgu2.float() - ref).abs().max().item() < 0.05 * max(1.0, ref.abs().max().item())
