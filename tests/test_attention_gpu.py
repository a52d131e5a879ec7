"""tcgen05 flash attention fwd/bwd vs the fp32 manual attention of reference gpt.py:230-234
(same dropout keep-mask, regenerated from the kernel's Philox stream)."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _attn_mask(B, H, S, p, rng, stream):
    """keep-mask [B,H,S,S] exactly as the attention kernels draw it (groups of 8 along kv)."""
    from distributed_llm_trainer_b200.ops import kernels as k

    m = torch.zeros(B * H * S, S, dtype=torch.uint8, device="cuda")
    k.attn_dropout_mask(m, p, rng, stream)
    return m.reshape(B, H, S, S).float()


def _p_eff(p):
    """drop probability the kernels realise (8-bit lanes) and rescale by"""
    return min(255, int(p * 256 + 0.5)) / 256.0


@pytest.mark.parametrize("B,S,H,p", [(2, 128, 2, 0.0), (1, 512, 3, 0.0), (2, 384, 2, 0.1), (1, 200, 2, 0.1),
                                     (2, 1024, 4, 0.1)])
def test_attention_fwd_bwd(B, S, H, p):
    from distributed_llm_trainer_b200.ops import kernels as k
    from distributed_llm_trainer_b200.ops import reference as R

    torch.manual_seed(0)
    h = H * 64
    qkv = (torch.randn(B * S, 3 * h, device="cuda") * 0.8).to(torch.bfloat16)
    rng = torch.tensor([42, 3], dtype=torch.int64, device="cuda")
    out = torch.zeros(B * S, h, device="cuda", dtype=torch.bfloat16)
    lse = torch.zeros(B, H, S, device="cuda")
    k.attn_fwd(qkv, out, lse, B, S, H, p, rng, 9)
    torch.cuda.synchronize()

    def split(t):
        return t.float().reshape(B, S, 3, H, 64).permute(2, 0, 3, 1, 4)  # [3,B,H,S,D]

    q, kk, v = [x.clone().requires_grad_(True) for x in split(qkv)]
    mask = _attn_mask(B, H, S, p, rng, 9) if p > 0 else None
    ref = R.causal_attention(q, kk, v, mask, _p_eff(p))                     # [B,H,S,D]
    ref_out = ref.permute(0, 2, 1, 3).reshape(B * S, h)
    err = (out.float() - ref_out).abs().max().item()
    assert err < 0.03, f"fwd err {err}"
    # lse (base 2, scaled) against the oracle
    sc = (q @ kk.transpose(-2, -1)) * 0.125
    sc = sc.masked_fill(torch.ones(S, S, dtype=torch.bool, device="cuda").triu(1), float("-inf"))
    ref_lse2 = torch.logsumexp(sc, dim=-1) * 1.4426950408889634
    assert (lse - ref_lse2).abs().max().item() < 0.02

    dout = (torch.randn(B * S, h, device="cuda") * 0.5).to(torch.bfloat16)
    ref_out.backward(dout.float())
    dqkv = torch.zeros(B * S, 3 * h, device="cuda", dtype=torch.bfloat16)
    delta = torch.zeros(B, H, S, device="cuda")
    dq_acc = torch.empty(B * S, h, device="cuda")
    k.attn_bwd(qkv, out, dout, lse, dqkv, delta, dq_acc, B, S, H, p, rng, 9)
    torch.cuda.synchronize()
    got = split(dqkv)
    for name, g, r_ in (("dq", got[0], q.grad), ("dk", got[1], kk.grad), ("dv", got[2], v.grad)):
        e = (g - r_).abs().max().item()
        scale = max(1.0, r_.abs().max().item())
        assert e < 0.04 * scale, f"{name} err {e} (scale {scale})"


@pytest.mark.parametrize("B,S,H,p", [(2, 384, 3, 0.1), (1, 200, 2, 0.0)])
def test_attention_bwd_with_fused_inverse_rope(B, S, H, p):
    """b200_attn_bwd_rope == b200_attn_bwd followed by rope_inplace(inverse) on the dQ|dK columns."""
    from distributed_llm_trainer_b200.ops import kernels as k

    torch.manual_seed(1)
    h = H * 64
    qkv = (torch.randn(B * S, 3 * h, device="cuda") * 0.8).to(torch.bfloat16)
    dout = (torch.randn(B * S, h, device="cuda") * 0.5).to(torch.bfloat16)
    rng = torch.tensor([7, 1], dtype=torch.int64, device="cuda")
    out = torch.zeros(B * S, h, device="cuda", dtype=torch.bfloat16)
    lse = torch.zeros(B, H, S, device="cuda")
    k.attn_fwd(qkv, out, lse, B, S, H, p, rng, 2)
    ang = torch.rand(S, 32, device="cuda") * 6.28
    cos, sin = ang.cos().contiguous(), ang.sin().contiguous()
    res = []
    for fused in (False, True):
        dqkv = torch.zeros(B * S, 3 * h, device="cuda", dtype=torch.bfloat16)
        delta = torch.full((B, H, S), float("nan"), device="cuda")
        dq_acc = torch.full((B * S, h), float("nan"), device="cuda")
        if fused:
            k.attn_bwd(qkv, out, dout, lse, dqkv, delta, dq_acc, B, S, H, p, rng, 2, rope_cos=cos, rope_sin=sin)
        else:
            k.attn_bwd(qkv, out, dout, lse, dqkv, delta, dq_acc, B, S, H, p, rng, 2)
            k.rope_inplace(dqkv, cos, sin, 2 * h, S, inverse=True)
        torch.cuda.synchronize()
        res.append((dqkv.float(), delta.clone()))
    assert torch.allclose(res[0][1], res[1][1], rtol=1e-5, atol=1e-5), "delta differs"
    err = (res[0][0] - res[1][0]).abs().max().item()
    scale = max(1.0, res[0][0].abs().max().item())
    assert err < 0.02 * scale, f"fused inverse-rope differs: {err} (scale {scale})"
    assert torch.equal(res[0][0][:, 2 * h:], res[1][0][:, 2 * h:]), "dV must be untouched by the rotary pass"
