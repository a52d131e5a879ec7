"""Stricter oracles for the fused engine:

* the BENCHMARK shape (small preset: h=768, 12 heads, V=50257, S=1024 -- 384-wide GEMM tiles, split-K LM-head dX,
  two logits chunks) against the fp32 PyTorch model, relative-L2 per parameter gradient;
* dropout: keep rate within 3 sigma per stream, independence across streams / step offsets, and an engine step with
  p = 0.1 against an fp32 restatement of the model that applies the KERNELS' masks -- forward loss and every gradient,
  with and without activation recompute (so fwd mask == bwd mask == recompute mask is asserted through the engine).
"""
import copy
import math

import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


def _rel_l2(a, b):
    return ((a.float() - b.float()).norm() / (b.float().norm() + 1e-12)).item()


def test_engine_at_benchmark_shape_matches_fp32():
    from distributed_llm_trainer_b200.models import GPT, GPTConfig

    cfg = GPTConfig.preset("small", dropout=0.0, attention_dropout=0.0)
    torch.manual_seed(0)
    model = GPT(cfg)
    ref = copy.deepcopy(model).cuda().float()
    ref.backend = "torch"
    model = model.cuda()
    eng = model.engine()
    eng.use_graphs = False
    ids = torch.randint(0, cfg.vocab_size, (2, 1024), device="cuda")
    ref.train()
    _, ref_loss = ref(ids, labels=ids)
    ref_loss.backward()
    model.train()
    eng.store.zero_grad()
    loss = eng.forward_backward(ids, ids, 1.0)
    torch.cuda.synchronize()
    assert abs(loss.item() - ref_loss.item()) < 2e-3 * ref_loss.item(), (loss.item(), ref_loss.item())
    worst = []
    for (n, p), (_, q) in zip(model.named_parameters(), ref.named_parameters()):
        e = _rel_l2(p.grad, q.grad)
        if e > (0.03 if p.dim() == 1 else 0.025):     # q/k projections sit at 2.0-2.1 % (bf16 dS through RoPE)
            worst.append((n, round(e, 4)))
    assert not worst, f"relative L2 of gradients above bound: {worst[:8]}"


@pytest.mark.parametrize("p", [0.1, 0.25])
def test_dropout_mask_statistics(p):
    from distributed_llm_trainer_b200.ops import kernels as k

    rows, cols = 2048, 1024
    n = rows * cols
    rng = torch.tensor([1234, 0], dtype=torch.int64, device="cuda")
    masks = {}
    for kind, fn, p_eff in (("elementwise", k.dropout_mask, round(p * 65536) / 65536),
                            ("attention", k.attn_dropout_mask, round(p * 256) / 256)):
        for stream, offset in ((3, 0), (4, 0), (3, 1)):
            rng[1] = offset
            m = torch.zeros(rows, cols, dtype=torch.uint8, device="cuda")
            fn(m, p, rng, stream)
            torch.cuda.synchronize()
            keep = m.float().mean().item()
            sigma = math.sqrt(p_eff * (1 - p_eff) / n)
            assert abs(keep - (1 - p_eff)) < 3 * sigma + 1e-9, (kind, stream, offset, keep, 1 - p_eff)
            # rows and columns individually, too (a generator that repeats per row would pass the global rate)
            assert (m.float().mean(0) - (1 - p_eff)).abs().max().item() < 6 * math.sqrt(p_eff * (1 - p_eff) / rows)
            assert (m.float().mean(1) - (1 - p_eff)).abs().max().item() < 6 * math.sqrt(p_eff * (1 - p_eff) / cols)
            masks[(kind, stream, offset)] = m
        a, b, c = (masks[(kind, 3, 0)].float(), masks[(kind, 4, 0)].float(), masks[(kind, 3, 1)].float())
        for other, name in ((b, "rng_stream"), (c, "step offset")):
            # independent masks agree on a fraction keep^2 + drop^2 of the positions
            agree = (a == other).float().mean().item()
            expect = (1 - p_eff) ** 2 + p_eff ** 2
            assert abs(agree - expect) < 5e-3, (kind, name, agree, expect)


def _oracle_forward(model, ids, masks, p_res, p_att_eff):
    """fp32 restatement of the reference forward (gpt.py:299-316, 388-455) with the kernels' keep-masks."""
    from distributed_llm_trainer_b200.ops import reference as R

    cfg = model.config
    B, S = ids.shape
    H, h = cfg.num_heads, cfg.hidden_size
    cos, sin = R.rope_tables(S, 64, device=ids.device)
    x = model.embed_tokens.weight[ids.reshape(-1)]
    for l, layer in enumerate(model.layers):
        att, mlp = layer.attention, layer.mlp
        xn, _ = R.rmsnorm(x, layer.input_layernorm.weight)
        w = torch.cat([att.q_proj.weight, att.k_proj.weight, att.v_proj.weight], 0)
        qkv = R.rope_packed(xn @ w.t(), cos, sin, S, 2 * h)
        q, k_, v = [t.reshape(B, S, H, 64).permute(0, 2, 1, 3) for t in qkv.split(h, dim=1)]
        a = R.causal_attention(q, k_, v, masks[8 * l], p_att_eff).permute(0, 2, 1, 3).reshape(B * S, h)
        x = x + (a @ att.o_proj.weight.t()) * masks[8 * l + 1] / (1 - p_res)
        xn, _ = R.rmsnorm(x, layer.post_attention_layernorm.weight)
        act = F.silu(xn @ mlp.gate_proj.weight.t()) * (xn @ mlp.up_proj.weight.t())
        x = x + (act @ mlp.down_proj.weight.t()) * masks[8 * l + 2] / (1 - p_res)
    xn, _ = R.rmsnorm(x, model.norm.weight)
    logits = (xn @ model.embed_tokens.weight.t()).reshape(B, S, -1)
    return F.cross_entropy(logits[:, :-1].reshape(-1, logits.shape[-1]), ids[:, 1:].reshape(-1))


@pytest.mark.parametrize("ckpt", [False, True])
def test_engine_with_dropout_matches_oracle_using_the_kernel_masks(ckpt):
    from distributed_llm_trainer_b200.models import GPT, GPTConfig
    from distributed_llm_trainer_b200.ops import kernels as k

    p = 0.1
    cfg = GPTConfig(vocab_size=1000, hidden_size=128, num_layers=2, num_heads=2, max_seq_len=256, dropout=p,
                    attention_dropout=p, gradient_checkpointing=ckpt)
    torch.manual_seed(0)
    model = GPT(cfg)
    ref = copy.deepcopy(model).cuda().float()
    model = model.cuda()
    eng = model.engine()
    eng.use_graphs = False
    eng.set_seed(99)
    B, S = 2, 256
    ids = torch.randint(0, cfg.vocab_size, (B, S), device="cuda")
    model.train()
    for step in range(2):                 # second step: the device-side offset has advanced, masks must follow it
        rng = eng.rng_state.clone()
        masks = {}
        for l in range(cfg.num_layers):
            ma = torch.zeros(B * cfg.num_heads * S, S, dtype=torch.uint8, device="cuda")
            k.attn_dropout_mask(ma, p, rng, 8 * l)
            masks[8 * l] = ma.reshape(B, cfg.num_heads, S, S).float()
            for s_ in (1, 2):
                m = torch.zeros(B * S, cfg.hidden_size, dtype=torch.uint8, device="cuda")
                k.dropout_mask(m, p, rng, 8 * l + s_)
                masks[8 * l + s_] = m.float()
        ref.zero_grad(set_to_none=True)
        ref_loss = _oracle_forward(ref, ids, masks, p, round(p * 256) / 256)
        ref_loss.backward()
        eng.store.zero_grad()
        loss = eng.forward_backward(ids, ids, 1.0)
        torch.cuda.synchronize()
        assert abs(loss.item() - ref_loss.item()) < 5e-3 * ref_loss.item(), (step, loss.item(), ref_loss.item())
        bad = [(n, round(_rel_l2(p_.grad, q.grad), 4)) for (n, p_), (_, q) in
               zip(model.named_parameters(), ref.named_parameters()) if _rel_l2(p_.grad, q.grad) > 0.03]
        assert not bad, f"step {step}: {bad}"
