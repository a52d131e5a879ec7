"""Property tests (hypothesis, CPU) of the host-side arithmetic every engine relies on: the flat parameter
layout, the shard geometry of DDP / FSDP, the LR schedules and the AdamW oracle.  A wrong offset here does not
crash -- it silently trains the wrong numbers -- so the invariants are stated and checked over random models."""
import math

import pytest
import torch
import torch.nn as nn

hypothesis = pytest.importorskip("hypothesis")
from hypothesis import given, settings, HealthCheck  # noqa: E402
from hypothesis import strategies as st  # noqa: E402

from distributed_llm_trainer_b200.ops.optim import adamw_reference_  # noqa: E402
from distributed_llm_trainer_b200.parallel.ddp import shard_ranges  # noqa: E402
from distributed_llm_trainer_b200.parallel.flat import ALIGN, plan_layout  # noqa: E402
from distributed_llm_trainer_b200.training.lr import cosine_lr  # noqa: E402

_SETTINGS = dict(max_examples=60, deadline=None, suppress_health_check=[HealthCheck.too_slow])


def _fake_named(layers, hidden, inter, vocab):
    """named_parameters() of a GPT-shaped model, in the reference's order (embedding, blocks, final norm)."""
    out = [("embed_tokens.weight", torch.empty(vocab, hidden))]
    for l in range(layers):
        p = f"layers.{l}."
        out += [(p + "input_layernorm.weight", torch.empty(hidden)),
                (p + "attention.q_proj.weight", torch.empty(hidden, hidden)),
                (p + "attention.k_proj.weight", torch.empty(hidden, hidden)),
                (p + "attention.v_proj.weight", torch.empty(hidden, hidden)),
                (p + "attention.o_proj.weight", torch.empty(hidden, hidden)),
                (p + "post_attention_layernorm.weight", torch.empty(hidden)),
                (p + "mlp.gate_proj.weight", torch.empty(inter, hidden)),
                (p + "mlp.up_proj.weight", torch.empty(inter, hidden)),
                (p + "mlp.down_proj.weight", torch.empty(hidden, inter))]
    out.append(("norm.weight", torch.empty(hidden)))
    return out


@settings(**_SETTINGS)
@given(layers=st.integers(0, 5), hidden=st.integers(1, 97), inter=st.integers(1, 211), vocab=st.integers(1, 300))
def test_plan_layout_invariants(layers, hidden, inter, vocab):
    named = _fake_named(layers, hidden, inter, vocab)
    slots, units, total = plan_layout(named)
    assert [s.name for s in slots] == [n for n, _ in named]            # returned in named_parameters() order
    assert [s.index for s in slots] == list(range(len(named)))
    assert total % ALIGN == 0 and len(units) == layers + 1
    # slots are ALIGN-aligned, disjoint, and inside their unit; units are contiguous, disjoint, and tile [0, total)
    spans = sorted((s.offset, s.offset + s.numel) for s in slots)
    for (a0, a1), (b0, _) in zip(spans, spans[1:]):
        assert a1 <= b0
    for s in slots:
        assert s.offset % ALIGN == 0 and s.numel == math.prod(s.shape)
        u = units[s.unit]
        assert u.lo <= s.offset and s.offset + s.numel <= u.hi
        assert s.decay == (("norm" not in s.name) and ("bias" not in s.name))
    assert units[0].lo == 0 and units[-1].hi == total
    for a, b in zip(units, units[1:]):
        assert a.hi == b.lo
    # the packed views the engine takes (q|k|v and gate|up) really are adjacent in memory
    by = {s.name: s for s in slots}
    for l in range(layers):
        p = f"layers.{l}."
        q, k_, v = (by[p + f"attention.{n}_proj.weight"] for n in "qkv")
        if (hidden * hidden) % ALIGN == 0:
            assert k_.offset == q.offset + q.numel and v.offset == k_.offset + k_.numel
        g, up = by[p + "mlp.gate_proj.weight"], by[p + "mlp.up_proj.weight"]
        if (inter * hidden) % ALIGN == 0:
            assert up.offset == g.offset + g.numel


@settings(**_SETTINGS)
@given(layers=st.integers(0, 4), hidden=st.integers(1, 64), inter=st.integers(1, 130), vocab=st.integers(1, 200),
       world=st.integers(1, 9))
def test_shard_ranges_partition_every_unit(layers, hidden, inter, vocab, world):
    _, units, total = plan_layout(_fake_named(layers, hidden, inter, vocab))
    ranges = shard_ranges(units, world)
    assert len(ranges) == world and all(len(r) == len(units) for r in ranges)
    covered = 0
    for ui, u in enumerate(units):
        cur = u.lo
        for r in range(world):
            a, b = ranges[r][ui]
            assert a == cur and a <= b <= u.hi                   # rank order, contiguous, inside the unit
            assert (a - u.lo) % ALIGN == 0                       # slices start on kernel-aligned offsets
            cur = b
            covered += b - a
        assert cur == u.hi                                       # nothing of the unit is left unowned
    assert covered == total
    # equal split up to alignment: no rank owns more than ceil(n/world) rounded up to ALIGN of any unit
    for ui, u in enumerate(units):
        cap = ((u.hi - u.lo + world - 1) // world + ALIGN - 1) // ALIGN * ALIGN
        assert all(ranges[r][ui][1] - ranges[r][ui][0] <= cap for r in range(world))


@settings(**_SETTINGS)
@given(base=st.floats(1e-6, 1.0), warm=st.integers(1, 500), extra=st.integers(1, 5000), clamp=st.booleans(),
       step=st.integers(0, 20000))
def test_cosine_lr_properties(base, warm, extra, clamp, step):
    mx = warm + extra
    lr = cosine_lr(step, base, warm, mx, clamp)
    assert 0.0 <= lr <= base * (1 + 1e-12)
    if step < warm:
        assert lr == pytest.approx(base * step / warm)
        assert cosine_lr(step + 1, base, warm, mx, clamp) >= lr          # warm-up is monotone
    elif step <= mx:
        assert lr >= 0.1 * base * (1 - 1e-9)
        if step < mx:
            assert cosine_lr(step + 1, base, warm, mx, clamp) <= lr * (1 + 1e-12)   # decay is monotone
    else:
        if clamp:                                                        # FSDP variant: pinned at the floor
            assert lr == pytest.approx(0.1 * base)
        else:                                                            # DDP variant: the cosine keeps going
            assert 0.1 * base * (1 - 1e-9) <= lr
    assert cosine_lr(warm, base, warm, mx, clamp) == pytest.approx(base)
    assert cosine_lr(mx, base, warm, mx, clamp) == pytest.approx(0.1 * base)


@settings(**dict(_SETTINGS, max_examples=25))
@given(n=st.integers(1, 400), steps=st.integers(1, 6), lr=st.floats(1e-5, 1e-2), wd=st.floats(0.0, 0.3),
       scale=st.floats(0.1, 1.0), seed=st.integers(0, 1000))
def test_adamw_oracle_matches_torch(n, steps, lr, wd, scale, seed):
    """The oracle the GPU AdamW kernels are tested against is itself torch.optim.AdamW (decoupled decay)."""
    g = torch.Generator().manual_seed(seed)
    p0 = torch.randn(n, generator=g, dtype=torch.float64)
    grads = [torch.randn(n, generator=g, dtype=torch.float64) for _ in range(steps)]
    ref = nn.Parameter(p0.clone())
    opt = torch.optim.AdamW([ref], lr=lr, betas=(0.9, 0.95), eps=1e-8, weight_decay=wd)
    p, m, v = p0.clone(), torch.zeros(n, dtype=torch.float64), torch.zeros(n, dtype=torch.float64)
    ones = torch.ones(n, dtype=torch.float64)
    for t, gr in enumerate(grads, 1):
        ref.grad = gr * scale
        opt.step()
        adamw_reference_(p, gr, m, v, ones, t, lr, 0.9, 0.95, 1e-8, wd, scale)
    assert torch.allclose(p, ref.detach(), rtol=1e-9, atol=1e-11)
