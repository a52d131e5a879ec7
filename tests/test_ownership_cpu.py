"""Slab ownership plan (parallel/ownership.py): partition properties that the fused reduce-scatter relies on."""
import pytest
import torch

from distributed_llm_trainer_b200.models import GPT, GPTConfig
from distributed_llm_trainer_b200.parallel.flat import plan_layout
from distributed_llm_trainer_b200.parallel.ownership import SLAB, OwnershipPlan


@pytest.mark.parametrize("world", [1, 2, 3, 4, 8])
@pytest.mark.parametrize("cfg", [dict(hidden_size=128, num_layers=2, num_heads=2, vocab_size=512),
                                 dict(hidden_size=192, num_layers=3, num_heads=3, vocab_size=1001, intermediate_size=640)])
def test_plan_partitions_every_parameter(world, cfg):
    with torch.device("meta"):
        model = GPT(GPTConfig(max_seq_len=64, **cfg))
    named = [(n, p) for n, p in model.named_parameters() if p.requires_grad]
    slots, units, total = plan_layout(named)
    plan = OwnershipPlan(slots, world)
    cover = torch.zeros(total, dtype=torch.int32)
    for r in range(world):
        rr = plan.ranges(r)
        assert rr == sorted(rr) and all(b > a and a % 4 == 0 and b % 4 == 0 for a, b in rr)
        for a, b in rr:
            cover[a:b] += 1
    for s in slots:                                  # every parameter element has exactly one owner
        assert torch.all(cover[s.offset: s.offset + s.numel] == 1), s.name
    assert int(cover.max()) == 1
    # packed groups: q|k|v and gate|up are one matrix each, slab boundaries are multiples of 128 rows
    L = cfg["num_layers"]
    assert set(plan.groups) == {"emb"} | {f"L{l}.{t}" for l in range(L) for t in ("qkv", "o", "gu", "d")}
    h = cfg["hidden_size"]
    assert plan.groups["L0.qkv"].rows == 3 * h and plan.groups["L1.gu"].rows == 2 * model.config.intermediate_size
    for g in plan.groups.values():
        rows = 0
        for o in range(world):
            a, b = g.row_range(o, world)
            assert b == a or (a % SLAB == 0 and (b % SLAB == 0 or b == g.rows))
            rows += b - a
            for s in range(*g.slab_range(o, world)):
                assert g.owner_of_slab(s, world) == o
        assert rows == g.rows
    # norm weights: owned whole
    assert {t.name for t in plan.smalls} == {n for n, p in named if p.dim() == 1}
    # load balance: no rank owns more than ~1.5x the mean once the model has a few matrices
    sizes = [sum(b - a for a, b in plan.ranges(r)) for r in range(world)]
    assert max(sizes) <= 1.6 * (sum(sizes) / world) + 128 * 1024
