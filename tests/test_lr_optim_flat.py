"""LR schedules, flat parameter store and the flat AdamW (CPU path == kernel oracle) vs torch / the reference."""
import math

import torch

from distributed_llm_trainer_b200.models import GPT, GPTConfig
from distributed_llm_trainer_b200.ops.optim import FlatAdamW
from distributed_llm_trainer_b200.parallel.ddp import shard_ranges
from distributed_llm_trainer_b200.parallel.flat import ALIGN, FlatParamStore
from distributed_llm_trainer_b200.training.lr import cosine_lr

from ._ref import import_ref


def test_lr_matches_both_reference_trainers():
    rd = import_ref("training.ddp_trainer")
    rf = import_ref("training.fsdp_trainer")

    class Shell:
        pass

    for warm, mx in ((10, 100), (1000, 50)):
        d = Shell()
        d.training_config = rd.TrainingConfig(max_steps=mx, warmup_steps=warm)
        f = Shell()
        f.training_config = rf.TrainingConfig(max_steps=mx, warmup_steps=warm)
        for step in (0, 1, 5, 9, 10, 11, 50, 99, 100, 150):
            if warm > mx and step >= warm:
                continue
            a = rd.DistributedTrainer.get_lr(d, step)
            assert math.isclose(a, cosine_lr(step, 6e-4, warm, mx, clamp=False), rel_tol=1e-12), step
            b = rf.FSDPTrainer.get_lr(f, step)
            assert math.isclose(b, cosine_lr(step, 3e-4, warm, mx, clamp=True), rel_tol=1e-12), step
    # un-clamped DDP schedule rises again after max_steps, clamped FSDP one does not
    assert cosine_lr(150, 1.0, 10, 100, clamp=False) > cosine_lr(100, 1.0, 10, 100, clamp=False)
    assert cosine_lr(150, 1.0, 10, 100, clamp=True) == cosine_lr(100, 1.0, 10, 100, clamp=True)


def _tiny():
    torch.manual_seed(0)
    return GPT(GPTConfig(vocab_size=300, hidden_size=64, num_layers=3, num_heads=1, max_seq_len=16, dropout=0.0,
                         attention_dropout=0.0))


def test_flat_store_layout_units_and_packing():
    m = _tiny()
    before = {n: p.detach().clone() for n, p in m.named_parameters()}
    st = FlatParamStore(m, with_shadow=True)
    assert st.total % ALIGN == 0 and len(st.units) == 4 and st.units[0].name == "root"
    for u, v in zip(st.units, st.units[1:]):
        assert u.hi == v.lo                                     # units are contiguous, unit-major
    for n, p in m.named_parameters():
        assert torch.equal(p, before[n]) and p.data_ptr() == st.view(st.master, n).data_ptr()
        assert p.grad.data_ptr() == st.view(st.grad, n).data_ptr()
    qkv = st.packed(st.master, [f"layers.1.attention.{x}_proj.weight" for x in "qkv"])
    assert qkv.shape == (192, 64) and torch.equal(qkv[64:128], m.layers[1].attention.k_proj.weight)
    d, nd = st.group_indices()
    names = [s.name for s in st.slots]
    assert all("norm" in names[i] for i in nd) and "embed_tokens.weight" == names[d[0]]
    assert len(d) == 1 + 7 * 3 and len(nd) == 2 * 3 + 1         # small: 85 / 25 by the same rule
    rr = shard_ranges(st.units, 4)
    for u_idx, u in enumerate(st.units):
        covered = sum(r[u_idx][1] - r[u_idx][0] for r in rr)
        assert covered == u.hi - u.lo and all(r[u_idx][0] % ALIGN == 0 for r in rr)


def test_flat_adamw_equals_torch_adamw_and_state_dict_layout():
    m1, m2 = _tiny(), _tiny()
    decay = [p for n, p in m2.named_parameters() if "norm" not in n and "bias" not in n]
    nodecay = [p for n, p in m2.named_parameters() if "norm" in n or "bias" in n]
    topt = torch.optim.AdamW([{"params": decay, "weight_decay": 0.1}, {"params": nodecay, "weight_decay": 0.0}],
                             lr=1e-3, betas=(0.9, 0.95))
    st = FlatParamStore(m1, with_shadow=False)
    fopt = FlatAdamW(st, lr=1e-3, betas=(0.9, 0.95), weight_decay=0.1, layout="ddp")
    ids = torch.randint(0, 300, (2, 16))
    for _ in range(3):
        for mm in (m1, m2):
            _, loss = mm(ids, labels=ids)
            loss.backward()
        fopt.step()
        st.zero_grad()
        topt.step()
        topt.zero_grad()
    for (n, p), (_, q) in zip(m1.named_parameters(), m2.named_parameters()):
        assert torch.allclose(p, q, atol=1e-6), n
    a, b = fopt.state_dict(), topt.state_dict()
    assert a["param_groups"][0]["params"] == b["param_groups"][0]["params"]
    assert a["param_groups"][1]["params"] == b["param_groups"][1]["params"]
    for k in b["state"]:
        assert torch.allclose(a["state"][k]["exp_avg"], b["state"][k]["exp_avg"], atol=1e-7)
        assert float(a["state"][k]["step"]) == float(b["state"][k]["step"])
    # torch's optimizer accepts our state dict, and ours accepts torch's
    topt.load_state_dict(a)
    fopt2 = FlatAdamW(FlatParamStore(_tiny(), with_shadow=False), lr=1e-3, layout="ddp")
    fopt2.load_state_dict(b)
    assert fopt2.step_count == 3 and torch.allclose(fopt2.exp_avg, fopt.exp_avg, atol=1e-7)


def test_sharded_ranges_optimizer_equals_full():
    m1, m2 = _tiny(), _tiny()
    s1, s2 = FlatParamStore(m1, with_shadow=False), FlatParamStore(m2, with_shadow=False)
    full = FlatAdamW(s1, lr=1e-3, layout="fsdp", decay_norms=True)
    rr = shard_ranges(s2.units, 3)
    shards = [FlatAdamW(s2, lr=1e-3, layout="fsdp", decay_norms=True, ranges=r) for r in rr]
    g = torch.randn(s1.total)
    for _ in range(2):
        s1.grad.copy_(g), s2.grad.copy_(g)
        full.step()
        for sh in shards:
            sh.step()
    assert torch.allclose(s1.master, s2.master, atol=1e-7)
    sd = full.state_dict()
    assert set(sd["state"].keys()) == {s.name for s in s1.slots} and len(sd["param_groups"]) == 1
    # a sharded optimizer loads only its slices of a full state dict
    sh = FlatAdamW(s2, lr=1e-3, layout="fsdp", ranges=rr[1])
    sh.load_state_dict(sd)
    (a, b), off = rr[1][0], sh.state_offsets[0]
    assert torch.allclose(sh.exp_avg[off: off + b - a], full.exp_avg[a:b])
