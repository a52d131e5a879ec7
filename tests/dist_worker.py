"""Worker launched under torchrun by test_distributed_cpu.py (gloo, CPU) and test_multi_gpu.py (NCCL bootstrap
+ NVLink kernels).  Each mode asserts on every rank and exits non-zero on failure.

    torchrun --nproc-per-node 2 tests/dist_worker.py <mode> [tmpdir]
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from distributed_llm_trainer_b200.models import GPT, GPTConfig  # noqa: E402
from distributed_llm_trainer_b200.training.config import FSDPConfig, TrainingConfig  # noqa: E402
from distributed_llm_trainer_b200.training.ddp_trainer import DistributedTrainer  # noqa: E402
from distributed_llm_trainer_b200.training.fsdp_trainer import FSDPTrainer  # noqa: E402

GPU = torch.cuda.is_available()
CFG = dict(vocab_size=512, hidden_size=128, num_layers=2, num_heads=2, max_seq_len=128, dropout=0.0,
           attention_dropout=0.0)
MP = "bf16" if GPU else "fp32"
BACKEND = "nccl" if GPU else "gloo"


def batches(n, rows, seed=11):
    g = torch.Generator().manual_seed(seed)
    return [torch.randint(0, 512, (rows, 128), generator=g) for _ in range(n)]


def close(a, b, tol, what):
    err = (a.float() - b.float()).abs().max().item()
    ref = b.float().abs().max().item() + 1e-12
    assert err <= tol * ref, f"{what}: max err {err:.3e} vs scale {ref:.3e} (tol {tol})"


def all_ranks_equal(t, what):
    ws = dist.get_world_size()
    got = [torch.empty_like(t) for _ in range(ws)]
    dist.all_gather(got, t.contiguous())
    for r in range(1, ws):
        assert torch.equal(got[0], got[r]), f"{what}: rank {r} differs from rank 0"


def single_process_reference(kind, world, ga, steps, strategy="FULL_SHARD"):
    """What one process computes for the same global batch: world*ga micro-steps per optimizer step."""
    os_rank = os.environ.pop("RANK")
    try:
        torch.manual_seed(0)
        model = GPT(GPTConfig(**CFG))
        if kind == "ddp":
            tr = _SoloDDP(model, ga * world)
        else:
            tr = _SoloFSDP(model, ga * world, strategy)
    finally:
        os.environ["RANK"] = os_rank
    return tr


class _Solo:
    """Build a trainer object without touching torch.distributed (bypasses _setup_distributed)."""


def _solo_common(cls, tr, model):
    tr.distributed, tr.rank, tr.world_size, tr.local_rank, tr.is_main_process = False, 0, 1, int(
        os.environ.get("LOCAL_RANK", 0)), False
    tr._setup_device()
    tr._setup_model(model)
    tr._setup_optimizer()
    tr.global_step, tr.tokens_seen = 0, 0
    tr.last_grad_norm = None
    return tr


def _SoloDDP(model, ga):
    tr = DistributedTrainer.__new__(DistributedTrainer)
    tr.model_config = GPTConfig(**CFG)
    tr.training_config = TrainingConfig(batch_size=2, gradient_accumulation_steps=ga, max_steps=50, warmup_steps=2,
                                        mixed_precision=MP, seed=0)
    return _solo_common(DistributedTrainer, tr, model)


def _SoloFSDP(model, ga, strategy):
    tr = FSDPTrainer.__new__(FSDPTrainer)
    tr.model_config = GPTConfig(**CFG)
    tr.training_config = TrainingConfig.fsdp_defaults(batch_size=2, gradient_accumulation_steps=ga, max_steps=50,
                                                      warmup_steps=2, seed=0)
    tr.fsdp_config = FSDPConfig(sharding_strategy=strategy, mixed_precision=MP, activation_checkpointing=True)
    return _solo_common(FSDPTrainer, tr, model)


def interleave(per_rank):
    """Global batch in the order the solo run must see it: micro-step m of rank r -> position r*ga + m is not
    required (sum is order independent); we simply concatenate the ranks' batches."""
    return torch.cat(per_rank, dim=0)


# ------------------------------------------------------------------------------------------------ modes
def mode_ddp(tmp, ga=2):
    steps = 3
    tcfg = TrainingConfig(batch_size=2, gradient_accumulation_steps=ga, max_steps=50, warmup_steps=2,
                          mixed_precision=MP, seed=0, backend=BACKEND, checkpoint_dir=tmp)
    tr = DistributedTrainer(GPTConfig(**CFG), tcfg)
    rank, world = tr.rank, tr.world_size
    solo = single_process_reference("ddp", world, ga, steps)
    solo.raw_model.load_state_dict(tr.raw_model.state_dict())
    if solo.engine is not None:
        solo.engine.store.refresh_shadow()
    data = [batches(steps, 2 * ga, seed=100 + r) for r in range(world)]
    for s in range(steps):
        out = tr.train_step({"input_ids": data[rank][s]})
        ref = solo.train_step({"input_ids": interleave([data[r][s] for r in range(world)])})
        lt = torch.tensor([out["loss"]], device=tr.device)
        dist.all_reduce(lt)
        assert abs(lt.item() / world - ref["loss"]) < (2e-2 if GPU else 1e-5) * abs(ref["loss"]), (lt.item() / world, ref)
        assert out["tokens"] == ref["tokens"]
    if tr.grad_sync is not None:
        tr.grad_sync.gather_master_and_state()
        all_ranks_equal(tr.engine.store.shadow, "bf16 shadow params")
    for (n, p), (_, q) in zip(tr.raw_model.named_parameters(), solo.raw_model.named_parameters()):
        close(p.detach(), q.detach(), 2e-2 if GPU else 1e-5, n)
        all_ranks_equal(p.detach(), n)
    # checkpoint: collective save, reference layout, reload into a fresh trainer
    path = os.path.join(tmp, "ddp.pt")
    tr.save_checkpoint(path)
    dist.barrier()
    ck = torch.load(path, map_location="cpu", weights_only=False)
    assert ck["global_step"] == steps and len(ck["optimizer"]["param_groups"]) == 2
    assert len(ck["optimizer"]["state"]) == len(list(tr.raw_model.parameters()))
    tr2 = DistributedTrainer(GPTConfig(**CFG), TrainingConfig(batch_size=2, gradient_accumulation_steps=ga,
                                                              max_steps=50, warmup_steps=2, mixed_precision=MP, seed=5,
                                                              backend=BACKEND, resume_from=path))
    nb = batches(1, 2 * ga, seed=500 + rank)[0]
    a, b = tr.train_step({"input_ids": nb}), tr2.train_step({"input_ids": nb})
    assert abs(a["loss"] - b["loss"]) < (1e-2 if GPU else 1e-5) * abs(a["loss"]), (a, b)


def mode_fsdp(tmp, strategy, offload=False):
    ga, steps = 2, 3
    tcfg = TrainingConfig.fsdp_defaults(batch_size=2, gradient_accumulation_steps=ga, max_steps=50, warmup_steps=2,
                                        seed=0, backend=BACKEND, checkpoint_dir=tmp)
    fcfg = FSDPConfig(sharding_strategy=strategy, mixed_precision=MP, activation_checkpointing=True,
                      cpu_offload=offload)
    tr = FSDPTrainer(GPTConfig(**CFG), tcfg, fcfg)
    rank, world = tr.rank, tr.world_size
    solo = single_process_reference("fsdp", world, ga, steps, strategy)
    if getattr(tr, "sharded", None) is not None:
        tr.sharded.gather_full_master()           # sharded fp32 master: materialise to read the initial weights
    solo.model.load_state_dict(tr.model.state_dict())
    if getattr(tr, "sharded", None) is not None:
        tr.sharded.release_full_master()
    if solo.engine is not None:
        solo.engine.store.refresh_shadow()
    data = [batches(steps, 2 * ga, seed=200 + r) for r in range(world)]
    for s in range(steps):
        out = tr.train_step({"input_ids": data[rank][s]})
        ref = solo.train_step({"input_ids": interleave([data[r][s] for r in range(world)])})
        lt = torch.tensor([out["loss"]], device=tr.device)
        dist.all_reduce(lt)
        assert abs(lt.item() / world - ref["loss"]) < (2e-2 if GPU else 1e-5) * abs(ref["loss"]), (lt.item() / world, ref)
    path = os.path.join(tmp, f"fsdp_{strategy}.pt")
    tr.save_checkpoint(path)          # gathers full fp32 parameters + optimizer state (collective)
    sharded = getattr(tr, "sharded", None)
    if sharded is not None:           # fp32 master weights are sharded: outside of save / load the module's
        assert all(p.numel() == 0 for p in tr.model.parameters())      # parameters hold no storage
        sharded.gather_full_master()
    for (n, p), (_, q) in zip(tr.model.named_parameters(), solo.model.named_parameters()):
        close(p.detach(), q.detach(), 2e-2 if GPU else 1e-5, n)
    if sharded is not None:
        sharded.release_full_master()
    ck = torch.load(path, map_location="cpu", weights_only=False)
    assert ck["fsdp_config"]["sharding_strategy"] == strategy
    assert set(ck["optimizer"]["state"]) == {n for n, _ in solo.model.named_parameters()}
    solo_sd = solo.optimizer.state_dict() if solo.optimizer.is_full else None
    if solo_sd is not None:
        for k, v in ck["optimizer"]["state"].items():
            close(v["exp_avg"], solo_sd["state"][k]["exp_avg"].cpu(), 5e-2 if GPU else 1e-5, f"exp_avg {k}")
    tr2 = FSDPTrainer(GPTConfig(**CFG), TrainingConfig.fsdp_defaults(batch_size=2, gradient_accumulation_steps=ga,
                                                                    max_steps=50, warmup_steps=2, seed=9,
                                                                    backend=BACKEND, resume_from=path), fcfg)
    nb = batches(1, 2 * ga, seed=600 + rank)[0]
    a, b = tr.train_step({"input_ids": nb}), tr2.train_step({"input_ids": nb})
    assert abs(a["loss"] - b["loss"]) < (1e-2 if GPU else 1e-5) * abs(a["loss"]), (a, b)


def mode_comm(tmp):
    """NVLink comm kernels vs torch.distributed collectives (GPU only)."""
    from distributed_llm_trainer_b200.ops import kernels as K
    from distributed_llm_trainer_b200.parallel.comm import SignalPads, SymmetricBuffer

    dist.init_process_group("nccl", device_id=torch.device(f"cuda:{int(os.environ['LOCAL_RANK'])}"))
    rank, world = dist.get_rank(), dist.get_world_size()
    dev = torch.device(f"cuda:{int(os.environ['LOCAL_RANK'])}")
    torch.cuda.set_device(dev)
    pads = SignalPads(dev)
    n = 128 * 1000 * world
    g = SymmetricBuffer(n, torch.float32, dev)
    torch.manual_seed(rank)
    g.local.copy_(torch.randn(n, device=dev))
    want = g.local.clone()
    dist.all_reduce(want)
    part = n // world
    lo, hi = rank * part, (rank + 1) * part
    sumsq = torch.zeros(1, device=dev)
    torch.cuda.synchronize()
    for it in range(20):                      # stress the epoch protocol
        pads.barrier()
    pads.barrier()
    K.reduce_scatter_f32(g.ptrs, g.local, lo, hi, 1.0 / world, sumsq, rank, world)
    pads.barrier()
    torch.cuda.synchronize()
    close(g.local[lo:hi], want[lo:hi] / world, 1e-6, "reduce_scatter_f32")
    assert abs(sumsq.item() - (want[lo:hi] / world).pow(2).sum().item()) < 1e-3 * sumsq.item()
    tot_in, tot = torch.zeros(8, device=dev), torch.zeros(8, device=dev)
    tot_in[0], tot_in[1] = float(rank + 1), 2.0
    for _ in range(5):
        pads.allreduce_scalars(tot_in, tot, 2)
    torch.cuda.synchronize()
    assert tot[0].item() == world * (world + 1) / 2 and tot[1].item() == 2.0 * world
    # all-gather (bf16) by P2P pull
    shard = SymmetricBuffer(4096, torch.bfloat16, dev)
    shard.local.copy_((torch.arange(4096, device=dev) % 251 + rank * 7).to(torch.bfloat16))
    full = torch.zeros(4096 * world, dtype=torch.bfloat16, device=dev)
    pads.barrier()
    K.all_gather_bf16(shard.ptrs, full, 4096, rank, world)
    pads.barrier()
    torch.cuda.synchronize()
    exp = [torch.empty(4096, dtype=torch.bfloat16, device=dev) for _ in range(world)]
    dist.all_gather(exp, shard.local.clone())
    assert torch.equal(full, torch.cat(exp))
    # fused AdamW + broadcast: every rank's shadow gets every rank's slice
    sh = SymmetricBuffer(n, torch.bfloat16, dev)
    master = torch.ones(n, device=dev)
    m, v = torch.zeros(part, device=dev), torch.zeros(part, device=dev)
    decay = torch.ones(n // 128, dtype=torch.uint8, device=dev)
    scale = torch.ones(1, device=dev)
    pads.barrier()
    K.adamw_bcast(master, g.local, m, v, sh.ptrs, rank, world, decay, lo, hi, 1, 1e-2, 0.9, 0.95, 1e-8, 0.1, scale)
    pads.barrier()
    torch.cuda.synchronize()
    mine = master[lo:hi].to(torch.bfloat16)
    got = [torch.empty(part, dtype=torch.bfloat16, device=dev) for _ in range(world)]
    dist.all_gather(got, mine)
    assert torch.equal(sh.local, torch.cat(got)), "adamw_bcast: shadow mismatch"
    # staging-buffer reduce-scatter with accumulation (FSDP)
    acc = torch.ones(part, device=dev)
    pads.barrier()
    K.reduce_scatter_f32_accum(g.ptrs, acc, lo, part, 0.5, rank, world)
    pads.barrier()
    torch.cuda.synchronize()
    gsum = g.local.clone()
    dist.all_reduce(gsum)
    close(acc, 1 + 0.5 * gsum[lo:hi], 1e-5, "reduce_scatter_f32_accum")
    dist.barrier()
    for b in (g, shard, sh, pads.buf):
        b.close()


def mode_aggemm(tmp):
    """All-gather fused into the GEMM: every rank owns a slice of a 'unit' that contains the weight; one launch
    gathers the weight rows over NVLink while it multiplies.  Checked against a plain matmul on the full weight."""
    from distributed_llm_trainer_b200.ops import kernels as K
    from distributed_llm_trainer_b200.ops import reference as R
    from distributed_llm_trainer_b200.parallel.comm import SignalPads, SymmetricBuffer

    local = int(os.environ["LOCAL_RANK"])
    dev = torch.device(f"cuda:{local}")
    torch.cuda.set_device(dev)
    dist.init_process_group("nccl", device_id=dev)
    rank, world = dist.get_rank(), dist.get_world_size()
    pads = SignalPads(dev)
    N, Kd, M, S = 1536, 512, 2048, 256
    b_off = 4096 + 128 * 3                      # the weight does not start on a shard boundary
    unit = b_off + N * Kd + 1000
    part = ((unit + world - 1) // world + 127) // 128 * 128
    part = max(part, 16384)
    torch.manual_seed(1234)                      # same full unit on every rank; each keeps only its slice
    full = (torch.randn(part * world) * 0.05).to(torch.bfloat16)
    shard = SymmetricBuffer(part, torch.bfloat16, dev)
    shard.local.copy_(full[rank * part:(rank + 1) * part])
    w_ref = full[b_off: b_off + N * Kd].view(N, Kd).to(dev)
    x = (torch.randn(M, Kd, device=dev)).to(torch.bfloat16)
    flags = torch.zeros((N + 127) // 128, dtype=torch.int32, device=dev)
    cos, sin = R.rope_tables(S, 64, device=dev)
    pads.barrier()
    torch.cuda.synchronize()
    for epoch, epi in ((1, K.EPI_BF16), (2, K.EPI_ROPE_BF16), (3, K.EPI_BF16)):
        gathered = torch.zeros(N, Kd, dtype=torch.bfloat16, device=dev)      # nothing local to start from
        out = torch.full((M, N), float("nan"), dtype=torch.bfloat16, device=dev)
        kw = dict(rope_cos=cos, rope_sin=sin, seq_len=S, rope_cols=1024) if epi == K.EPI_ROPE_BF16 else {}
        K.gemm(x, gathered, out, epilogue=epi, ag=(shard.ptrs, part, b_off, flags, epoch, rank, world, 2 + epoch % 2), **kw)
        torch.cuda.synchronize()
        assert torch.equal(gathered, w_ref), f"epoch {epoch}: gathered weight differs from the owners' rows"
        ref = x.float() @ w_ref.float().t()
        if epi == K.EPI_ROPE_BF16:
            ref = R.rope_packed(ref, cos, sin, S, 1024)
        close(out, ref, 2e-2, f"fused all-gather GEMM (epoch {epoch})")
        assert int(flags.min().item()) == epoch
    pads.barrier()
    torch.cuda.synchronize()
    dist.barrier()
    shard.close()
    pads.buf.close()


def mode_rsgemm(tmp):
    """GEMM -> reduce-scatter over real NVLink peers: every rank runs the same wgrad GEMM on its own data and the
    128-row slabs of the result are reduce-added into their owners' buffers; after a barrier each owner must hold
    the all-reduced rows.  Same for the owner-directed embedding scatter."""
    from distributed_llm_trainer_b200.ops import kernels as K
    from distributed_llm_trainer_b200.parallel.comm import SignalPads, SymmetricBuffer, wrap_pointer
    from distributed_llm_trainer_b200.parallel.ownership import SlabGroup

    local = int(os.environ["LOCAL_RANK"])
    dev = torch.device(f"cuda:{local}")
    torch.cuda.set_device(dev)
    dist.init_process_group("nccl", device_id=dev)
    rank, world = dist.get_rank(), dist.get_world_size()
    pads = SignalPads(dev)
    for n_out, k_in, T, rot in ((2304, 768, 2048, 0), (768, 768, 1024, 1), (1000, 256, 512, 3)):
        g = SlabGroup("t", 0, n_out, k_in, rot % world, ("t",))
        buf = SymmetricBuffer(n_out * k_in, torch.float32, dev)       # full-size layout, owned rows in place
        bases = []
        for o in range(world):
            a, b = g.row_range(o, world)
            bases.append(buf.raw_ptrs[o] + 4 * a * k_in if b > a else None)
        rs = K.make_rs_args(bases, g.rot)
        torch.manual_seed(77 + rank)
        dy = (torch.randn(T, n_out, device=dev) * 0.3).to(torch.bfloat16)
        x = (torch.randn(T, k_in, device=dev) * 0.3).to(torch.bfloat16)
        want = dy.float().t() @ x.float()
        dist.all_reduce(want)
        pads.barrier()
        for rep in (1, 2):
            K.gemm(dy, x, buf.local.view(n_out, k_in), a_mn=True, b_mn=True, epilogue=K.EPI_F32_ATOMIC, split_k=rep,
                   tile_n=256 if k_in % 256 == 0 else 128, M=n_out, N=k_in, K=T, rs=rs)
            pads.barrier()
            torch.cuda.synchronize()
            a, b = g.row_range(rank, world)
            if b > a:
                got = buf.local.view(n_out, k_in)[a:b]
                err = ((got - rep * want[a:b]).norm() / (rep * want[a:b]).norm()).item()
                assert err < 2e-3, f"gemm_rs {n_out}x{k_in} rep {rep}: rel L2 {err:.3e}"
            mine = torch.zeros(n_out, dtype=torch.bool, device=dev)
            mine[a:b] = True
            assert torch.all(buf.local.view(n_out, k_in)[~mine] == 0), "rows this rank does not own were written"
        # embedding scatter into the same owners
        buf.local.zero_()
        pads.barrier()
        ids = torch.randint(0, n_out, (4096,), device=dev)
        dx = torch.randn(4096, k_in, device=dev)
        ref = torch.zeros(n_out, k_in, device=dev)
        ref.index_add_(0, ids, dx)
        dist.all_reduce(ref)
        K.embed_bwd_rs(ids, dx, rs, n_out)
        pads.barrier()
        torch.cuda.synchronize()
        a, b = g.row_range(rank, world)
        if b > a:
            close(buf.local.view(n_out, k_in)[a:b], ref[a:b], 1e-4, "embed_bwd_rs")
        dist.barrier()
        buf.close()
    pads.buf.close()


def mode_nvls(tmp):
    """NVLS multicast: the AdamW parameter broadcast as one multimem.st per element and the small-tensor sum over
    ranks as one multimem.ld_reduce, against the per-rank store / load forms of the same kernels."""
    from distributed_llm_trainer_b200.ops import kernels as K
    from distributed_llm_trainer_b200.parallel.comm import MulticastBuffer, SignalPads

    local = int(os.environ["LOCAL_RANK"])
    dev = torch.device(f"cuda:{local}")
    torch.cuda.set_device(dev)
    dist.init_process_group("nccl", device_id=dev)
    rank, world = dist.get_rank(), dist.get_world_size()
    assert MulticastBuffer.available()
    pads = SignalPads(dev)
    n = 128 * 64 * world
    sh = MulticastBuffer(n, torch.bfloat16, dev)
    acc = MulticastBuffer(n, torch.float32, dev)
    part = n // world
    lo = rank * part
    torch.manual_seed(5)                                  # same "model" everywhere; every rank updates its part
    master = torch.randn(n, device=dev)
    grad = torch.randn(n, device=dev)
    m, v = torch.zeros(part, device=dev), torch.zeros(part, device=dev)
    decay = torch.ones(n // 128, dtype=torch.uint8, device=dev)
    scale = torch.ones(1, device=dev)
    table, nch = K.make_chunk_table([(lo, lo, 0, lo, part, -1)], dev)
    pads.barrier()
    K.adamw_chunks(master, grad, m, v, sh.ptrs, rank, world, decay, table, nch, 1, 1e-2, 0.9, 0.95, 1e-8, 0.1, scale,
                   zero_grad=False, multicast_ptr=sh.mc_ptr)
    pads.barrier()
    torch.cuda.synchronize()
    mine = master[lo: lo + part].to(torch.bfloat16)
    got = [torch.empty(part, dtype=torch.bfloat16, device=dev) for _ in range(world)]
    dist.all_gather(got, mine)
    assert torch.equal(sh.local, torch.cat(got)), "multimem.st broadcast: some rank's copy differs"
    # sum over ranks inside the switch + clearing every copy
    torch.manual_seed(100 + rank)
    acc.local.copy_(torch.randn(n, device=dev))
    want = acc.local.clone()
    dist.all_reduce(want)
    dst = torch.zeros(n, device=dev)
    pads.barrier()
    K.pull_reduce_chunks(acc.ptrs, dst, table, nch, rank, world, zero_remote=True, multicast_ptr=acc.mc_ptr)
    pads.barrier()
    torch.cuda.synchronize()
    close(dst[lo: lo + part], want[lo: lo + part], 1e-5, "multimem.ld_reduce")
    assert torch.all(acc.local == 0), "multimem.st of zeros did not clear this rank's copy"
    dist.barrier()
    sh.close()
    acc.close()
    pads.buf.close()


def main():
    mode = sys.argv[1]
    tmp = sys.argv[2] if len(sys.argv) > 2 else "/tmp/b200_dist_test"
    os.makedirs(tmp, exist_ok=True)
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    if mode == "ddp":
        mode_ddp(tmp)
    elif mode == "ddp:ga1":      # a step of ONE micro-step: nothing accumulated locally, the wgrads deliver directly
        mode_ddp(tmp, ga=1)
    elif mode == "ddp:ga3":      # store / add / deliver micro-steps all present
        mode_ddp(tmp, ga=3)
    elif mode == "rsgemm":
        mode_rsgemm(tmp)
    elif mode == "nvls":
        mode_nvls(tmp)
    elif mode.startswith("fsdp"):
        parts = mode.split(":")
        opt = parts[2] if len(parts) > 2 else ""
        if opt == "fuseag":       # QKV rows gathered inside the QKV GEMM (b200_gemm2_ag_bf16), rest prefetched
            os.environ["B200_FSDP_FUSE_AG"] = "1"
        if opt == "smgather":     # P2P-pull SM kernel instead of copy-engine gathers
            os.environ["B200_FSDP_CE_GATHER"] = "0"
        if opt == "everymicro":   # every micro-step's wgrad GEMMs deliver to the owners (no local accumulation)
            os.environ["B200_FSDP_ACCUM_LOCAL"] = "0"
        if opt == "nographs":
            os.environ["B200_FSDP_GRAPHS"] = "0"
        mode_fsdp(tmp, parts[1], offload=opt == "offload")
    elif mode == "comm":
        mode_comm(tmp)
    elif mode == "aggemm":
        mode_aggemm(tmp)
    else:
        raise SystemExit(f"unknown mode {mode}")
    if dist.is_initialized():
        dist.barrier()
        dist.destroy_process_group()
    print(f"rank {os.environ.get('RANK')} mode {mode} OK")


if __name__ == "__main__":
    main()
