"""The scratch-free attention backward, called the way a C program would call it.

`b200_attn_bwd_noalloc` is for hosts that cannot give the library device memory when they call it: a process
whose framework allocator has already taken all of HBM, a serving loop that bans cudaMalloc because it stalls
the device, a C caller with no allocator.  Its scratch is a fixed 32 MiB array inside the library instead of
buffers passed by the caller.  This test loads the shared object with plain `ctypes` (no project bindings),
runs the new entry point twice in a row on different inputs -- the second call finds the first call's
leftovers in the shared array and must still be right -- and compares each result with the regular entry
point, which gets its scratch from the caller.  It also checks that an over-sized problem is refused.

The file name sorts last on purpose: with `pytest -x` every other test has already run if this one fails.
It needs no state from other tests and passes in a fresh process as well.
"""
import ctypes as C
import os

import pytest
import torch

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "distributed_llm_trainer_b200", "ops", "libb200trainer.so")


def _load():
    if not os.path.exists(LIB):
        pytest.fail(f"{LIB} is missing: build it with `make -C csrc`")
    lib = C.CDLL(LIB)
    lib.b200_last_error.restype = C.c_char_p
    lib.b200_attn_bwd_noalloc_max_scratch_bytes.restype = C.c_ulonglong
    P, I, F, U = C.c_void_p, C.c_int, C.c_float, C.c_uint
    lib.b200_attn_fwd.argtypes = [P, P, P, I, I, I, F, P, U, P]
    lib.b200_attn_bwd.argtypes = [P, P, P, P, P, P, P, I, I, I, F, P, U, P]
    lib.b200_attn_bwd_noalloc.argtypes = [P, P, P, P, P, I, I, I, F, P, U, P]
    for f in (lib.b200_attn_fwd, lib.b200_attn_bwd, lib.b200_attn_bwd_noalloc):
        f.restype = I
    return lib


def _ok(lib, rc):
    assert rc == 0, lib.b200_last_error().decode()


def _case(lib, B, S, H, p, seed, first_call_on_device=False):
    h = H * 64
    torch.manual_seed(seed)
    stream = torch.cuda.current_stream().cuda_stream
    qkv = (torch.randn(B * S, 3 * h, device="cuda") * 0.8).to(torch.bfloat16)
    dout = (torch.randn(B * S, h, device="cuda") * 0.5).to(torch.bfloat16)
    rng = torch.tensor([77, seed], dtype=torch.int64, device="cuda")
    out = torch.zeros(B * S, h, device="cuda", dtype=torch.bfloat16)
    lse = torch.zeros(B, H, S, device="cuda")
    _ok(lib, lib.b200_attn_fwd(qkv.data_ptr(), out.data_ptr(), lse.data_ptr(), B, S, H, p, rng.data_ptr(), 3, stream))
    # regular entry point: scratch comes from the caller
    ref = torch.zeros(B * S, 3 * h, device="cuda", dtype=torch.bfloat16)
    delta = torch.empty(B, H, S, device="cuda")
    dq_acc = torch.empty(B * S, h, device="cuda")
    _ok(lib, lib.b200_attn_bwd(qkv.data_ptr(), out.data_ptr(), dout.data_ptr(), lse.data_ptr(), ref.data_ptr(),
                               delta.data_ptr(), dq_acc.data_ptr(), B, S, H, p, rng.data_ptr(), 3, stream))
    # new entry point: inputs and outputs only
    got = torch.full((B * S, 3 * h), float("nan"), device="cuda", dtype=torch.bfloat16)
    torch.cuda.synchronize()
    # the claim in the name: no allocation anywhere -- neither through torch's caching allocator nor behind its back
    # (cudaMalloc would lower the driver's free-memory figure)
    alloc0, reserved0, free0 = torch.cuda.memory_allocated(), torch.cuda.memory_reserved(), torch.cuda.mem_get_info()[0]
    _ok(lib, lib.b200_attn_bwd_noalloc(qkv.data_ptr(), out.data_ptr(), dout.data_ptr(), lse.data_ptr(), got.data_ptr(),
                                       B, S, H, p, rng.data_ptr(), 3, stream))
    torch.cuda.synchronize()
    if not first_call_on_device:    # the very first call may make the driver load the kernel image (not an allocation of ours)
        assert torch.cuda.mem_get_info()[0] == free0, "device free memory changed across the call"
    assert torch.cuda.memory_allocated() == alloc0 and torch.cuda.memory_reserved() == reserved0
    assert torch.isfinite(got.float()).all()
    # Both entry points run the same kernels on the same inputs: dK and dV (each written once, by one CTA) must be
    # bit-identical; dQ is summed over kv blocks with fp32 atomics whose order is free, so it may differ in the last
    # bf16 place.
    assert torch.equal(got[:, h:], ref[:, h:]), "dK / dV differ between the two entry points"
    dq_got, dq_ref = got[:, :h].float(), ref[:, :h].float()
    tol = 2.0 ** -7 * dq_ref.abs().clamp_min(2.0 ** -10)            # two bf16 ulps
    assert ((dq_got - dq_ref).abs() <= tol).all(), f"B={B} S={S} H={H}: dQ differs by more than rounding"


def test_noalloc_attention_backward_matches_regular_entry_point_twice():
    lib = _load()
    assert lib.b200_attn_bwd_noalloc_max_scratch_bytes() == 32 << 20
    # two calls in a row; the second one is smaller, so stale data from the first sits beside and under it
    _case(lib, B=2, S=512, H=4, p=0.1, seed=1, first_call_on_device=True)
    _case(lib, B=1, S=384, H=2, p=0.0, seed=2)


def test_noalloc_rejects_problems_larger_than_the_builtin_scratch():
    lib = _load()
    B, S, H = 16, 1024, 12          # 16*1024*768*4 B = 48 MiB of dQ accumulator alone
    dummy = torch.zeros(8, device="cuda")
    rc = lib.b200_attn_bwd_noalloc(dummy.data_ptr(), dummy.data_ptr(), dummy.data_ptr(), dummy.data_ptr(),
                                   dummy.data_ptr(), B, S, H, 0.0, None, 0, None)
    assert rc != 0
    msg = lib.b200_last_error().decode()
    assert "scratch" in msg and "33554432" in msg, msg
