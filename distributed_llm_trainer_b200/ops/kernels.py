"""Python bindings of the C ABI in ``include/b200_trainer.h``: thin, allocation-free wrappers that
take torch tensors, pass raw device pointers + the current CUDA stream, and raise on any error.

Nothing here falls back to PyTorch: the pure-PyTorch statements of the same ops live in
``ops/reference.py`` and are used only as test oracles and by CPU-only processes.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional

import os
import torch

from . import _lib

EPI_BF16, EPI_F32, EPI_F32_ATOMIC, EPI_ROPE_BF16, EPI_RESID_DROPOUT, EPI_SWIGLU, EPI_SWIGLU_BWD = range(7)

# launch counter: bench.py reports how many of OUR kernels ran inside the timed region
launch_count = 0
# route plain / fp32-accumulate GEMMs to the CTA-pair kernel (set False to A/B against the 1-CTA kernel)
use_pair_kernel = True


class GemmArgs(C.Structure):
    _fields_ = [
        ("A", C.c_void_p), ("B", C.c_void_p), ("C", C.c_void_p),
        ("M", C.c_int), ("N", C.c_int), ("K", C.c_int),
        ("lda", C.c_int), ("ldb", C.c_int), ("ldc", C.c_int),
        ("a_mn_major", C.c_int), ("b_mn_major", C.c_int),
        ("epilogue", C.c_int), ("split_k", C.c_int), ("tile_n", C.c_int), ("max_ctas", C.c_int),
        ("rope_cos", C.c_void_p), ("rope_sin", C.c_void_p), ("seq_len", C.c_int), ("rope_cols", C.c_int),
        ("residual", C.c_void_p), ("dropout_p", C.c_float), ("rng_state", C.c_void_p), ("rng_stream", C.c_uint),
        ("aux", C.c_void_p), ("ld_aux", C.c_int), ("half_n", C.c_int), ("raster", C.c_int),
    ]


class AgArgs(C.Structure):
    """``b200_ag_args``: where the rows of a sharded weight live on every rank (fused all-gather GEMM)."""
    _fields_ = [
        ("shard_ptrs", C.c_void_p), ("part_elems", C.c_longlong), ("b_offset_elems", C.c_longlong),
        ("flags", C.c_void_p), ("epoch", C.c_uint), ("rank", C.c_int), ("world", C.c_int), ("gather_pairs", C.c_int),
        ("epoch_dev", C.c_void_p),
    ]


class RsArgs(C.Structure):
    """``b200_rs_args``: where every rank keeps its rows of a weight gradient (GEMM -> reduce-scatter)."""
    _fields_ = [("owner_base", C.c_void_p), ("world", C.c_int), ("rot", C.c_int), ("prev", C.c_void_p)]


def make_rs_args(owner_base_ptrs, rot: int, prev_ptr: Optional[int] = None):
    """(RsArgs, keep-alive pointer array) for ``gemm(..., rs=...)`` / ``embed_bwd_rs``; ``None`` entries = owners
    without rows.  ``prev_ptr``: this rank's own fp32 accumulator of the same matrix (sent along: owner += acc + prev)."""
    arr = (C.c_void_p * len(owner_base_ptrs))(*[p if p else None for p in owner_base_ptrs])
    return RsArgs(owner_base=C.cast(arr, C.c_void_p), world=len(owner_base_ptrs), rot=rot, prev=prev_ptr), arr


_P, _I, _F, _LL, _U = C.c_void_p, C.c_int, C.c_float, C.c_longlong, C.c_uint
_SIGS = {
    "b200_gemm2_rs_bf16": [C.POINTER(GemmArgs), C.POINTER(RsArgs), _P],
    "b200_sumsq_chunks": [_P, _P, _I, _P, _F, _P],
    "b200_adamw_chunks": [_P, _P, _P, _P, _P, _P, _I, _I, _P, _P, _I, _I, _F, _F, _F, _F, _F, _P, _I, _I, _P],
    "b200_pull_reduce_chunks": [_P, _P, _P, _P, _I, _I, _I, _I, _P],
    "b200_cast_scatter_chunks": [_P, _P, _P, _I, _I, _P, _I, _P],
    "b200_embed_bwd_rs": [_P, _P, _P, _I, _I, _I, _I, _I, _P],
    "b200_gemm2_ag_bf16": [C.POINTER(GemmArgs), C.POINTER(AgArgs), _P],
    "b200_gemm_bf16": [C.POINTER(GemmArgs), _P],
    "b200_gemm2_bf16": [C.POINTER(GemmArgs), _P],
    "b200_embed_fwd": [_P, _P, _P, _I, _I, _I, _P],
    "b200_embed_bwd": [_P, _P, _P, _I, _I, _I, _P],
    "b200_rmsnorm_fwd": [_P, _P, _P, _P, _I, _I, _F, _P],
    "b200_rmsnorm_bwd": [_P, _P, _P, _P, _P, _P, _I, _I, _P],
    "b200_rmsnorm_bwd_dropout": [_P, _P, _P, _P, _P, _P, _I, _I, _P, _F, _P, _U, _P],
    "b200_resid_dropout_fwd": [_P, _P, _P, _I, _I, _F, _P, _U, _P],
    "b200_dropout_bwd_cast": [_P, _P, _I, _I, _F, _P, _U, _P],
    "b200_dropout_mask": [_P, _I, _I, _F, _P, _U, _P],
    "b200_rope_inplace": [_P, _P, _P, _I, _I, _I, _I, _I, _P],
    "b200_cast_f32_bf16": [_P, _P, _LL, _P],
    "b200_swiglu_fwd": [_P, _P, _I, _I, _P],
    "b200_swiglu_bwd": [_P, _P, _P, _I, _I, _P],
    "b200_attn_fwd": [_P, _P, _P, _I, _I, _I, _F, _P, _U, _P],
    "b200_attn_dropout_mask": [_P, _I, _I, _F, _P, _U, _P],
    "b200_attn_bwd": [_P, _P, _P, _P, _P, _P, _P, _I, _I, _I, _F, _P, _U, _P],
    "b200_attn_bwd_rope": [_P, _P, _P, _P, _P, _P, _P, _I, _I, _I, _F, _P, _U, _P, _P, _P],
    "b200_ce_fwd_bwd": [_P, _P, _P, _I, _I, _I, _F, _F, _P],
    "b200_adamw": [_P, _P, _P, _P, _P, _P, _LL, _LL, _I, _F, _F, _F, _F, _F, _P, _P],
    "b200_sumsq": [_P, _LL, _LL, _P, _P],
    "b200_clip_coef": [_P, _F, _F, _P, _P, _P],
    "b200_rng_advance": [_P, C.c_ulonglong, _P],
    "b200_comm_barrier": [_P, _I, _I, _P, _P],
    "b200_reduce_scatter_f32": [_P, _P, _LL, _LL, _F, _P, _I, _I, _I, _P],
    "b200_adamw_bcast": [_P, _P, _P, _P, _P, _I, _I, _P, _LL, _LL, _I, _F, _F, _F, _F, _F, _P, _I, _P],
    "b200_allreduce_scalars": [_P, _I, _I, _P, _P, _P, _I, _P],
    "b200_all_gather_bf16": [_P, _P, _LL, _I, _I, _I, _P],
    "b200_memcpy_async": [_P, _P, C.c_size_t, _P],
    "b200_all_gather_bf16_excl": [_P, _P, _LL, _I, _I, _LL, _LL, _I, _P],
    "b200_reduce_scatter_bf16_accum": [_P, _P, _LL, _LL, _F, _I, _I, _I, _P],
    "b200_reduce_scatter_f32_accum": [_P, _P, _LL, _LL, _F, _I, _I, _I, _P],
}
_fn_cache = {}


def _fn(name: str):
    f = _fn_cache.get(name)
    if f is None:
        f = getattr(_lib.lib(), name)
        f.argtypes = _SIGS[name]
        f.restype = C.c_int
        _fn_cache[name] = f
    return f


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def _call(name: str, *args) -> None:
    global launch_count
    rc = _fn(name)(*args)
    if rc != 0:
        _lib.check(rc, name)
    launch_count += 1


def _ptr(t: Optional[torch.Tensor]) -> Optional[int]:
    return None if t is None else t.data_ptr()


# ----------------------------------------------------------------------------------------------- GEMM
def gemm(a: torch.Tensor, b: torch.Tensor, c: torch.Tensor, *, a_mn: bool = False, b_mn: bool = False,
         epilogue: int = EPI_BF16, split_k: int = 1, tile_n: int = 0, max_ctas: int = 0,
         rope_cos=None, rope_sin=None, seq_len: int = 0, rope_cols: int = 0,
         residual=None, dropout_p: float = 0.0, rng_state=None, rng_stream: int = 0,
         aux=None, half_n: int = 0, M: Optional[int] = None, N: Optional[int] = None, K: Optional[int] = None,
         pair: Optional[bool] = None, ag=None, rs=None, raster: int = 0) -> None:
    """``c = a @ b`` on tcgen05.  ``pair=True`` selects the CTA-pair kernel (``cta_group::2``, 256x256 tiles,
    TMA-store epilogue); ``None`` lets ``use_pair_kernel`` decide for the epilogues that kernel implements.

    ``a``: K-major ``[M, K]`` or (``a_mn``) MN-major ``[K, M]``; ``b``: K-major ``[N, K]`` or (``b_mn``)
    MN-major ``[K, N]``.  Row-major 2-D tensors with unit inner stride; leading dimensions come from
    ``stride(0)`` so column/row slices of packed buffers work without copies.
    """
    assert a.dtype == torch.bfloat16 and b.dtype == torch.bfloat16 and a.stride(1) == 1 and b.stride(1) == 1
    if M is None:
        M = a.shape[1] if a_mn else a.shape[0]
    if K is None:
        K = a.shape[0] if a_mn else a.shape[1]
    if N is None:
        N = b.shape[1] if b_mn else b.shape[0]
    args = GemmArgs(
        A=a.data_ptr(), B=b.data_ptr(), C=c.data_ptr(), M=M, N=N, K=K,
        lda=a.stride(0), ldb=b.stride(0), ldc=c.stride(0), a_mn_major=int(a_mn), b_mn_major=int(b_mn),
        epilogue=epilogue, split_k=split_k, tile_n=tile_n, max_ctas=max_ctas,
        rope_cos=_ptr(rope_cos), rope_sin=_ptr(rope_sin), seq_len=seq_len, rope_cols=rope_cols,
        residual=_ptr(residual), dropout_p=dropout_p, rng_state=_ptr(rng_state), rng_stream=rng_stream,
        aux=_ptr(aux), ld_aux=0 if aux is None else aux.stride(0), half_n=half_n, raster=raster,
    )
    if ag is not None:
        # all-gather fused into the GEMM: ``b`` is the local gathered destination the kernel fills while it multiplies
        # ``epoch``: an int, or a device int64 tensor whose low 32 bits are the epoch (bumped by the caller with
        # rng_advance before the launch: replayable from a CUDA graph)
        shard_ptrs, part_elems, b_offset_elems, flags, epoch, rank, world, gather_pairs = ag
        epoch_dev = epoch.data_ptr() if isinstance(epoch, torch.Tensor) else None
        agargs = AgArgs(shard_ptrs=C.cast(shard_ptrs, C.c_void_p), part_elems=part_elems, b_offset_elems=b_offset_elems,
                        flags=flags.data_ptr(), epoch=0 if epoch_dev else (epoch & 0xFFFFFFFF), rank=rank, world=world,
                        gather_pairs=gather_pairs, epoch_dev=epoch_dev)
        _call("b200_gemm2_ag_bf16", C.byref(args), C.byref(agargs), _stream())
        return
    if rs is not None:
        # reduce-scatter fused into the GEMM: ``rs`` = (RsArgs, keep-alive) from make_rs_args; ``c`` is only a layout
        # template (leading dimension) -- every 128-row slab of the result is reduce-added into its owner's buffer
        _call("b200_gemm2_rs_bf16", C.byref(args), C.byref(rs[0]), _stream())
        return
    if pair is None:
        pair = (use_pair_kernel and M >= 256 and N >= 128 and not (a_mn and epilogue > EPI_F32_ATOMIC)
                and not (epilogue == EPI_SWIGLU and (half_n % 128 or b_mn)))
    _call("b200_gemm2_bf16" if pair else "b200_gemm_bf16", C.byref(args), _stream())


# ----------------------------------------------------------------------------------------------- elementwise
def embed_fwd(ids, weight_bf16, out_f32):
    rows = ids.numel()
    _call("b200_embed_fwd", ids.data_ptr(), weight_bf16.data_ptr(), out_f32.data_ptr(), rows, weight_bf16.shape[1],
          weight_bf16.shape[0], _stream())


def embed_bwd(ids, dx_f32, dweight_f32):
    rows = ids.numel()
    _call("b200_embed_bwd", ids.data_ptr(), dx_f32.data_ptr(), dweight_f32.data_ptr(), rows, dweight_f32.shape[1],
          dweight_f32.shape[0], _stream())


def embed_bwd_rs(ids, dx_f32, rs, vocab: int):
    """``embed_bwd`` whose destination rows live with their owners (``rs`` from make_rs_args over ``vocab`` rows)."""
    rows, hidden = dx_f32.shape
    _call("b200_embed_bwd_rs", ids.data_ptr(), dx_f32.data_ptr(), rs[0].owner_base, rows, hidden, vocab, rs[0].world,
          rs[0].rot, _stream())


def rmsnorm_fwd(x_f32, w_bf16, y_bf16, rstd, eps: float):
    rows, hidden = x_f32.shape
    _call("b200_rmsnorm_fwd", x_f32.data_ptr(), w_bf16.data_ptr(), y_bf16.data_ptr(), rstd.data_ptr(), rows, hidden,
          eps, _stream())


def rmsnorm_bwd(dy_bf16, x_f32, w_bf16, rstd, dx_accum, dw_accum, dy_next=None, p_next: float = 0.0, rng_state=None,
                rng_stream: int = 0):
    """``dx_accum += d/dx``, ``dw_accum += d/dw``; with ``dy_next`` also writes the dropout-masked bf16 copy of
    the updated ``dx_accum`` (the following ``dropout_bwd_cast`` fused into this pass)."""
    rows, hidden = x_f32.shape
    if dy_next is None:
        _call("b200_rmsnorm_bwd", dy_bf16.data_ptr(), x_f32.data_ptr(), w_bf16.data_ptr(), rstd.data_ptr(),
              dx_accum.data_ptr(), dw_accum.data_ptr(), rows, hidden, _stream())
    else:
        _call("b200_rmsnorm_bwd_dropout", dy_bf16.data_ptr(), x_f32.data_ptr(), w_bf16.data_ptr(), rstd.data_ptr(),
              dx_accum.data_ptr(), dw_accum.data_ptr(), rows, hidden, dy_next.data_ptr(), p_next, _ptr(rng_state),
              rng_stream, _stream())


def resid_dropout_fwd(y_bf16, residual, out, p, rng_state, rng_stream):
    rows, cols = y_bf16.shape
    _call("b200_resid_dropout_fwd", y_bf16.data_ptr(), residual.data_ptr(), out.data_ptr(), rows, cols, p,
          _ptr(rng_state), rng_stream, _stream())


def dropout_bwd_cast(dx_f32, dy_bf16, p, rng_state, rng_stream):
    rows, cols = dx_f32.shape
    _call("b200_dropout_bwd_cast", dx_f32.data_ptr(), dy_bf16.data_ptr(), rows, cols, p, _ptr(rng_state), rng_stream,
          _stream())


def dropout_mask(mask_u8, p, rng_state, rng_stream):
    rows, cols = mask_u8.shape
    _call("b200_dropout_mask", mask_u8.data_ptr(), rows, cols, p, _ptr(rng_state), rng_stream, _stream())


def rope_inplace(qkv_bf16, cos, sin, rope_cols, seq_len, inverse=False):
    rows = qkv_bf16.shape[0]
    _call("b200_rope_inplace", qkv_bf16.data_ptr(), cos.data_ptr(), sin.data_ptr(), rows, qkv_bf16.stride(0),
          rope_cols, seq_len, int(inverse), _stream())


def cast_f32_bf16(src_f32, dst_bf16):
    _call("b200_cast_f32_bf16", src_f32.data_ptr(), dst_bf16.data_ptr(), src_f32.numel(), _stream())


def swiglu_fwd(gu_bf16, act_bf16):
    rows, inter = act_bf16.shape
    _call("b200_swiglu_fwd", gu_bf16.data_ptr(), act_bf16.data_ptr(), rows, inter, _stream())


def swiglu_bwd(dact_bf16, gu_bf16, dgu_bf16):
    rows, inter = dact_bf16.shape
    _call("b200_swiglu_bwd", dact_bf16.data_ptr(), gu_bf16.data_ptr(), dgu_bf16.data_ptr(), rows, inter, _stream())


# ----------------------------------------------------------------------------------------------- attention
def attn_fwd(qkv_bf16, out_bf16, lse, batch, seq, heads, dropout_p, rng_state, rng_stream):
    _call("b200_attn_fwd", qkv_bf16.data_ptr(), out_bf16.data_ptr(), lse.data_ptr(), batch, seq, heads, dropout_p,
          _ptr(rng_state), rng_stream, _stream())


def attn_dropout_mask(mask_u8, p, rng_state, rng_stream):
    """keep-mask ``[B*H*S, S]`` (uint8) exactly as the attention kernels draw it (tests / oracles)."""
    rows, cols = mask_u8.shape
    _call("b200_attn_dropout_mask", mask_u8.data_ptr(), rows, cols, p, _ptr(rng_state), rng_stream, _stream())


def attn_bwd(qkv_bf16, out_bf16, dout_bf16, lse, dqkv_bf16, delta, dq_accum, batch, seq, heads, dropout_p, rng_state,
             rng_stream, rope_cos=None, rope_sin=None):
    """With rotary tables the dQ / dK columns of ``dqkv`` come out rotated back (inverse RoPE fused in)."""
    if rope_cos is None:
        _call("b200_attn_bwd", qkv_bf16.data_ptr(), out_bf16.data_ptr(), dout_bf16.data_ptr(), lse.data_ptr(),
              dqkv_bf16.data_ptr(), delta.data_ptr(), dq_accum.data_ptr(), batch, seq, heads, dropout_p, _ptr(rng_state),
              rng_stream, _stream())
    else:
        _call("b200_attn_bwd_rope", qkv_bf16.data_ptr(), out_bf16.data_ptr(), dout_bf16.data_ptr(), lse.data_ptr(),
              dqkv_bf16.data_ptr(), delta.data_ptr(), dq_accum.data_ptr(), batch, seq, heads, dropout_p, _ptr(rng_state),
              rng_stream, rope_cos.data_ptr(), rope_sin.data_ptr(), _stream())


# ----------------------------------------------------------------------------------------------- loss / optimizer
def ce_fwd_bwd(logits_bf16, labels, loss_sum, vocab, loss_scale, grad_scale):
    rows = logits_bf16.shape[0]
    _call("b200_ce_fwd_bwd", logits_bf16.data_ptr(), labels.data_ptr(), loss_sum.data_ptr(), rows, vocab,
          logits_bf16.stride(0), loss_scale, grad_scale, _stream())


def adamw_step(master, grad, exp_avg, exp_avg_sq, shadow, decay_chunks, lo, hi, step, lr, beta1, beta2, eps, wd,
               grad_scale):
    _call("b200_adamw", master.data_ptr(), grad.data_ptr(), exp_avg.data_ptr(), exp_avg_sq.data_ptr(), _ptr(shadow),
          decay_chunks.data_ptr(), lo, hi, step, lr, beta1, beta2, eps, wd, _ptr(grad_scale), _stream())


def adamw_step_ptr(master_ptr, grad_ptr, m_ptr, v_ptr, shadow_ptr, decay_ptr, lo, hi, step, lr, beta1, beta2, eps, wd,
                   grad_scale_ptr):
    """Raw-pointer form: sharded engines index ``grad``/``shadow`` through pre-offset base pointers."""
    _call("b200_adamw", master_ptr, grad_ptr, m_ptr, v_ptr, shadow_ptr, decay_ptr, lo, hi, step, lr, beta1, beta2,
          eps, wd, grad_scale_ptr, _stream())


def sumsq(x_f32, lo, hi, out):
    _call("b200_sumsq", x_f32.data_ptr(), lo, hi, out.data_ptr(), _stream())


def clip_coef(sumsq_t, max_norm, pre_scale, scale_out, norm_out=None):
    _call("b200_clip_coef", sumsq_t.data_ptr(), max_norm, pre_scale, scale_out.data_ptr(), _ptr(norm_out), _stream())


def rng_advance(rng_state, delta=1):
    _call("b200_rng_advance", rng_state.data_ptr(), delta, _stream())


# ----------------------------------------------------------------------------------------------- NVLink comm
def ptr_array(ptrs):
    """Host array of device pointers (one per rank) in the form the comm entry points take."""
    return (C.c_void_p * len(ptrs))(*ptrs)


def comm_barrier(pad_ptrs, rank, world, epoch_counter):
    _call("b200_comm_barrier", pad_ptrs, rank, world, epoch_counter.data_ptr(), _stream())


def reduce_scatter_f32(grad_ptrs, dst, lo, hi, scale, sumsq, rank, world, max_ctas=0):
    _call("b200_reduce_scatter_f32", grad_ptrs, dst.data_ptr(), lo, hi, scale, _ptr(sumsq), rank, world, max_ctas,
          _stream())


def adamw_bcast(master, grad, exp_avg, exp_avg_sq, shadow_ptrs, rank, world, decay_chunks, lo, hi, step, lr, beta1,
                beta2, eps, wd, grad_scale, max_ctas=0):
    _call("b200_adamw_bcast", master.data_ptr(), grad.data_ptr(), exp_avg.data_ptr(), exp_avg_sq.data_ptr(),
          shadow_ptrs, rank, world, decay_chunks.data_ptr(), lo, hi, step, lr, beta1, beta2, eps, wd,
          _ptr(grad_scale), max_ctas, _stream())


def allreduce_scalars(pad_ptrs, rank, world, epoch_counter, inp, out, n):
    _call("b200_allreduce_scalars", pad_ptrs, rank, world, epoch_counter.data_ptr(), inp.data_ptr(), out.data_ptr(), n,
          _stream())


CHUNK_ELEMS = 4096


def make_chunk_table(runs, device) -> "tuple[torch.Tensor, int]":
    """Device table of ``b200_chunk`` from runs ``(g, pg, mv, sh, n, dst)`` (element offsets; n % 4 == 0), each split
    into pieces of at most CHUNK_ELEMS elements.  Returns (uint8 tensor holding the table, number of chunks)."""
    import numpy as np

    dt = np.dtype([("g", "<i8"), ("pg", "<i8"), ("mv", "<i8"), ("sh", "<i8"), ("n", "<i4"), ("dst", "<i4")])
    assert dt.itemsize == 40
    rows = []
    for g, pg, mv, sh, n, dst in runs:
        assert n % 4 == 0, "chunk runs must be multiples of 4 elements"
        o = 0
        while o < n:
            m = min(CHUNK_ELEMS, n - o)
            rows.append((g + o, pg + o, mv + o, sh + o, m, dst))
            o += m
    arr = np.array(rows, dtype=dt) if rows else np.zeros(0, dtype=dt)
    t = torch.from_numpy(arr.view(np.uint8).copy()) if rows else torch.zeros(0, dtype=torch.uint8)
    return t.to(device), len(rows)


def sumsq_chunks(grad_f32, table, n_chunks, out, sq_scale: float = 1.0):
    _call("b200_sumsq_chunks", grad_f32.data_ptr(), table.data_ptr(), n_chunks, out.data_ptr(), sq_scale, _stream())


def adamw_chunks(master, grad, exp_avg, exp_avg_sq, param_ptrs, rank, world, decay_chunks, table, n_chunks, step, lr,
                 beta1, beta2, eps, wd, grad_scale, zero_grad: bool = True, max_ctas: int = 0, multicast_ptr: int = 0):
    """``multicast_ptr``: NVLS multicast address of the parameter buffer (``MulticastBuffer.mc_ptr``); chunks with
    dst = -1 are then broadcast by one ``multimem.st`` instead of one store per rank."""
    _call("b200_adamw_chunks", master.data_ptr(), grad.data_ptr(), exp_avg.data_ptr(), exp_avg_sq.data_ptr(), param_ptrs,
          multicast_ptr or None, rank, world, decay_chunks.data_ptr(), table.data_ptr(), n_chunks, step, lr, beta1, beta2, eps, wd,
          _ptr(grad_scale), int(zero_grad), max_ctas, _stream())


def cast_scatter_chunks(master, grad_to_zero, param_ptrs, rank, world, table, n_chunks):
    _call("b200_cast_scatter_chunks", master.data_ptr(), _ptr(grad_to_zero), param_ptrs, rank, world, table.data_ptr(),
          n_chunks, _stream())


def pull_reduce_chunks(src_ptrs, dst_f32, table, n_chunks, rank, world, zero_remote: bool = True, multicast_ptr: int = 0):
    _call("b200_pull_reduce_chunks", src_ptrs, multicast_ptr or None, dst_f32.data_ptr(), table.data_ptr(), n_chunks, rank,
          world, int(zero_remote), _stream())


def memcpy_async(dst_ptr: int, src_ptr: int, nbytes: int) -> None:
    """Copy-engine transfer on the current stream between raw device addresses (peer mappings included)."""
    rc = _fn("b200_memcpy_async")(dst_ptr, src_ptr, C.c_size_t(nbytes), _stream())   # not a kernel: not counted
    if rc != 0:
        _lib.check(rc, "b200_memcpy_async")


def all_gather_bf16(shard_ptrs, dst, shard_elems, rank, world, max_ctas=0, exclude=(0, 0)):
    """``exclude=(lo, hi)``: elements of ``dst`` to leave alone (filled by a fused all-gather GEMM instead)."""
    _call("b200_all_gather_bf16_excl", shard_ptrs, dst.data_ptr(), shard_elems, rank, world, exclude[0], exclude[1],
          max_ctas, _stream())


def reduce_scatter_bf16_accum(full_grad_ptrs, dst_shard, src_offset, n, scale, rank, world, max_ctas=0):
    _call("b200_reduce_scatter_bf16_accum", full_grad_ptrs, dst_shard.data_ptr(), src_offset, n, scale, rank, world,
          max_ctas, _stream())


def reduce_scatter_f32_accum(full_grad_ptrs, dst_shard, src_offset, n, scale, rank, world, max_ctas=0):
    _call("b200_reduce_scatter_f32_accum", full_grad_ptrs, dst_shard.data_ptr(), src_offset, n, scale, rank, world,
          max_ctas, _stream())
