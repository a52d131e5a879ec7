"""NVLink communication runtime: peer-mapped ("symmetric") buffers, pointer tables, signal pads.

Replaces what NCCL provides underneath torch DDP / FSDP in the reference (communicator init
ddp_trainer.py:99, fsdp_trainer.py:128) for the hot collectives: every rank allocates a
buffer with the library's raw allocator, exports a CUDA IPC handle, and maps every peer's buffer into
its own address space, after which kernels load / store peer memory directly over NVLink
(``csrc/comm.cu``).  ``torch.distributed`` (NCCL) is used only for the bootstrap exchange of the
64-byte handles and for cold paths (checkpoint gathers).
"""
from __future__ import annotations

import ctypes as C
from typing import List, Optional

import torch
import torch.distributed as dist

from ..ops import _lib
from ..ops import kernels as K

_DTYPE_STR = {torch.float32: "<f4", torch.int32: "<i4", torch.uint8: "|u1", torch.int16: "<i2"}


class _RawCudaArray:
    """Minimal ``__cuda_array_interface__`` carrier so torch can wrap a raw device pointer zero-copy."""

    def __init__(self, ptr: int, numel: int, typestr: str):
        self.__cuda_array_interface__ = {"shape": (numel,), "typestr": typestr, "data": (ptr, False), "version": 2,
                                         "strides": None}


def wrap_pointer(ptr: int, numel: int, dtype: torch.dtype, device: torch.device) -> torch.Tensor:
    if dtype == torch.bfloat16:
        return torch.as_tensor(_RawCudaArray(ptr, numel, "<i2"), device=device).view(torch.bfloat16)
    return torch.as_tensor(_RawCudaArray(ptr, numel, _DTYPE_STR[dtype]), device=device)


class SymmetricBuffer:
    """One allocation per rank, all of them mapped into every rank.  ``local`` is a torch view of this
    rank's copy; ``ptrs`` is the host pointer table (index = rank) the comm entry points take."""

    def __init__(self, numel: int, dtype: torch.dtype, device: torch.device, group=None):
        lib = _lib.lib()
        self.group = group
        self.rank = dist.get_rank(group)
        self.world = dist.get_world_size(group)
        self.numel, self.dtype, self.device = numel, dtype, device
        itemsize = torch.empty((), dtype=dtype).element_size()
        self.nbytes = (numel * itemsize + 255) // 256 * 256
        p = C.c_void_p()
        _lib.check(lib.b200_device_malloc(C.byref(p), C.c_size_t(self.nbytes)), "b200_device_malloc")
        self._own_ptr = p.value
        handle = (C.c_ubyte * 64)()
        _lib.check(lib.b200_ipc_export(C.c_void_p(self._own_ptr), handle), "b200_ipc_export")
        handles: List[Optional[bytes]] = [None] * self.world
        dist.all_gather_object(handles, bytes(handle), group=group)
        raw: List[int] = []
        self._opened: List[int] = []
        for r, h in enumerate(handles):
            if r == self.rank:
                raw.append(self._own_ptr)
                continue
            q = C.c_void_p()
            buf = (C.c_ubyte * 64).from_buffer_copy(h)
            _lib.check(lib.b200_ipc_open(buf, C.byref(q)), f"b200_ipc_open(rank {r})")
            raw.append(q.value)
            self._opened.append(q.value)
        self.raw_ptrs = raw
        self.ptrs = K.ptr_array(raw)
        self.local = wrap_pointer(self._own_ptr, numel, dtype, device)
        dist.barrier(group=group)

    def offset_ptrs(self, byte_offset: int):
        return K.ptr_array([p + byte_offset for p in self.raw_ptrs])

    def close(self) -> None:
        """Collective: every rank unmaps its peers' allocations, then (after a barrier) frees its own."""
        lib = _lib.lib()
        torch.cuda.synchronize()
        for p in self._opened:
            lib.b200_ipc_close(C.c_void_p(p))
        self._opened = []
        if dist.is_initialized() and self.world > 1:
            dist.barrier(group=self.group)
        if self._own_ptr:
            lib.b200_device_free(C.c_void_p(self._own_ptr))
            self._own_ptr = 0


class MulticastBuffer:
    """Symmetric buffer that ALSO has an NVLS multicast mapping: one ``multimem.st`` to ``mc_ptr + offset`` is
    replicated by the NVSwitch into every rank's copy (reference: the broadcast half of DDP's all-reduce,
    ddp_trainer.py:167-172 -- here one store per element leaves the GPU instead of world-1).  The driver objects
    (cuMemCreate / cuMulticastCreate / cuMulticastBindMem, handle exchange between the ranks' processes) come from
    ``torch.distributed._symmetric_memory``; the kernels that use the mapping are ours (``csrc/comm.cu``).
    Same surface as ``SymmetricBuffer``; ``available()`` says whether this process model can have one."""

    @staticmethod
    def available() -> bool:
        import os

        if os.environ.get("B200_NVLS", "1") == "0":
            return False
        try:
            import torch.distributed._symmetric_memory as symm  # noqa: F401
        except Exception:
            return False
        return True

    def __init__(self, numel: int, dtype: torch.dtype, device: torch.device, group=None):
        import torch.distributed._symmetric_memory as symm

        self.group = group
        self.rank = dist.get_rank(group)
        self.world = dist.get_world_size(group)
        self.numel, self.dtype, self.device = numel, dtype, device
        self._t = symm.empty(numel, dtype=dtype, device=device)
        g = group if group is not None else dist.group.WORLD
        self._hdl = symm.rendezvous(self._t, g.group_name)
        self.mc_ptr = int(getattr(self._hdl, "multicast_ptr", 0) or 0)
        if self.mc_ptr == 0:
            raise RuntimeError("symmetric memory came without a multicast mapping")
        self._t.zero_()
        self.raw_ptrs = [int(p) for p in self._hdl.buffer_ptrs]
        self.ptrs = K.ptr_array(self.raw_ptrs)
        self.local = self._t
        torch.cuda.synchronize()
        dist.barrier(group=group)

    def offset_ptrs(self, byte_offset: int):
        return K.ptr_array([p + byte_offset for p in self.raw_ptrs])

    def close(self) -> None:
        torch.cuda.synchronize()
        if dist.is_initialized() and self.world > 1:
            dist.barrier(group=self.group)
        self._hdl = None
        self._t = None
        self.local = None


class SignalPads:
    """Per-rank flag area + the epoch counter of pad operations.

    The counter lives in device memory and is bumped by the pad kernels themselves, so barrier /
    scalar-all-reduce launches carry no changing host argument and can sit inside a captured CUDA graph.
    Invariant: all ranks issue the same sequence of pad operations, and never two of them in flight at once.
    The second half is enforced here: a pad operation issued on a different stream than the previous one is
    ordered behind it with an event (two concurrent ones would both bump the epoch counter and deadlock every
    rank until the 20 s device-side timeout); inside a stream capture the check raises instead."""

    def __init__(self, device: torch.device, group=None):
        nbytes = _lib.lib().b200_comm_pad_bytes()
        self.buf = SymmetricBuffer(nbytes // 4, torch.int32, device, group)
        self.rank, self.world = self.buf.rank, self.buf.world
        self.counter = torch.zeros(1, dtype=torch.int32, device=device)
        self._last_stream: Optional[int] = None
        self._last_event: Optional[torch.cuda.Event] = None

    def _serialize(self) -> None:
        cur = torch.cuda.current_stream()
        if self._last_stream is not None and self._last_stream != cur.cuda_stream:
            if torch.cuda.is_current_stream_capturing():
                raise RuntimeError("SignalPads: pad operation captured on a different stream than the previous one")
            if self._last_event is not None:
                cur.wait_event(self._last_event)
        self._last_stream = cur.cuda_stream

    def _mark(self) -> None:
        if not torch.cuda.is_current_stream_capturing():
            self._last_event = torch.cuda.Event()
            self._last_event.record(torch.cuda.current_stream())

    def barrier(self) -> None:
        """Device-side barrier on the current stream (no host sync)."""
        self._serialize()
        K.comm_barrier(self.buf.ptrs, self.rank, self.world, self.counter)
        self._mark()

    def allreduce_scalars(self, inp: torch.Tensor, out: torch.Tensor, n: int) -> None:
        self._serialize()
        K.allreduce_scalars(self.buf.ptrs, self.rank, self.world, self.counter, inp, out, n)
        self._mark()
