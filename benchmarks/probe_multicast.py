"""Is NVLS multicast (multimem.* through a multicast mapping) available to this process model on the box?
Probes torch.distributed._symmetric_memory (CUDA driver multicast objects shared between the ranks' processes).

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port P benchmarks/probe_multicast.py
"""
import json
import os

import torch
import torch.distributed as dist

local = int(os.environ.get("LOCAL_RANK", 0))
dev = torch.device(f"cuda:{local}")
torch.cuda.set_device(dev)
dist.init_process_group("nccl", device_id=dev)
rank, world = dist.get_rank(), dist.get_world_size()
out = {"world": world}
try:
    import torch.distributed._symmetric_memory as symm

    out["has_multicast_support"] = bool(symm.has_multicast_support(dev.type, dev.index)) if hasattr(symm, "has_multicast_support") else None
    t = symm.empty(1 << 20, dtype=torch.float32, device=dev)
    hdl = symm.rendezvous(t, dist.group.WORLD.group_name)
    out["buffer_ptrs"] = [hex(p) for p in hdl.buffer_ptrs]
    out["signal_pad_ptrs"] = len(hdl.signal_pad_ptrs)
    out["multicast_ptr"] = hex(hdl.multicast_ptr) if getattr(hdl, "multicast_ptr", 0) else 0
    out["buffer_size"] = hdl.buffer_size
except Exception as e:  # noqa: BLE001
    out["error"] = f"{type(e).__name__}: {e}"[:400]
import ctypes

try:
    cu = ctypes.CDLL("libcuda.so.1")
    v = ctypes.c_int()
    cu.cuInit(0)
    d = ctypes.c_int()
    cu.cuDeviceGet(ctypes.byref(d), local)
    CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED = 132
    cu.cuDeviceGetAttribute(ctypes.byref(v), CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, d)
    out["CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED"] = v.value
    CU_DEVICE_ATTRIBUTE_HANDLE_TYPE_FABRIC_SUPPORTED = 128
    cu.cuDeviceGetAttribute(ctypes.byref(v), CU_DEVICE_ATTRIBUTE_HANDLE_TYPE_FABRIC_SUPPORTED, d)
    out["HANDLE_TYPE_FABRIC_SUPPORTED"] = v.value
except Exception as e:  # noqa: BLE001
    out["driver_probe_error"] = str(e)[:200]
if rank == 0:
    print(json.dumps(out))
dist.barrier()
dist.destroy_process_group()
