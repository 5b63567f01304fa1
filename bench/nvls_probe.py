"""Does this box expose NVLink SHARP / multicast (NVLS) to user code? Run with torchrun on >= 2 GPUs.
Probes (1) the driver attribute CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, (2) torch symmetric memory's multicast
pointer, (3) what NCCL reports (NCCL_DEBUG=INFO | grep NVLS). Prints one JSON object on rank 0."""
import ctypes
import json
import os

import torch
import torch.distributed as dist


def main():
    rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    out = {"world": world}
    try:
        cu = ctypes.CDLL("libcuda.so.1")
        v = ctypes.c_int(-1)
        cu.cuInit(0)
        rc = cu.cuDeviceGetAttribute(ctypes.byref(v), 132, local)   # CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED
        out["cu_multicast_supported"] = {"rc": rc, "value": v.value}
        v2 = ctypes.c_int(-1)
        rc = cu.cuDeviceGetAttribute(ctypes.byref(v2), 128, local)  # HANDLE_TYPE_FABRIC_SUPPORTED
        out["cu_fabric_handle_supported"] = {"rc": rc, "value": v2.value}
    except Exception as e:
        out["driver_probe_error"] = str(e)
    if world > 1:
        dist.init_process_group("nccl", rank=rank, world_size=world)
        try:
            import torch.distributed._symmetric_memory as symm_mem

            t = symm_mem.empty(4096, dtype=torch.float32, device="cuda")
            hdl = symm_mem.rendezvous(t, dist.group.WORLD)
            out["symm_mem_multicast_ptr"] = int(getattr(hdl, "multicast_ptr", 0) or 0)
            out["symm_mem_buffer_ptrs"] = len(getattr(hdl, "buffer_ptrs", []))
        except Exception as e:
            out["symm_mem_error"] = str(e)[:300]
        x = torch.ones(1 << 20, device="cuda")
        dist.all_reduce(x)
        torch.cuda.synchronize()
        dist.barrier()
    if rank == 0:
        print("NVLS_PROBE " + json.dumps(out))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
