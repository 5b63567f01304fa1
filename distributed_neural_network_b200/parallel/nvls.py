"""NVLink SHARP (NVLS) plumbing: a symmetric buffer that every rank maps into ONE multicast object.

The multicast address lets a kernel store once and have NVSwitch deliver to all GPUs (``multimem.st``), or load
once and have the switch return the SUM of all GPUs' copies (``multimem.ld_reduce``). Creating the multicast object
needs the CUDA VMM API and a file-descriptor exchange between processes; ``torch.distributed._symmetric_memory``
already does exactly that, so it is used for the allocation and rendezvous only -- the kernels that use the
addresses are ours (csrc/allreduce.cu). Availability is probed at run time (north star: "an NVLS-multicast path
where the host exposes it"); when anything is missing the caller keeps the peer-memory variants.
No counterpart in the reference, whose only collective is the rank-0 star of data_parallelism_train.py:226-244.
"""
from __future__ import annotations

from typing import Optional, Tuple

import torch

from ..ops import native


def try_create(device: torch.device) -> Optional[Tuple[torch.Tensor, int]]:
    """(local buffer, multicast pointer) or None. Collective: every rank of the default group must call it."""
    try:
        import torch.distributed as dist
        import torch.distributed._symmetric_memory as symm_mem

        if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() < 2:
            return None
        if "nccl" not in str(dist.get_backend()):
            return None
        buf = symm_mem.empty(native.NVLS_FLOATS, dtype=torch.float32, device=device)
        hdl = symm_mem.rendezvous(buf, dist.group.WORLD)
        mc = int(getattr(hdl, "multicast_ptr", 0) or 0)
        if mc == 0:
            return None
        buf.zero_()
        torch.cuda.synchronize(device)
        dist.barrier()
        return buf, mc
    except Exception:
        return None
