"""NVLink SHARP (NVLS): the symmetric buffer that every rank binds into ONE multicast object.

The multicast address lets a kernel store once and have NVSwitch deliver to all GPUs (``multimem.st``), or load
once and have the switch return the SUM of all GPUs' copies (``multimem.ld_reduce``); the exchange variant built on
them is csrc/exchange.cuh (variant 2). The object is made by the runtime's own group code (csrc/nvls_group.cu:
``cuMulticastCreate`` on rank 0, the handle passed to the other processes as a file descriptor over a Unix socket,
``cuMulticastAddDevice`` / ``cuMemCreate`` / ``cuMulticastBindMem`` / two mappings on every rank); this module is the
collective choreography around it -- the two barriers the driver's rules demand and the all-or-nobody decision -- on
whatever control plane the job has (mpi4py, gloo or NCCL process group: ``bootstrap.Comm``). Availability is probed
at run time (north star: "an NVLS-multicast path where the host exposes it"); when anything is missing every rank
keeps the peer-memory variants.
No counterpart in the reference, whose only collective is the rank-0 star of data_parallelism_train.py:226-244.
"""
from __future__ import annotations

import os
import socket
import time
from typing import Optional, Tuple

import torch

from ..ops import native

last_error: str = ""      # why the last try_create on this rank returned None ("" = it did not, or another rank failed)


def try_create(device: torch.device, comm, timeout_s: float = 10.0) -> Optional[Tuple["native.NvlsGroup", int]]:
    """(group, multicast pointer) or None. Collective: every rank of ``comm`` must call it."""
    global last_error
    last_error = ""
    world, rank = comm.world, comm.rank
    idx = device.index if device.index is not None else torch.cuda.current_device()
    mine = (socket.gethostname(), bool(world >= 2 and device.type == "cuda" and native.nvls_supported(idx)))
    seen = comm.allgather(mine)
    if not all(ok for _, ok in seen) or len({h for h, _ in seen}) != 1:   # descriptors do not cross hosts
        last_error = "no multicast support on this device/driver" if not mine[1] else "ranks on several hosts, or a rank without support"
        return None
    name = comm.bcast(f"{os.getpid()}-{time.time_ns():x}" if rank == 0 else None, 0)
    group = None
    try:
        group = native.NvlsGroup(rank, world, idx, native.NVLS_FLOATS * 4, name, timeout_s)
    except RuntimeError as e:
        group, last_error = None, str(e)
    if not all(comm.allgather(group is not None)):       # also the barrier "every device has been added"
        if group is not None:
            group.close()
        return None
    try:
        group.map()
        mapped = True
    except RuntimeError as e:
        mapped, last_error = False, str(e)
    if not all(comm.allgather(mapped)):                  # also the barrier "every rank has bound its memory"
        group.close()
        return None
    return group, group.multicast_ptr
