"""Does this box expose NVLink SHARP / multicast (NVLS) to user code? Run with bin/mpiexec or torchrun on >= 2 GPUs.
Reports (1) the driver attribute CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED as the runtime reads it, (2) whether the
runtime's own multicast group (csrc/nvls_group.cu) can be built across the ranks of the job, and why not if not.
Prints one JSON object on rank 0. (What NCCL does itself: NCCL_DEBUG=INFO | grep NVLS.)"""
import json
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from distributed_neural_network_b200.ops import native  # noqa: E402
from distributed_neural_network_b200.parallel import bootstrap, nvls  # noqa: E402


def main():
    comm = bootstrap.init()
    dev = torch.device("cuda", comm.local_rank % torch.cuda.device_count())
    torch.cuda.set_device(dev)
    mine = {"rank": comm.rank, "device": dev.index, "multicast_supported": native.nvls_supported(dev.index)}
    made = nvls.try_create(dev, comm) if comm.world > 1 else None
    mine["group"] = made is not None
    mine["why_not"] = nvls.last_error
    if made is not None:
        mine["bytes"] = made[0].nbytes
    everyone = comm.allgather(mine)
    if comm.rank == 0:
        print("NVLS_PROBE " + json.dumps({"world": comm.world, "ranks": everyone}))
    comm.barrier()
    if made is not None:
        made[0].close()
    comm.close()


if __name__ == "__main__":
    main()
