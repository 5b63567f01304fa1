"""Worker for the dead-rank test: rank 1 exits without a word after the first barrier; rank 0 goes on to the next."""
import os
import sys
import time
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))

from distributed_neural_network_b200.parallel import bootstrap  # noqa: E402


def main():
    out = sys.argv[1]
    comm = bootstrap.init(prefer_mpi=False)
    comm.barrier()
    if comm.rank == 1:
        os._exit(0)                      # no destroy_process_group, no goodbye: a crash as the others see it
    t0 = time.time()
    try:
        comm.allgather("still here")
        verdict = "completed?!"
    except bootstrap.PeerLost as e:
        verdict = f"PeerLost after {time.time() - t0:.1f}s: {e}"
    with open(out, "w") as f:
        f.write(verdict)
    os._exit(0)


if __name__ == "__main__":
    main()
