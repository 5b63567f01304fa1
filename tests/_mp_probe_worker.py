"""Worker for the 128-byte-line probe: ranks pair up (0-1, 2-3, ...) and play ping-pong over NVLink."""
import os
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))

import torch  # noqa: E402

from distributed_neural_network_b200.trainer import TrainConfig, Trainer  # noqa: E402


def main():
    out_dir, iters, n_ctas = sys.argv[1], int(sys.argv[2]), int(sys.argv[3])
    os.makedirs(out_dir, exist_ok=True)   # every rank writes its result here; nobody may rely on another rank's mkdir
    tr = Trainer(TrainConfig(device="cuda", synthetic=True, quiet=True, sync_mode="step-allreduce", seed=3,
                             peer_timeout_s=20.0, log_dir=out_dir + "/log", calibrate_exchange=False))
    res = {"rank": tr.rank, "world": tr.world}
    peer = tr.rank ^ 1
    if peer < tr.world:
        tr.comm.barrier()
        res["probe"] = tr.engine.ll128_probe(peer, iters=iters, n_ctas=n_ctas)
        res["ll16_one_way_us"] = tr.engine.pingpong(peer, iters=2000)
    tr.comm.barrier()
    torch.save(res, f"{out_dir}/rank{tr.rank}.pt")
    tr.close()
    tr.comm.close()


if __name__ == "__main__":
    main()
