"""Device-timed latency of the stand-alone peer-memory all-reduce per message size and variant (incl. NVLS), max over
ranks. Each cell is a burst of back-to-back calls enqueued from C between two CUDA events (the round-1 table timed a
Python launch loop and measured the interpreter).
    bin/mpiexec -n N python bench/allreduce_latency.py [iters] > profiles/allreduce_latency_nN.json"""
import json
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from distributed_neural_network_b200.ops import flat as L  # noqa: E402
from distributed_neural_network_b200.parallel import nvls  # noqa: E402
from distributed_neural_network_b200.trainer import TrainConfig, Trainer  # noqa: E402


def main():
    iters = int(sys.argv[1]) if len(sys.argv) > 1 else 400
    tr = Trainer(TrainConfig(device="cuda", synthetic=True, quiet=True, sync_mode="step-allreduce", seed=1,
                             peer_timeout_s=30.0))
    eng, world = tr.engine, tr.world
    tr._align_sequence()
    src = torch.randn(L.FLAT_NUMEL, device="cuda")
    dst = torch.empty(L.FLAT_NUMEL, device="cuda")
    variants = ["one-shot", "one-shot-128", "two-shot"] + (["nvls"] if tr.nvls_ok else [])
    table = {}
    for n in (1024, 4096, 16384, L.FLAT_NUMEL):
        row = {}
        for v in variants:
            us = eng.allreduce_timed(src[:n], dst[:n], v, iters)
            row[v] = round(max(tr.comm.allgather(us)), 2)
        table[f"{n * 4}_bytes"] = row
    st = int(eng.read_ctl().status)
    if tr.rank == 0:
        print(json.dumps({"world": world, "iters": iters, "unit": "us per all-reduce, device-timed burst, max over ranks",
                          "nvls": tr.nvls_ok, "nvls_why_not": nvls.last_error, "peer_timeouts": st, "table": table},
                         indent=1))
    tr.close()
    tr.comm.close()


if __name__ == "__main__":
    main()
