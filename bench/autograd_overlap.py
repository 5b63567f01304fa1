"""Layer-at-a-time training step at a large per-GPU batch: per-bucket gradient exchange on a side stream overlapped
with backward (parallel/bucketed.py) against the same three all-reduces enqueued after backward. Run under torchrun.

Prints, per batch size, microseconds per step (device-timed, max over ranks) for: overlap, no overlap, and no
exchange at all (what a perfectly hidden exchange would cost)."""
import json
import os
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from distributed_neural_network_b200.ops import functional as Fn  # noqa: E402
from distributed_neural_network_b200.parallel.bucketed import BucketedGradientAverager  # noqa: E402
from distributed_neural_network_b200.trainer import TrainConfig, Trainer  # noqa: E402


def main():
    world = int(os.environ.get("WORLD_SIZE", "1"))
    tr = Trainer(TrainConfig(batch_size=16, sync_mode="step-allreduce" if world > 1 else "none", synthetic=True, quiet=True,
                             seed=1, device="cuda", peer_timeout_s=60.0, calibrate_exchange=False))
    model, eng = tr.model, tr.engine
    out = {"world": world}
    for B in (256, 1024, 4096):
        x = torch.randn(B, 3, 32, 32, device="cuda")
        y = torch.randint(0, 10, (B,), device="cuda")
        res = {}
        for mode in ("overlap", "after-backward", "no-exchange"):
            avg = None
            if world > 1 and mode != "no-exchange":
                tr._align_sequence()
                tr._apply_membership()
                avg = BucketedGradientAverager(model, world, engine=eng, variant=tr._standalone_variant(),
                                               overlap=mode == "overlap")

            def step():
                model.zero_grad()
                loss = Fn.cross_entropy(model(x), y)
                if avg is not None:
                    avg.arm()
                loss.backward()
                if avg is not None:
                    avg.finish()
                Fn.sgd_momentum_(model.flat_params.data, model.flat_grads, tr.momentum, 1e-4, 0.9, False)

            for _ in range(10):
                step()
            torch.cuda.synchronize()
            tr.comm.barrier()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            for _ in range(50):
                step()
            b.record()
            torch.cuda.synchronize()
            us = max(tr.comm.allgather(a.elapsed_time(b) * 1e3 / 50))
            res[mode] = round(us, 1)
            if avg is not None:
                avg.close()
            tr.comm.barrier()
        out[f"B{B}_us_per_step"] = res
    if tr.rank == 0:
        print(json.dumps(out))
        Path("gpurun_out").mkdir(exist_ok=True)
        Path(f"gpurun_out/autograd_overlap_n{world}.json").write_text(json.dumps(out, indent=1))
    tr.close()
    tr.comm.close()


if __name__ == "__main__":
    main()
