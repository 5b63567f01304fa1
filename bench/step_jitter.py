"""Per-step timeline of the last 1024 training steps on every rank (ring mode of the update-kernel trace):
where does the step time of a multi-GPU run go -- compute jitter between ranks or the exchange itself?

Run under torchrun; every rank writes gpurun_out/jitter_w{world}_r{rank}.npy with one row per step:
[dependency wait passed, last local gradient done, last exchange done, first CTA entered, SGD stored, fence
passed, counter bumped] (ns, that GPU's globaltimer). Environment: DNN_DEBUG_FLAGS (engine A/B switches), DNN_EXCHANGE
(one-shot | two-shot), DNN_CLUSTER (cluster size of the forward/backward kernel).
"""
import os
import sys
from pathlib import Path

import numpy as np
import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from distributed_neural_network_b200.trainer import TrainConfig, Trainer  # noqa: E402


def main():
    B = 16
    world = int(os.environ.get("WORLD_SIZE", "1"))
    tr = Trainer(TrainConfig(batch_size=B, sync_mode="step-allreduce" if world > 1 else "none", synthetic=True, quiet=True,
                             seed=1, device="cuda", peer_timeout_s=60.0))
    eng = tr.engine
    if os.environ.get("DNN_CLUSTER"):                       # CTAs per image in the forward/backward kernel (default: auto = 6)
        eng.set_cluster_size(int(os.environ["DNN_CLUSTER"]))
    variant = os.environ.get("DNN_EXCHANGE", "")
    if world > 1 and variant:
        eng.set_exchange_variant({"one-shot": 0, "two-shot": 1}[variant])
    dx = torch.rand(64, B, 3, 32, 32, device="cuda") * 2 - 1
    dy = torch.randint(0, 10, (64, B), device="cuda")
    for _ in range(4):
        eng.train_device(dx, dy)
    torch.cuda.synchronize()
    tr.comm.barrier()
    ut = torch.zeros(1024 * 122 * 8, dtype=torch.int64, device="cuda")
    eng.set_update_trace(ut)
    extra = int(os.environ.get("DNN_DEBUG_FLAGS", "0"))   # 8: forward/backward waits for grid completion (no counter)
    eng.set_debug_flags(4 | extra)
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    host = os.environ.get("DNN_HOST")                     # the host-fed path (per-step inputs from pinned memory) instead
    if host:
        hx, hy = dx.cpu().pin_memory(), dy.cpu().pin_memory()
        eng.train_host(hx, hy, None)
        torch.cuda.synchronize()
        tr.comm.barrier()
    a.record()
    for _ in range(32):                 # 2048 steps; the ring keeps the last 1024
        if host:
            eng.train_host(hx, hy, None)
        else:
            eng.train_device(dx, dy)
    b.record()
    torch.cuda.synchronize()
    eng.set_debug_flags(0)
    eng.set_update_trace(None)
    step_us = a.elapsed_time(b) * 1e3 / 2048
    u = ut.view(1024, 122, 8).cpu().numpy()
    first = int(eng.read_ctl().step) % 1024          # oldest entry of the ring
    u = np.roll(u, -first, axis=0)
    rows = np.stack([u[:, :, 1].max(1), u[:, :, 2].max(1), u[:, :, 3].max(1), u[:, :, 0].min(1), u[:, :, 4].max(1),
                     u[:, :, 5].max(1), u[:, :, 6].max(1)], axis=1)
    out = Path("gpurun_out")
    out.mkdir(exist_ok=True)
    np.save(out / f"jitter_w{world}_r{tr.rank}.npy", rows)
    dep, grad, xd = rows[:, 0], rows[:, 1], rows[:, 2]
    period = np.diff(xd) / 1e3
    comp = (grad[1:] - xd[:-1]) / 1e3                   # previous update finished -> my local gradients done
    wait = (xd - grad) / 1e3
    print(f"rank {tr.rank}/{world} flags={extra}: {step_us:.2f} us/step | period mean {period.mean():.2f} p50 {np.median(period):.2f} "
          f"p90 {np.percentile(period, 90):.2f} | compute mean {comp.mean():.2f} std {comp.std():.2f} p90 "
          f"{np.percentile(comp, 90):.2f} | exchange mean {wait.mean():.2f} p50 {np.median(wait):.2f} p90 "
          f"{np.percentile(wait, 90):.2f}", flush=True)
    if host:   # where do the slow steps sit inside a 64-step call (groups of 4, 16, 16, 16, 12 steps)?
        pos = (np.arange(len(period)) + 1 + (int(eng.read_ctl().step) - len(xd))) % 64
        by = [float(np.mean(period[pos == k])) for k in range(64)]
        print(f"rank {tr.rank}/{world} host-fed: period p50 {np.median(period):.2f} p90 {np.percentile(period, 90):.2f} p99 "
              f"{np.percentile(period, 99):.2f} max {period.max():.1f}; mean period by position in the call: " +
              " ".join(f"{v:.0f}" for v in by), flush=True)
    sgd, fen, end = rows[:, 4], rows[:, 5], rows[:, 6]
    print(f"rank {tr.rank}/{world}: first update CTA enters -> dependency passed {np.mean(dep - rows[:, 3]) / 1e3:.2f}, -> local "
          f"gradients done {np.mean(grad - dep) / 1e3:.2f} (per-slice mean {np.mean(u[:, :, 2] - u[:, :, 1]) / 1e3:.2f}); "
          f"update CTAs enter over {np.mean(u[:, :, 0].max(1) - u[:, :, 0].min(1)) / 1e3:.2f} us", flush=True)
    print(f"rank {tr.rank}/{world}: tail of the update kernel (us): exchange done -> SGD stored {np.mean(sgd - xd) / 1e3:.2f}, "
          f"-> fence passed {np.mean(fen - sgd) / 1e3:.2f}, -> counter bumped {np.mean(end - fen) / 1e3:.2f}, "
          f"-> first CTA of the next update kernel enters {np.mean(rows[1:, 3] - end[:-1]) / 1e3:.2f}", flush=True)
    tr.comm.barrier()
    tr.close()
    tr.comm.close()


if __name__ == "__main__":
    main()
