"""Timeline of the update kernel (and of the whole step) at N ranks: where does a multi-GPU step spend its time?
Run with torchrun; prints rank 0's view."""
import os
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from distributed_neural_network_b200.trainer import TrainConfig, Trainer  # noqa: E402


def main():
    B = 16
    world = int(os.environ.get("WORLD_SIZE", "1"))
    tr = Trainer(TrainConfig(batch_size=B, sync_mode="step-allreduce" if world > 1 else "none", synthetic=True, quiet=True,
                             seed=1, device="cuda", peer_timeout_s=60.0))
    eng = tr.engine
    eng.set_debug_flags(int(os.environ.get("DNN_DEBUG_FLAGS", "0")))
    cs, _ = eng.cluster_info(B)
    dx = torch.rand(64, B, 3, 32, 32, device="cuda") * 2 - 1
    dy = torch.randint(0, 10, (64, B), device="cuda")
    eng.train_device(dx, dy)
    torch.cuda.synchronize()
    tr.comm.barrier()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(8):
        eng.train_device(dx, dy)
    b.record()
    torch.cuda.synchronize()
    if tr.rank == 0:
        print(f"  step time {a.elapsed_time(b) * 1e3 / 512:.2f} us (debug flags {os.environ.get('DNN_DEBUG_FLAGS', '0')})")
    tr.comm.barrier()
    ut = torch.zeros(122 * 8, dtype=torch.int64, device="cuda")
    ft = torch.zeros(B * cs * 32, dtype=torch.int64, device="cuda")
    eng.set_update_trace(ut)
    eng.set_trace(ft)
    eng.train_device(dx[:8], dy[:8])      # the stamps of the LAST step survive
    torch.cuda.synchronize()
    eng.set_update_trace(None)
    eng.set_trace(None)
    tr.comm.barrier()
    if world > 1 and tr.rank < 2:
        print(f"  rank {tr.rank}: LL one-way latency {eng.pingpong(1 - tr.rank, 2000):.3f} us")
    if tr.rank < 2:
        u = ut.view(122, 8).cpu().double()
        f = ft.view(B * cs, 32).cpu().double()
        f_entry, f_end = f[:, 14], f[:, 15]
        t0 = f_entry.min()
        print(f"rank {tr.rank} world={world} cs={cs}  (ns relative to the first fwd/bwd CTA entry of the last step)")
        print(f"  fwd/bwd: last CTA leaves its image loop at {f_end.max()-t0:8.0f}")
        print(f"  update : CTAs enter at {u[:,0].min()-t0:8.0f}..{u[:,0].max()-t0:8.0f}; pass the dependency wait at "
              f"{u[:,1].min()-t0:8.0f}..{u[:,1].max()-t0:8.0f}")
        print(f"           local gradient done at {u[:,2].min()-t0:8.0f}..{u[:,2].max()-t0:8.0f} "
              f"(slowest slice {int(u[:,2].argmax())}); exchange done at {u[:,3].min()-t0:8.0f}..{u[:,3].max()-t0:8.0f}")
        print(f"           per-slice exchange time mean {(u[:,3]-u[:,2]).mean():6.0f} max {(u[:,3]-u[:,2]).max():6.0f} ns")
        t = ft.view(B * cs, 32).cpu()
        names = ["cluster_wait", "img+weights", "conv1", "wait p1", "conv2", "wait p2", "fc1(+wait)", "fc2(+wait)",
                 "fc3,CE,dx3,dx2(+wait)", "dx1(+wait)", "conv2 wgrad", "dgrad2", "conv1 wgrad"]
        d = (t[:, 1:14] - t[:, 0:13]).float()
        print(f"  rank {tr.rank} fwd/bwd stages, cycles mean/max: " + "; ".join(
            f"{names[k]} {d[:, k].mean():.0f}/{d[:, k].max():.0f}" for k in range(13)))
        pr = t[:, 16:21].double()
        print(f"  rank {tr.rank} prologue cycles: " + ", ".join(f"{(pr[:, k+1]-pr[:, k]).mean():.0f}" for k in range(4))
              + f", to loop start {(t[:, 0] - t[:, 20]).double().mean():.0f}; CTA entry spread {f_entry.max()-f_entry.min():.0f} ns; "
              f"entry->exit per CTA mean {(f_end-f_entry).mean():.0f} max {(f_end-f_entry).max():.0f} ns; "
              f"image loop mean {(t[:, 13]-t[:, 0]).double().mean()/1.965:.0f} ns")
    tr.close()
    tr.comm.close()


if __name__ == "__main__":
    main()
