"""Fault drill (BASELINE.json config 5): one rank drops out mid-epoch at N GPUs, the survivors keep training.

Run under torchrun / bin/mpiexec with N >= 2 ranks. Two scenarios, each from a fresh set of identical replicas:
  announced  rank K runs S steps, writes its leave marker (Engine.leave) and stops; the others run S + T steps.
  silent     rank K simply stops after S steps (no marker). The survivors' exchange gives up on it after
             `peer_timeout`, flags it in StepCtl::status, and the host (which looks at the status every CHUNK steps)
             takes it out of the membership mask; then the steps are fast again.
After either, the epoch boundary of the Trainer runs: membership epoch bumped, the returning rank copies parameters +
momentum from a survivor, adopts its sequence number, and one more common step shows everybody bit-identical.
What is recorded per scenario (rank 0's view unless said otherwise): microseconds per step for the steps around the
event from the update kernel's ring trace, the divisor (ranks in the sum) before/after, status words, whether the
survivors are bit-identical, whether everybody is after the rejoin, wall time (no hang: the whole drill is bounded).
reference: the sleep-only simulation of data_parallelism_train.py:41-46,117,141.
"""
import json
import os
import sys
import time
from pathlib import Path

import numpy as np
import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from distributed_neural_network_b200.trainer import TrainConfig, Trainer  # noqa: E402

B, S, T, CHUNK = 16, 300, 300, 10


def step_times(ut, end_step, n):
    """per-step [period us, divisor] for the last n steps before end_step, from the ring trace"""
    u = ut.view(1024, 122, 8).cpu().numpy()
    idx = [(end_step - n + k) % 1024 for k in range(n)]
    done = u[idx, :, 6].max(axis=1)                     # last CTA bumped the counter
    div = u[idx, :, 7]
    return np.diff(done) / 1e3, div.min(axis=1)[1:], div.max(axis=1)[1:]


def scenario(name, tr, x, y, k, timeout_s):
    eng, comm, rank, world = tr.engine, tr.comm, tr.rank, tr.world
    out = {"scenario": name, "world": world, "dropped_rank": k, "steps_before": S, "steps_after": T}
    full = (1 << world) - 1
    eng.set_membership(full, int(timeout_s * 1e9))
    eng.set_exchange_variant("one-shot")                 # a membership-safe variant (leave markers, per-step present set)
    ut = torch.zeros(1024 * 122 * 8, dtype=torch.int64, device="cuda")
    eng.set_update_trace(ut)
    eng.set_debug_flags(4)
    torch.cuda.synchronize()
    comm.barrier()
    t0 = time.time()
    mask, statuses = full, []
    if rank == k:
        eng.train_device(x[:S], y[:S])
        if name == "announced":
            eng.leave()
        torch.cuda.synchronize()
        tr.dropped, tr.steps_this_epoch = True, S
    else:
        done = 0
        while done < S + T:
            c = min(CHUNK, S + T - done)
            eng.train_device(x[done:done + c], y[done:done + c])
            done += c
            if name == "silent" and done >= S:          # the host looks at the status word every CHUNK steps
                ctl = eng.read_ctl()
                if ctl.status & 1:
                    gone = (ctl.status >> 8) & 0xffff
                    statuses.append({"after_step": done, "status": int(ctl.status), "gave_up_on_mask": int(gone)})
                    mask &= ~gone
                    eng.set_membership(mask, int(timeout_s * 1e9))   # also clears the sticky status
        torch.cuda.synchronize()
        tr.dropped, tr.steps_this_epoch = False, S + T
    wall = time.time() - t0
    eng.set_debug_flags(0)
    eng.set_update_trace(None)
    end = int(eng.read_ctl().step)
    if rank != k:
        per, dmin, dmax = step_times(ut, end, min(1000, S + T))
        ev = len(per) - T                                # index of the first step without rank k
        lo, hi = max(0, ev - 25), min(len(per), ev + 25)
        out.update(step_us_around_event=[round(float(v), 2) for v in per[lo:hi]],
                   divisor_around_event=[int(v) for v in dmin[lo:hi]],
                   median_step_us_before=float(np.median(per[:ev - 1])),
                   median_step_us_after=float(np.median(per[ev + CHUNK + 1:])),
                   divisor_before=int(dmax[ev - 2]), divisor_after=int(dmin[-1]),
                   slowest_step_us=float(per.max()), host_saw=statuses, final_mask=mask)
    flat = tr.model.flat_params.detach()
    sums = comm.allgather((rank, float(flat.double().sum()), float(flat.double().abs().sum()), wall))
    surv = [s for s in sums if s[0] != k]
    out["survivors_identical"] = all(s[1:3] == surv[0][1:3] for s in surv)
    out["dropped_rank_differs"] = [s for s in sums if s[0] == k][0][1:3] != surv[0][1:3]
    out["wall_s_max"] = max(s[3] for s in sums)
    # membership epoch at the boundary: the returning rank adopts a survivor's state; then one common step
    tr.membership.active = set(range(world))
    ep0 = tr.membership.epoch
    tr._epoch_boundary()
    eng.train_device(x[:8], y[:8])
    torch.cuda.synchronize()
    sums = comm.allgather((float(flat.double().sum()), float(flat.double().abs().sum())))
    out["all_identical_after_rejoin"] = all(s == sums[0] for s in sums)
    out["membership_epoch_bumps"] = tr.membership.epoch - ep0
    out["status_after"] = int(eng.read_ctl().status)
    return out


def main():
    world = int(os.environ.get("WORLD_SIZE", "1"))
    assert world >= 2, "run with at least 2 ranks"
    k = int(os.environ.get("DRILL_RANK", str(world // 2)))
    tr = Trainer(TrainConfig(batch_size=B, sync_mode="step-allreduce", synthetic=True, quiet=True, seed=7, device="cuda",
                             peer_timeout_s=30.0, failure_mode="drop", failure_probability=1e-9,
                             calibrate_exchange=False))
    g = torch.Generator().manual_seed(50 + tr.rank)
    x = (torch.rand((S + T, B, 3, 32, 32), generator=g) * 2 - 1).cuda()
    y = torch.randint(0, 10, (S + T, B), generator=g).cuda()
    for _ in range(2):
        tr.engine.train_device(x[:64], y[:64])
    res = [scenario("announced", tr, x, y, k, timeout_s=30.0),
           scenario("silent", tr, x, y, k, timeout_s=0.02)]
    if tr.rank == 0:
        out = {"what": "fault drill: one rank drops mid-epoch, survivors continue (step-allreduce, one-shot exchange)",
               "batch_per_gpu": B, "chunk_between_host_status_checks": CHUNK, "results": res}
        Path("gpurun_out").mkdir(exist_ok=True)
        Path(f"gpurun_out/fault_drill_n{world}.json").write_text(json.dumps(out, indent=1))
        for r in res:
            print({kk: v for kk, v in r.items() if kk not in ("step_us_around_event", "divisor_around_event")}, flush=True)
    tr.comm.barrier()
    tr.close()
    tr.comm.close()


if __name__ == "__main__":
    main()
