#!/usr/bin/env python
"""Headline benchmark: CIFAR-10 CNN training throughput (images/s) at N GPUs, batch 16 per GPU.

    python bench.py                                  # 1 GPU, defaults
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port P bench.py --gpus N --steps K --warmup W
    python bench.py --impl reference --gpus N ...    # the unmodified reference (CPU, see baseline/)

Metric and config are the ones BASELINE.json names: data-parallel CNN (models/model.py), batch
16 per GPU (weak scaling), CIFAR-10-shaped synthetic data, random-init weights, a full optimizer
step (forward, loss, backward, gradient exchange, SGD-momentum) inside the timed region.

Numbers per run:
  value  device-timed (CUDA events around exactly K steps, max over ranks). Inputs come from a
         device-resident pool of distinct batches that is larger than the 126 MB L2 and is cycled,
         so no step re-reads its inputs from cache; the 248 KB of weights legitimately stay hot.
  e2e    the same K steps through the public host-fed API (`Engine.train_host`): every step's
         inputs are copied host->device from pinned memory and the step's loss is copied
         device->host; wall-clock, sync on both sides, max over ranks.
  allreduce_us / exchange_roofline_frac (N > 1)
         the other half of BASELINE.json's metric: exposed gradient exchange per step (from the
         update kernel's own time stamps, in a separate traced pass of the same steps) and the
         fraction of the NVLink roofline (bytes per direction at 900 GB/s) it reaches.

How the timed region is aligned (the whole region is 0.4 ms at the driver's 20 steps, so start-up
skew between ranks would BE the measurement): everything that touches Python imports, NVML or
the allocator happens before the barrier; after barrier + synchronize every rank enqueues, in one
burst with no host work in between, A untimed steps, the start event, the K timed steps and the
stop event. The exchange inside the untimed steps is the device-side barrier: when the start
event fires, all GPUs are within a few microseconds of each other and the launch queue is deep.
One JSON line on stdout (rank 0).
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

METRIC = "CIFAR-10 CNN images/sec (max over ranks), batch 16/GPU, full training step"
PER_GPU_BATCH = 16
# reference's best committed aggregate throughput: 2 children (3 procs), bs 16 -> 3949 img/s
# (log/log_epochs25_proc3_children.txt:2). Largest denominator available.
BASELINE_IMG_S = 3949.0


# --------------------------------------------------------------------------- clocks
class ClockSampler:
    """Samples SM clock and throttle reasons of one GPU with NVML while a region runs."""

    REASONS = {0x1: "gpu_idle", 0x2: "applications_clocks_setting", 0x4: "sw_power_cap", 0x8: "hw_slowdown",
               0x10: "sync_boost", 0x20: "sw_thermal_slowdown", 0x40: "hw_thermal_slowdown",
               0x80: "hw_power_brake_slowdown", 0x100: "display_clock_setting"}

    def __init__(self, cuda_index: int, period_s: float = 0.0005):
        self.samples, self.reasons, self.max_mhz, self.ok = [], set(), None, False
        self.period_s = period_s
        self.t_lo = self.t_hi = None      # perf_counter window of the region the samples are reported for
        self._stop = threading.Event()
        self._t = None
        try:
            import pynvml
            import torch

            pynvml.nvmlInit()
            self._nv = pynvml
            try:  # CUDA_VISIBLE_DEVICES renumbers CUDA devices but not NVML's: go through the UUID
                uuid = "GPU-" + str(torch.cuda.get_device_properties(cuda_index).uuid)
                self._h = pynvml.nvmlDeviceGetHandleByUUID(uuid.encode() if hasattr(uuid, "encode") else uuid)
            except Exception:
                self._h = pynvml.nvmlDeviceGetHandleByIndex(cuda_index)
            self.max_mhz = int(pynvml.nvmlDeviceGetMaxClockInfo(self._h, pynvml.NVML_CLOCK_SM))
            self.ok = True
        except Exception as e:  # NVML missing: the JSON says so instead of inventing numbers
            self.err = str(e)

    def _loop(self):
        nv = self._nv
        while not self._stop.is_set():
            try:
                mhz = int(nv.nvmlDeviceGetClockInfo(self._h, nv.NVML_CLOCK_SM))
                try:
                    bits = int(nv.nvmlDeviceGetCurrentClocksEventReasons(self._h))
                except Exception:
                    bits = int(nv.nvmlDeviceGetCurrentClocksThrottleReasons(self._h))
                self.samples.append((time.perf_counter(), mhz, bits))
            except Exception:
                pass
            time.sleep(self.period_s)

    def __enter__(self):
        if self.ok:
            self._t = threading.Thread(target=self._loop, daemon=True)
            self._t.start()
        return self

    def __exit__(self, *a):
        self._stop.set()
        if self._t:
            self._t.join()

    def nvlink_kib(self):
        """(transmitted, received) KiB over all NVLinks of this GPU since the driver loaded (payload counters), or None
        where NVML does not expose them (self.nvlink_note says what it answered)."""
        if not self.ok:
            return None
        nv = self._nv
        ids = (nv.NVML_FI_DEV_NVLINK_THROUGHPUT_DATA_TX, nv.NVML_FI_DEV_NVLINK_THROUGHPUT_DATA_RX)
        try:
            vals = nv.nvmlDeviceGetFieldValues(self._h, [(ids[0], 0xffffffff), (ids[1], 0xffffffff)])   # all links
            if all(v.nvmlReturn == 0 for v in vals):
                return tuple(int(v.value.ullVal) for v in vals)
            tot = [0, 0]
            for link in range(18):                                                                       # link by link
                vals = nv.nvmlDeviceGetFieldValues(self._h, [(ids[0], link), (ids[1], link)])
                if any(v.nvmlReturn != 0 for v in vals):
                    self.nvlink_note = f"NVML field NVLINK_THROUGHPUT_DATA: return code {vals[0].nvmlReturn} (link {link})"
                    return None
                tot[0] += int(vals[0].value.ullVal)
                tot[1] += int(vals[1].value.ullVal)
            return tuple(tot)
        except Exception as e:
            self.nvlink_note = f"NVML NVLink counters: {e}"
            return None

    def window(self, t_lo: float, t_hi: float, t_timed: float = None):
        self.t_lo, self.t_hi, self.t_timed = t_lo, t_hi, t_timed

    def summary(self):
        """Median SM clock and the union of throttle reasons over the samples taken while the GPU ran the warm-up,
        the alignment steps and the timed steps; `samples_in_timed_burst` of them fall between the host's enqueue
        of the alignment steps and the end of the final synchronize."""
        if not self.ok or not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": [], "note": "no NVML samples"}
        inside = [x for x in self.samples if self.t_lo is None or self.t_lo <= x[0] <= self.t_hi]
        use = inside or self.samples
        mhz = sorted(x[1] for x in use)
        reasons = set()
        for _, _, bits in use:
            for b, name in self.REASONS.items():
                if bits & b and name != "gpu_idle":
                    reasons.add(name)
        timed = [x for x in use if getattr(self, "t_timed", None) is not None and x[0] >= self.t_timed]
        return {"sm_mhz": mhz[len(mhz) // 2], "sm_min_mhz": mhz[0], "sm_max_mhz": self.max_mhz,
                "reasons": sorted(reasons), "samples": len(use), "samples_in_timed_burst": len(timed),
                "sm_mhz_in_timed_burst": sorted(x[1] for x in timed)[len(timed) // 2] if timed else None,
                "period_ms": self.period_s * 1e3,
                "window": "warm-up steps + alignment steps + timed steps (one continuous load)"}


# --------------------------------------------------------------------------- our arm
def run_ours(args) -> int:
    import torch

    from distributed_neural_network_b200.ops import native
    from distributed_neural_network_b200.utils.tracing import nvtx_range
    from distributed_neural_network_b200.trainer import (NVLINK_BYTES_PER_S, TrainConfig, Trainer,
                                                         exchange_bytes_per_direction)

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        if world == 1 and args.gpus > 1:  # convenience: re-launch ourselves under torchrun
            cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={args.gpus}",
                   "--master-addr", "127.0.0.1", "--master-port", str(29500 + os.getpid() % 500), __file__,
                   "--gpus", str(args.gpus), "--steps", str(args.steps), "--warmup", str(args.warmup),
                   "--batch", str(args.batch), "--exchange", args.exchange, "--prep", str(args.prep),
                   "--align", str(args.align), "--pool", str(args.pool)] + \
                  (["--precision", args.precision] if args.precision else [])
            return subprocess.call(cmd)
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}")
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (the native sm_100a engine has no CPU path)")
    native.require()
    torch.cuda.set_device(local)
    B, K, W = args.batch, args.steps, max(3, args.warmup)

    cfg = TrainConfig(batch_size=B, sync_mode="step-allreduce" if world > 1 else "none", synthetic=True, quiet=True,
                      seed=1234, lr=0.001, momentum=0.9, device="cuda", peer_timeout_s=60.0,
                      allreduce_variant=args.exchange, precision=args.precision)
    trainer = Trainer(cfg)            # public API: model + engine + peer group + control plane
    eng, comm = trainer.engine, trainer.comm
    if os.environ.get("DNN_DEBUG_FLAGS_BENCH"):           # engine A/B switches (include/dnnb200.h), development only
        eng.set_debug_flags(int(os.environ["DNN_DEBUG_FLAGS_BENCH"]))
    dev = trainer.device
    variant = trainer._exchange_variant() if world > 1 else 0
    variant_name = {0: "one-shot push, 16-byte LL lines (value, seq, value, seq)",
                    1: "two-shot (owner-reduce + broadcast), 16-byte LL lines",
                    2: "NVLS in-switch reduction (multimem.ld_reduce)",
                    3: "one-shot push, 128-byte lines (31 values + 1 flag word per warp-wide store)"}[variant]

    # ---- data: pools of distinct synthetic batches, larger than L2 -------------------------------
    pool = args.pool
    g = torch.Generator(device="cpu").manual_seed(100 + rank)
    hx = (torch.rand((pool, B, 3, 32, 32), generator=g) * 2 - 1).pin_memory()
    hy = torch.randint(0, 10, (pool, B), generator=g).pin_memory()
    hl = torch.zeros(pool).pin_memory()
    dx, dy = hx.to(dev), hy.to(dev)
    pool_bytes = dx.numel() * 4

    cur = {"d": 0, "h": 0}                      # the pools are walked round-robin ACROSS calls, so a batch is
                                                # re-read only after the other pool-1 batches (> L2) went by
    def run_device(n):
        while n > 0:
            c = min(pool - cur["d"], n)
            eng.train_device(dx[cur["d"]:cur["d"] + c], dy[cur["d"]:cur["d"] + c])
            cur["d"] = (cur["d"] + c) % pool
            n -= c

    def run_host(n):
        while n > 0:
            c = min(pool - cur["h"], n)
            eng.train_host(hx[cur["h"]:cur["h"] + c], hy[cur["h"]:cur["h"] + c],
                           None if os.environ.get("DNN_E2E_NO_LOSS") else hl[cur["h"]:cur["h"] + c])   # (A/B knob)
            cur["h"] = (cur["h"] + c) % pool
            n -= c

    # ---- preparation: everything slow happens BEFORE the barrier ----------------------------------
    A = max(W, args.align)                      # untimed steps enqueued directly in front of the timed ones
    prep = max(W, args.prep)                    # loads every kernel of the path, runs full engine calls
    clk = ClockSampler(torch.cuda.current_device())          # NVML init + handle lookup: milliseconds
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    utrace = torch.zeros(1024 * 122 * 8, dtype=torch.int64, device=dev) if world > 1 else None
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    clk.__enter__()                             # sampling starts with the warm-up: the same kernels at the same load, so the
    t_load = time.perf_counter()                # 0.4 ms timed region is not judged by the one sample that falls into it
    run_device(prep)
    run_host(prep)
    e0.record(); e1.record()                    # first use of an event allocates it
    flush.zero_()                               # 256 MiB written: whatever the upload or the warm-up left in L2 is gone
    torch.cuda.synchronize()

    # ---- device-timed region -------------------------------------------------------------------
    with nvtx_range("bench/device-timed"):
        comm.barrier()
        torch.cuda.synchronize()
        l0 = eng.launches
        t_burst = time.perf_counter()
        run_device(A)                           # untimed; its exchanges align the GPUs of all ranks on the device
        e0.record()
        run_device(K)
        e1.record()
        torch.cuda.synchronize()
        t_end = time.perf_counter()
    clk.__exit__(None, None, None)
    clk.window(t_load, t_end, t_burst)
    comm.barrier()
    launches = eng.launches - l0 - 2 * A
    ms = e0.elapsed_time(e1)
    ctl = eng.read_ctl()

    # ---- end-to-end region (pinned host -> device every step, loss -> host every step) -----------
    comm.barrier()
    torch.cuda.nvtx.range_push("bench/end-to-end")
    run_host(A)                                 # the last of these ends at the same moment on every GPU ...
    torch.cuda.synchronize()                    # ... so every host leaves this synchronize within microseconds
    t0 = time.perf_counter()
    run_host(K)
    torch.cuda.synchronize()
    t1 = time.perf_counter()
    torch.cuda.nvtx.range_pop()
    comm.barrier()
    e2e_s = t1 - t0
    loss_mean = float(hl[:min(pool, prep + A + K)].mean())

    # ---- exchange time per step, from the update kernel's own stamps (separate traced pass) -----------
    xch = None
    if world > 1:
        n_tr = min(1024, max(K, 256))
        eng.set_update_trace(utrace)
        eng.set_debug_flags(4)                  # ring mode: keep the stamps of the last 1024 steps
        comm.barrier()
        nvl0 = clk.nvlink_kib()
        run_device(A + n_tr)
        torch.cuda.synchronize()
        nvl1 = clk.nvlink_kib()
        eng.set_debug_flags(0)
        eng.set_update_trace(None)
        end = int(eng.read_ctl().step)
        u = utrace.view(1024, 122, 8).cpu().numpy()
        idx = [(end - n_tr + k) % 1024 for k in range(n_tr)]
        grad_done = u[idx, :, 2].max(axis=1)    # last slice has its local batch gradient
        xch_done = u[idx, :, 3].max(axis=1)     # last slice has the sum of all ranks
        wait = np.sort((xch_done - grad_done) / 1e3)
        xch = {"median_us": float(wait[len(wait) // 2]), "p90_us": float(wait[int(len(wait) * 0.9)]),
               "mean_us": float(wait.mean()), "steps": n_tr,
               # what the NVLink counters of this GPU saw per step during the traced pass (payload bytes, all links)
               "nvlink_tx_bytes_per_step": (nvl1[0] - nvl0[0]) * 1024 / (A + n_tr) if nvl0 and nvl1 else None,
               "nvlink_rx_bytes_per_step": (nvl1[1] - nvl0[1]) * 1024 / (A + n_tr) if nvl0 and nvl1 else None,
               "nvlink_note": getattr(clk, "nvlink_note", None)}
        comm.barrier()

    stats = comm.allgather({"ms": ms, "e2e_s": e2e_s, "launches": launches, "clk": clk.summary(),
                            "status": int(ctl.status), "loss": loss_mean, "xch": xch})
    if rank == 0:
        ms_max = max(s["ms"] for s in stats)
        e2e_max = max(s["e2e_s"] for s in stats)
        value = world * B * K / (ms_max / 1e3)
        e2e_value = world * B * K / e2e_max
        cs, _ = eng.cluster_info(B)
        # the other half of the metric: exposed exchange per step and its NVLink roofline
        xout = {}
        if world > 1:
            per_dir = exchange_bytes_per_direction(variant, world)
            t_roof_us = per_dir / NVLINK_BYTES_PER_S * 1e6
            xs = [s_["xch"] for s_ in stats]
            med = max(x["median_us"] for x in xs)
            xout = {"allreduce_us": round(med, 3),
                    "allreduce_us_p90": round(max(x["p90_us"] for x in xs), 3),
                    "exchange_roofline_frac": round(t_roof_us / med, 4) if med > 0 else None,
                    "exchange": {"bytes_per_direction_per_gpu": per_dir, "roofline_us_at_900GBps": round(t_roof_us, 3),
                                 "definition": "per step: (last slice holds the sum of all ranks) - (last slice holds its "
                                               "local batch gradient), globaltimer stamps of the update kernel; median over "
                                               f"{xs[0]['steps']} traced steps run after the timed region, max over ranks; "
                                               "compute at peak is negligible, so the roofline is the NVLink time",
                                 "per_rank_median_us": [round(x["median_us"], 3) for x in xs],
                                 "nvlink_counters_bytes_per_step_per_gpu": {
                                     "tx": xs[0]["nvlink_tx_bytes_per_step"], "rx": xs[0]["nvlink_rx_bytes_per_step"],
                                     "note": xs[0].get("nvlink_note") or
                                             "NVML NVLINK_THROUGHPUT_DATA_TX/RX of rank 0's GPU over the traced pass "
                                             "(KiB granularity), to set against bytes_per_direction_per_gpu"}}}
        out = {
            "metric": METRIC, "value": round(value, 1), "unit": "images/s", "n_gpus": world, "steps": K,
            "warmup": prep + A, "warmup_requested": args.warmup,
            "warmup_note": f"{prep} steps before the barrier (kernels loaded, full engine calls) + {A} untimed steps "
                           "enqueued in the same burst as the timed ones (device-side alignment of the ranks)",
            "ms_per_step": round(ms_max / K, 6), "higher_is_better": True, "scaling": "weak",
            "vs_baseline": round(value / BASELINE_IMG_S, 2),
            "baseline_ref": "3949 img/s: reference's best committed aggregate (3 procs = 2 children, bs16, 8-core CPU; "
                            "BASELINE.md section 2)",
            "dtype": ("tf32 tensor-core operands for conv1/conv2 forward and the conv2 data gradient (tcgen05, fp32 "
                      "accumulate); FC layers, weight gradients, softmax, SGD and master weights fp32"
                      if eng.precision == "tf32" and cs > 1 else
                      "fp32 (FFMA, fp32 accumulate, fp32 master weights) except the conv2 data gradient, which runs on "
                      "the tensor cores (tcgen05, TF32 operands, fp32 accumulate): forward, loss and all gradients but "
                      "conv1's are the fp32 ones"
                      if eng.precision == "tf32-dgrad" and cs > 1 else
                      "fp32 (FFMA, fp32 accumulate, fp32 master weights; reference is fp32)"),
            "data": "synthetic CIFAR-10-shaped (3x32x32 in [-1,1], 10 classes), random-init weights",
            "config": {"model": "models/model.py Network (LeNet-5 style CNN, 62006 params)",
                       "global_batch": world * B, "per_gpu_batch": B, "seq_len": None,
                       "parallelism": f"dp{world}", "optimizer": "SGD lr=0.001 momentum=0.9 (inside the timed step)",
                       "exchange": "none (1 GPU)" if world == 1 else
                                   variant_name + " over NVLink peer memory, fused into the update kernel; chosen " +
                                   ("by --exchange" if args.exchange != "auto" else
                                    "by the start-up measurement of every usable variant (exchange_table)"),
                       "l2": f"inputs cycle through a pool of {pool} distinct batches = {pool_bytes >> 20} MiB on the "
                             f"device (> 126 MB L2), walked round-robin across all calls; 256 MiB written to flush L2 "
                             f"after the warm-up, before the burst",
                       "cluster_size": cs, "kernels_per_step": 2},
            "clocks": stats[0]["clk"] if world == 1 else
                      {"sm_mhz": min((s["clk"]["sm_mhz"] or 0) for s in stats), "sm_max_mhz": stats[0]["clk"]["sm_max_mhz"],
                       "reasons": sorted({r for s in stats for r in s["clk"]["reasons"]}),
                       "samples": min(s["clk"].get("samples", 0) for s in stats),
                       "per_rank_sm_mhz": [s["clk"]["sm_mhz"] for s in stats]},
            "e2e": {"value": round(e2e_value, 1), "unit": "images/s",
                    "h2d_bytes_per_step": B * 3 * 32 * 32 * 4 + B * 8, "d2h_bytes_per_step": 4,
                    "ms_per_step": round(e2e_max * 1e3 / K, 6),
                    "how": "Engine.train_host: every step's images+labels are copied from pinned host memory inside the "
                           "region (cudaMemcpyAsync on a copy stream into a 32-slot device ring, two steps per copy; the "
                           "kernels wait for a sequence number the copy stream writes behind the data, no stream events), "
                           "2 kernels per step, and the update kernel stores each step's mean loss into the caller's "
                           "pinned host buffer (device->host over PCIe); wall clock with synchronize on both sides, "
                           "max over ranks",
                    "mean_loss_read_back": round(max(s["loss"] for s in stats), 4)},
            "gpu_launches": int(sum(s["launches"] for s in stats) // world),
            "gpu_launches_note": "kernels of libdnnb200.so per rank inside the device-timed region "
                                 "(fwd_bwd[_cluster]_kernel + update_kernel per step)",
            "peer_timeouts": [s["status"] for s in stats],
            "impl": "ours",
        }
        out.update(xout)
        if trainer.exchange_table:
            out["exchange_table"] = trainer.exchange_table
        print(json.dumps(out), flush=True)
    trainer.close()
    comm.close()
    return 0


# --------------------------------------------------------------------------- reference arm
def run_reference(args) -> int:
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0  # the reference's ranks are spawned by rank 0 below (it is a CPU program)
    ref = ROOT / "baseline" / "_ref" / "data_parallelism_train.py"
    if not ref.exists():
        try:
            sys.path.insert(0, str(ROOT / "baseline"))
            import install_reference

            install_reference.ensure()
        except Exception:
            pass
    if not ref.exists():
        print(json.dumps({"impl": "reference", "unavailable": "baseline/_ref is empty: the reference has no "
                          "setup.py/pyproject.toml (pip install fails) and its sources were not copied in"}))
        return 0
    N, K, B = args.gpus, args.steps, args.batch
    K = min(K, args.ref_max_steps)                      # a CPU reference step takes milliseconds, not microseconds
    warm_epochs = 1
    P = N + 1                                           # N children + the parent that never trains (reference Q1)
    work = ROOT / "baseline" / "_work" / f"n{N}_{os.getpid()}"
    work.mkdir(parents=True, exist_ok=True)
    (work / "log").mkdir(exist_ok=True)
    import socket

    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    threads = max(1, (os.cpu_count() or P) // P)
    procs = []
    train_len = N * B * K                               # each child gets exactly K batches of B per epoch
    for r in range(P):
        env = dict(os.environ)
        env.update({"REFSHIM_RANK": str(r), "REFSHIM_SIZE": str(P), "REFSHIM_PORT": str(port + 1),
                    "REFSHIM_THREADS": str(threads), "OMP_NUM_THREADS": str(threads), "CUDA_VISIBLE_DEVICES": ""})
        cmd = [sys.executable, str(ROOT / "baseline" / "run_reference.py"), "data_parallelism_train.py",
               str(work / f"rank{r}.json"), str(train_len), str(B), "--", "--nb-proc", str(P), "--batch-size", str(B),
               "--epochs", str(warm_epochs + 1)]
        procs.append(subprocess.Popen(cmd, cwd=work, env=env, stdout=subprocess.DEVNULL, stderr=subprocess.PIPE))
    errs = []
    for p in procs:
        try:
            _, err = p.communicate(timeout=args.ref_timeout)
        except subprocess.TimeoutExpired:
            p.kill()
            _, err = p.communicate()
            errs.append("timeout")
        if p.returncode != 0:
            errs.append(err.decode(errors="replace")[-400:])
    if errs:
        print(json.dumps({"impl": "reference", "unavailable": "reference run failed: " + " | ".join(errs)[:600]}))
        return 0
    # per child: epoch e runs from its e-th recv (weights arrive) to its e-th send (weights leave)
    dur = []
    for r in range(1, P):
        ev = json.loads((work / f"rank{r}.json").read_text())["events"]
        recvs = [t for what, t, _ in ev if what == "recv"]
        sends = [t for what, t, _ in ev if what == "send"]
        dur.append(sends[warm_epochs] - recvs[warm_epochs])
    t_max = max(dur)
    value = N * B * K / t_max
    out = {"metric": METRIC, "value": round(value, 1), "unit": "images/s", "n_gpus": N, "steps": K,
           "warmup": warm_epochs * K, "ms_per_step": round(t_max * 1e3 / K, 4), "higher_is_better": True,
           "scaling": "weak", "vs_baseline": round(value / BASELINE_IMG_S, 3), "dtype": "fp32",
           "data": "synthetic CIFAR-10-shaped uint8 images through the reference's own transform + DataLoader",
           "config": {"model": "reference models/model.py Network", "global_batch": N * B, "per_gpu_batch": B,
                      "parallelism": f"{N} children + 1 parent (mpi4py star, weights averaged once per epoch)",
                      "device": "CPU -- the reference's distributed scripts never move the model to a GPU",
                      "threads_per_process": threads,
                      "shims": "mpi4py -> local sockets, neptune -> no-op, CIFAR-10 download -> synthetic; "
                               "the reference script itself runs unmodified from baseline/_ref"},
           "timing": "wall clock inside each child from receiving the weights to sending them back (its whole epoch "
                     "of K steps), max over children; a CPU program has no device timeline",
           # The reference's only timeline IS end to end: the child's epoch from the arrival of the weights to their
           # departure includes its DataLoader (worker process, uint8 -> float transform, collation), forward, loss,
           # backward, optimizer step and loss.item() of every step (data_parallelism_train.py:190-207). Nothing moves
           # between host and device because the reference's distributed path never leaves the CPU.
           "e2e": {"value": round(value, 1), "unit": "images/s", "ms_per_step": round(t_max * 1e3 / K, 4),
                   "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0,
                   "how": "wall clock recv->send of the child's epoch incl. the reference's DataLoader + transform, "
                          "max over children (same region as `value`: a CPU program has no separate device time)"},
           "clocks": {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "note": "CPU program: no GPU clocks apply"},
           "gpu_launches": 0,
           "impl": "reference"}
    print(json.dumps(out), flush=True)
    return 0


def main() -> int:
    ap = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawDescriptionHelpFormatter)
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20000)
    ap.add_argument("--warmup", type=int, default=200)
    ap.add_argument("--batch", type=int, default=PER_GPU_BATCH, help="per-GPU batch (the headline config is 16)")
    ap.add_argument("--pool", type=int, default=1024, help="distinct batches cycled through (1024 x 16 = 201 MB > L2)")
    ap.add_argument("--prep", type=int, default=512, help="untimed steps before the barrier (at least --warmup)")
    ap.add_argument("--align", type=int, default=128,
                    help="untimed steps enqueued directly in front of the timed ones (at least --warmup)")
    ap.add_argument("--impl", choices=["ours", "reference"], default="ours")
    ap.add_argument("--exchange", choices=["auto", "one-shot", "one-shot-128", "two-shot", "nvls"], default="auto",
                    help="gradient exchange variant (auto: every usable variant is timed at start-up, the fastest runs)")
    ap.add_argument("--precision", choices=["fp32", "tf32", "tf32-dgrad"], default=None,
                    help="conv stages of the fused step: fp32 = FFMA (default, measured faster), tf32 = tcgen05 tensor cores")
    ap.add_argument("--ref-max-steps", type=int, default=400)
    ap.add_argument("--ref-timeout", type=float, default=1500.0)
    a = ap.parse_args()
    return run_reference(a) if a.impl == "reference" else run_ours(a)


if __name__ == "__main__":
    sys.exit(main())
