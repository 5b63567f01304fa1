"""Epoch loop and synchronisation modes shared by the three entry points.

Reference orchestration: ``main``/``run_child``/``run_parent``/``eval`` of
data_parallelism_train.py:56-254 and model_replication_train.py:31-148, and the script body of
single_proc_train.py:16-107. What the modes mean here:

``step-allreduce`` (default, new)
    every rank trains its shard; gradients are averaged EVERY step inside the update kernel
    (peer-memory all-reduce fused with 1/W and SGD-momentum). Replicas stay bit-identical, so
    there is nothing to average at the epoch boundary.
``epoch-average`` (reference parity)
    each trainer runs plain local SGD for one epoch with a fresh optimizer (momentum zeroed,
    reference :187), then the PARAMETERS are averaged (reference :238-240) -- with the same
    all-reduce kernel in "mean only" mode instead of pickled state_dicts through rank 0.
    ``reference_topology=True`` also reproduces who trains: rank 0 is a parent that only
    evaluates, ranks 1..P-1 train contiguous shards of ``len // (P-1)`` (reference :49-53,66).
``replication`` (reference model_replication_train.py)
    like ``epoch-average`` but every trainer sees the full training set with its own shuffle.
``none``
    single process (reference single_proc_train.py).

On CUDA the step runs through the native engine (two kernels per step, data resident in HBM or
streamed from pinned memory); on CPU the same algorithm runs in plain PyTorch with gloo for the
exchange, which is what the no-GPU tests and the "plumbing" configuration of BASELINE.json use.
"""
from __future__ import annotations

import os
import time
from dataclasses import dataclass
from typing import Dict, List, Optional, Tuple

import numpy as np
import torch
import torch.nn.functional as F

from .models.network import Network
from .ops import flat as L
from .ops import functional as Fn
from .ops import native
from .parallel import bootstrap, nvls
from .parallel.membership import FailureSimulator, Membership
from .utils import checkpoint as ckpt
from .utils import data as D
from .utils import logs
from .utils.tracing import nvtx_range


@dataclass
class TrainConfig:
    lr: float = 0.001                 # reference default (data_parallelism_train.py:260)
    momentum: float = 0.9             # :261
    batch_size: int = 16              # :262
    epochs: int = 25                  # :263
    nb_proc: int = 4                  # :264 -- labels the log files only, as in the reference
    failure_probability: float = 0.0  # :265
    failure_duration: float = 0.0     # :267
    failure_mode: str = "sleep"       # sleep = reference behaviour; drop = membership epoch
    sync_mode: str = "step-allreduce"
    reference_topology: bool = False
    device: str = "auto"              # auto | cpu | cuda
    data_root: str = "./data"
    synthetic: Optional[bool] = None  # None: real CIFAR-10 if present, else synthetic
    train_size: Optional[int] = None  # truncate / size the (synthetic) training set
    test_size: Optional[int] = None
    seed: Optional[int] = None        # the reference has none; set for reproducible runs
    log_dir: str = "log"
    jsonl: Optional[str] = None
    save: Optional[str] = None
    resume: Optional[str] = None
    input_pipeline: str = "device"    # device: dataset resident in HBM; host: per-step H2D from pinned memory
    allreduce_variant: str = "auto"        # auto (measured at start-up) | one-shot | one-shot-128 | two-shot | nvls
    calibrate_exchange: bool = True        # auto: time every usable variant on this world size and take the fastest
    calibrate_steps: int = 300
    precision: Optional[str] = None        # conv stages of the fused step: fp32 (FFMA; parity with PyTorch fp32 and, at
                                           # these GEMM sizes, the faster one) | tf32 (tcgen05 tensor cores);
                                           # None = native.default_precision()
    step_path: str = "fused"               # fused: the two-kernel step engine (default) | autograd: Network.forward +
                                           # loss.backward() through the layer kernels, gradients averaged bucket by
                                           # bucket on a side stream while backward runs (large batches per GPU)
    overlap_exchange: Optional[bool] = None  # autograd path: exchange buckets while backward runs (True) or after it
                                           # (False, and None). Off unless asked for: the overlap won only at 2 GPUs with
                                           # 4096 images each (1161 vs 1171 us) and lost everywhere else, at 8 GPUs at
                                           # every size -- profiles/autograd_overlap_n{2,8}.json
    exchange_timing: bool = True           # per-epoch "allreduce_us" in the JSONL (the reference books its
                                           # mpi_communication_time_children, data_parallelism_train.py:209-213)
    peer_timeout_s: float = 30.0
    eval_batch_size: Optional[int] = None  # the reference evaluates with --batch-size too (Q11)
    neptune: bool = False
    quiet: bool = False

    def __post_init__(self):
        if self.sync_mode not in ("step-allreduce", "epoch-average", "replication", "none"):
            raise ValueError(f"unknown sync mode {self.sync_mode!r}")
        if self.step_path not in ("fused", "autograd"):
            raise ValueError(f"unknown step path {self.step_path!r}")
        self.lr, self.momentum = float(self.lr), float(self.momentum)   # the reference forgets the casts (Q5)
        self.batch_size, self.epochs = int(self.batch_size), int(self.epochs)


def exchange_bytes_per_direction(variant: int, world: int) -> int:
    """Bytes one GPU sends (= receives) over NVLink per step for the flat gradient, per exchange variant: the
    numerator of the exchange roofline (bytes over NVLink at 900 GB/s per direction, BASELINE.json)."""
    vec = L.FLAT_NUMEL * 4
    if world <= 1:
        return 0
    return {0: (world - 1) * vec * 2,                    # one-shot, 16-byte lines carry 8 bytes of payload
            1: 2 * (world - 1) * vec * 2 // world,       # two-shot: reduce-scatter + all-gather, 16-byte lines
            2: (world - 1) * vec,                        # NVLS: every peer's ld_reduce pulls my vector once
            3: (world - 1) * 122 * 17 * 128}[int(variant)]   # one-shot, 128-byte lines (31 values + flag), 17 per slice


NVLINK_BYTES_PER_S = 900e9   # per direction per GPU (NVLink 5)


def reference_global_avg_loss(total_losses: float, n_children: int) -> float:
    """The reference's 'Global Average Training Loss' (quirk Q4: it divides the summed batch
    losses by 10 keys x children, data_parallelism_train.py:233,248). Kept for parity tests."""
    return total_losses / float(10 * n_children)


class Trainer:
    def __init__(self, cfg: TrainConfig, comm: Optional[bootstrap.Comm] = None, model: Optional[Network] = None,
                 run=None):
        self.cfg = cfg
        self.comm = comm if comm is not None else bootstrap.init()
        self.rank, self.world = self.comm.rank, self.comm.world
        self.timers = logs.Timers()
        self.run = run if run is not None else (logs.init_run(enabled=cfg.neptune) if self.rank == 0 else logs.NullRun())
        self.jsonl = logs.JsonlLog(cfg.jsonl if self.rank == 0 else None)
        self.device = self._pick_device()
        if cfg.seed is not None:
            torch.manual_seed(cfg.seed + self.rank)
        self.gen = torch.Generator().manual_seed((cfg.seed if cfg.seed is not None else int(time.time())) + 1000 * self.rank)

        # who trains / who evaluates
        self.parent_only = cfg.reference_topology and self.world > 1 and self.rank == 0
        self.trainers: List[int] = list(range(1, self.world)) if (cfg.reference_topology and self.world > 1) \
            else list(range(self.world))
        self.is_trainer = self.rank in self.trainers
        self.membership = Membership(self.world, active=set(self.trainers))

        self.model = (model if model is not None else Network()).to(self.device)
        self.momentum = torch.zeros(L.FLAT_NUMEL, dtype=torch.float32, device=self.device)
        self.start_epoch = 0
        if cfg.resume:
            meta = ckpt.load(cfg.resume, self.model, self.momentum)
            self.start_epoch = int(meta.get("epoch", -1)) + 1
        # replicas must start identical: rank 0's weights win, as the parent's do in the reference
        # (data_parallelism_train.py:118,137)
        if self.world > 1:
            flat0 = self.comm.bcast(self.model.flat_params.detach().cpu() if self.rank == 0 else None, root=0)
            self.model.flat_params.copy_(flat0.to(self.device))

        self.engine: Optional[native.Engine] = None
        self.group: Optional[native.Group] = None
        self.first_step = True
        self.nvls_ok = False
        self._nvls_probed = False
        self.ll128_ok = False                           # 128-byte lines arrive whole on this job's links (probed below)
        self._utrace: Optional[torch.Tensor] = None
        self._averager = None                           # autograd path: per-bucket exchange overlapped with backward
        self.last_exchange_us: Optional[float] = None  # median exposed exchange per step over the epoch's last steps
        self.exchange_choice: Optional[str] = None     # measured winner (calibrate_exchange)
        self.exchange_table: Dict = {}
        if self.device.type == "cuda":
            self.engine = native.Engine(self.model.flat_params, self.momentum, None, max(cfg.batch_size, 1),
                                        cfg.lr, cfg.momentum, precision=cfg.precision)
            if self.world > 1:
                self.group = native.Group(self.rank, self.world)
                handles = self.comm.allgather(self.group.handle)
                self.group.connect(handles)
                self.engine.attach_group(self.group)
                self._apply_membership()
                # NVLS (in-switch reduction) where the host exposes NVLink multicast; all ranks or nobody
                made = nvls.try_create(self.device, self.comm)
                if all(self.comm.allgather(made is not None)):
                    self.nvls_buf, self.nvls_mc = made
                    self.engine.attach_nvls(self.nvls_buf, self.nvls_mc)
                    self.nvls_ok = True
                self._nvls_probed = True
                self._apply_membership()
                self.ll128_ok = self._probe_ll128()
        self.failure = FailureSimulator(cfg.failure_probability, cfg.failure_duration, cfg.failure_mode, cfg.seed,
                                        self.rank)
        self.dropped = False
        self.saw_timeout = False
        self.steps_this_epoch = 0
        self.history: List[Dict] = []
        if (self.engine is not None and self.world > 1 and cfg.sync_mode == "step-allreduce"
                and cfg.allreduce_variant == "auto" and cfg.calibrate_exchange and not cfg.reference_topology):
            self.calibrate_exchange()

    # ------------------------------------------------------------------ setup helpers
    def _pick_device(self) -> torch.device:
        want = self.cfg.device
        if want == "cpu" or (want == "auto" and not torch.cuda.is_available()):
            # P ranks on one host: give each its share of the cores. (The reference lets every rank
            # run a fully threaded torch on the same 8 cores -- the negative scaling of its Table 1.)
            if self.world > 1 and "OMP_NUM_THREADS" not in os.environ:
                torch.set_num_threads(max(1, (os.cpu_count() or 1) // self.world))
            return torch.device("cpu")
        idx = int(os.environ.get("DNNB200_DEVICE", os.environ.get("LOCAL_RANK", self.comm.local_rank)))
        idx %= max(1, torch.cuda.device_count())
        torch.cuda.set_device(idx)
        return torch.device("cuda", idx)

    def _log(self, *a) -> None:
        if not self.cfg.quiet:
            print(*a, flush=True)

    def _sync(self) -> None:
        if self.device.type == "cuda":
            torch.cuda.synchronize()

    def _can_drop(self) -> bool:
        return self.cfg.failure_mode == "drop" and self.cfg.failure_probability > 0

    def _probe_ll128(self, iters: int = 300, n_ctas: int = 8) -> bool:
        """The one-shot-128 exchange relies on a warp-wide 128-byte store becoming visible in a peer's memory as a
        unit (31 values + 1 flag word; tests/test_protocol_model.py shows what goes wrong otherwise). Ranks pair up
        (0-1, 2-3, ...) and play ping-pong with such lines; one torn line anywhere, or a pair that did not finish,
        and the variant is not offered to this job. Collective."""
        ok = self.world <= 8
        peer = self.rank ^ 1
        self.comm.barrier()
        if ok and peer < self.world:
            try:
                torn, _, gave_up = self.engine.ll128_probe(peer, iters=iters, n_ctas=n_ctas)
                ok = torn == 0 and not gave_up
            except RuntimeError:
                ok = False
        return all(self.comm.allgather(ok))

    def _usable_variants(self) -> List[str]:
        """Exchange variants this run may use. The one-shot forms tolerate ranks leaving mid-epoch (leave markers);
        two-shot has static slice ownership and the NVSwitch adds EVERY member of the multicast group, so both need
        a membership that cannot change (no drop-mode failure drill, every rank a stepping trainer for NVLS)."""
        names = ["one-shot"] + (["one-shot-128"] if self.world <= 8 and self.ll128_ok else [])   # (parks 7 peers' lines)
        if not self._can_drop():
            names.append("two-shot")
            if self.nvls_ok and len(self.membership.active) == self.world:
                names.append("nvls")
        return names

    def _exchange_variant(self) -> int:
        """Variant of the exchange fused into the update kernel (numbering of include/dnnb200.h: 0 one-shot, 1 two-shot,
        2 NVLS, 3 one-shot with 128-byte lines)."""
        v = self.cfg.allreduce_variant
        usable = self._usable_variants()
        if v != "auto":
            if v == "nvls" and not self.nvls_ok and self._nvls_probed:
                raise RuntimeError("--allreduce-variant nvls: NVLink multicast is not available to this job")
            return native.variant_id(v if v in usable else "one-shot")
        if self.exchange_choice in usable:
            return native.variant_id(self.exchange_choice)
        # not measured (calibration switched off): the variant that tolerates ranks leaving and wins up to 4 GPUs
        # (profiles/exchange_table_n*.json); two-shot wins at 8 but needs a static membership
        return native.variant_id("one-shot")

    def _exchange_two_shot(self) -> bool:
        return self._exchange_variant() == 1

    def _standalone_variant(self):
        """Variant of the stand-alone all-reduce (parameter averaging once per epoch; not on the step path)."""
        v = self.cfg.allreduce_variant
        if v == "nvls":
            if not self.nvls_ok:
                raise RuntimeError("--allreduce-variant nvls: NVLink multicast is not available to this job")
            return 2
        if v == "auto":   # the measured winner of the fused exchange where there is one; else by world size (round-1 table)
            if self.exchange_choice in ("one-shot", "one-shot-128", "two-shot"):
                return native.variant_id(self.exchange_choice)
            return 1 if self.world >= 6 else 0
        if v == "one-shot-128" and self.world > 8:
            return 0
        return native.variant_id(v)

    @torch.no_grad()
    def calibrate_exchange(self, steps: Optional[int] = None) -> Dict:
        """Measure, on THIS world size and batch size, every usable exchange variant inside the real step
        (forward/backward + update kernel, device-timed, max over ranks) and keep the fastest. The message is
        the 248 KB flat gradient whatever the batch size, so it is latency-bound and the best variant depends on
        the number of ranks, not on a formula: the table is measured, not assumed. Collective: every rank calls
        it at the same point. Parameters and momentum are restored afterwards; only the step counter moves on."""
        if self.engine is None or self.world == 1:
            return {}
        steps = int(steps or self.cfg.calibrate_steps)
        B, eng = self.cfg.batch_size, self.engine
        with nvtx_range("calibrate/exchange variants"):
            return self._calibrate_exchange(steps, B, eng)

    def _calibrate_exchange(self, steps: int, B: int, eng) -> Dict:
        keep_p, keep_m = self.model.flat_params.detach().clone(), self.momentum.clone()
        g = torch.Generator(device="cpu").manual_seed(4242 + self.rank)
        nb = 64
        x = (torch.rand((nb, B, 3, 32, 32), generator=g) * 2 - 1).to(self.device)
        y = torch.randint(0, 10, (nb, B), generator=g).to(self.device)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        table: Dict[str, float] = {}
        self.engine.set_membership(self.membership.mask, int(self.cfg.peer_timeout_s * 1e9))
        stepping = self.is_trainer and self.rank in self.membership.active   # a parent that never trains only listens
        for name in (self._usable_variants() if stepping else []):
            eng.set_exchange_variant(name)
            eng.train_device(x, y)                       # warm-up; also the device-side alignment of the ranks
            e0.record()
            done = 0
            while done < steps:
                c = min(nb, steps - done)
                eng.train_device(x[:c], y[:c])
                done += c
            e1.record()
            torch.cuda.synchronize()
            table[name] = e0.elapsed_time(e1) * 1e3 / steps
        per_rank = [t for t in self.comm.allgather(table) if t]
        worst = {k: max(t[k] for t in per_rank) for k in per_rank[0]}    # a step is as slow as its slowest rank
        self.exchange_choice = min(worst, key=worst.get)
        self.exchange_table = {"world": self.world, "batch_size": B, "steps": steps, "message_bytes": L.FLAT_NUMEL * 4,
                               "us_per_step_max_over_ranks": {k: round(v, 3) for k, v in worst.items()},
                               "chosen": self.exchange_choice}
        ctl = eng.read_ctl()
        if ctl.status & 1:
            self.saw_timeout = True
        self.model.flat_params.copy_(keep_p)
        self.momentum.copy_(keep_m)
        eng.reset_loss()
        self._sync()
        self.comm.barrier()
        self._apply_membership()
        if self.rank == 0:
            self.jsonl.write(event="exchange_calibration", **self.exchange_table)
            self._log(f"exchange variants, us/step at {self.world} ranks: {self.exchange_table['us_per_step_max_over_ranks']}"
                      f" -> {self.exchange_choice}")
        return self.exchange_table

    def _apply_membership(self, local_only: bool = False) -> None:
        """Program the exchange kernels: who is in the sum, how long to wait for a silent peer."""
        if self.engine is None or self.world == 1:
            return
        mask = (1 << self.rank) if local_only else self.membership.mask
        self.engine.set_membership(mask, int(self.cfg.peer_timeout_s * 1e9))
        self.engine.set_exchange_variant(0 if local_only else self._exchange_variant())

    # ------------------------------------------------------------------ data
    def load_data(self) -> Tuple[Optional[D.TensorDataset], Optional[D.TensorDataset]]:
        """Per-role data, timed like the reference (:66-96): trainers the train shard, rank 0 the test set."""
        cfg = self.cfg
        t0 = time.time()
        shard = test = None
        use_syn = cfg.synthetic
        if self.is_trainer:
            if use_syn:
                full = D.synthetic(cfg.train_size or D.TRAIN_SIZE, seed=0)
            else:
                full = D.cifar10(cfg.data_root, True, allow_synthetic=(use_syn is None),
                                 synthetic_size=cfg.train_size, verbose=self.rank == self.trainers[0] and not cfg.quiet)
                if cfg.train_size and len(full) > cfg.train_size:
                    full = D.Subset(full, range(cfg.train_size)).materialize()
            if cfg.sync_mode == "replication" or self.world == 1:
                part = D.Subset(full, range(len(full)))
            elif cfg.reference_topology:
                part = D.partition_dataset(full, self.rank, self.world)
            else:
                part = D.partition_symmetric(full, self.rank, self.world)
            shard = part.materialize()
            self._log("(Loaded Train Dataset for worker {0} of length {1})".format(self.rank, len(shard)))
        if self.rank == 0:
            if use_syn:
                test = D.synthetic(cfg.test_size or D.TEST_SIZE, seed=1)
            else:
                test = D.cifar10(cfg.data_root, False, allow_synthetic=(use_syn is None), synthetic_size=cfg.test_size,
                                 verbose=not cfg.quiet)
                if cfg.test_size and len(test) > cfg.test_size:
                    test = D.Subset(test, range(cfg.test_size)).materialize()
        if self.device.type == "cuda":
            if shard is not None:
                shard = shard.to(self.device) if cfg.input_pipeline == "device" else shard.pin_memory()
            if test is not None:
                test = test.to(self.device)
            self._sync()
        self.timers.data_loading_time += time.time() - t0
        return shard, test

    # ------------------------------------------------------------------ one epoch of training
    def train_epoch(self, shard: D.TensorDataset, drop_at: Optional[int] = None) -> Tuple[float, int]:
        """One pass over ``shard``; returns (sum of per-batch mean losses, number of batches) like the
        reference's ``total_loss``/``total_batches`` (:188-203). ``drop_at``: leave the group after
        that many steps (failure drill)."""
        cfg = self.cfg
        B = cfg.batch_size
        local_sgd = cfg.sync_mode in ("epoch-average", "replication", "none") or self.world == 1
        if cfg.sync_mode in ("epoch-average", "replication"):
            self.momentum.zero_()      # the reference builds a new optimizer every epoch (:187)
            self.first_step = True
        if cfg.step_path == "autograd":
            return self._train_epoch_autograd(shard, B, local_sgd, drop_at)
        if self.engine is not None:
            return self._train_epoch_cuda(shard, B, local_sgd, drop_at)
        return self._train_epoch_cpu(shard, B, local_sgd, drop_at)

    def _train_epoch_autograd(self, shard, B, local_sgd, drop_at) -> Tuple[float, int]:
        """The layer-at-a-time step: forward and backward through the autograd ops (layer kernels on CUDA), the
        gradient averaged bucket by bucket on a side stream while backward is still running
        (parallel/bucketed.py), one SGD-momentum kernel over the flat vector. Same arithmetic as the reference's
        loop body (data_parallelism_train.py:195-203) with the exchange moved from the epoch boundary into the step."""
        if drop_at is not None:
            raise ValueError("the drop-mode failure drill runs on the fused step path (leave markers live in its exchange)")
        from .parallel.bucketed import BucketedGradientAverager

        model, cfg = self.model, self.cfg
        model.train()
        avg = None
        if not local_sgd and self.world > 1:
            if self.engine is not None:
                self._align_sequence()
                self._apply_membership()
            if self._averager is None:
                self._averager = BucketedGradientAverager(
                    model, len(self.membership.active), engine=self.engine, allreduce_fn=self.comm.allreduce_sum_,
                    variant=self._standalone_variant() if self.engine is not None else "one-shot",
                    overlap=bool(cfg.overlap_exchange))
            avg = self._averager
        total = torch.zeros((), device=self.device)
        nb = 0
        for x, y in D.batches(shard, B, shuffle=True, generator=self.gen):
            model.zero_grad()
            loss = Fn.cross_entropy(model(x), y)
            if avg is not None:
                avg.arm()
            loss.backward()
            if avg is not None:
                avg.finish()
            Fn.sgd_momentum_(model.flat_params.data, model.flat_grads, self.momentum, cfg.lr, cfg.momentum, self.first_step)
            self.first_step = False
            total += loss.detach()
            nb += 1
        self.steps_this_epoch = nb
        return float(total), nb

    def _train_epoch_cuda(self, shard, B, local_sgd, drop_at) -> Tuple[float, int]:
        eng = self.engine
        self._apply_membership(local_only=local_sgd)
        eng.reset_loss()
        pinned = None
        steps = len(shard) // B
        x, y, tail_xy = D.epoch_tensor(shard, B, shuffle=True, generator=self.gen)
        run_steps = steps if drop_at is None else min(steps, drop_at)
        # exchange timing: the update kernel stamps the LAST steps of the epoch (at most 256), the rest runs untraced
        n_tail = min(run_steps, 256) if (self.cfg.exchange_timing and not local_sgd and self.world > 1
                                         and self.cfg.input_pipeline != "host") else 0
        if self.cfg.input_pipeline == "host":
            # per-step H2D copies from pinned memory through the engine's 32-slot ring
            if run_steps:
                # (kept alive until the synchronising read of the control block below: train_host returns while the
                # last copies are still in flight, and the copy stream is not one PyTorch's pinned allocator knows)
                pinned = (x[:run_steps].pin_memory(), y[:run_steps].pin_memory())
                eng.train_host(pinned[0], pinned[1], None)
        elif run_steps:
            if run_steps > n_tail:
                eng.train_device(x[:run_steps - n_tail], y[:run_steps - n_tail])
            if n_tail:
                if self._utrace is None:
                    self._utrace = torch.zeros(1024 * 122 * 8, dtype=torch.int64, device=self.device)
                eng.set_update_trace(self._utrace)
                eng.set_debug_flags(4)
                eng.train_device(x[run_steps - n_tail:run_steps], y[run_steps - n_tail:run_steps])
                eng.set_debug_flags(0)
                eng.set_update_trace(None)
                end = int(eng.read_ctl().step)
                u = self._utrace.view(1024, 122, 8)[[(end - n_tail + k) % 1024 for k in range(n_tail)]].cpu().numpy()
                wait = np.sort((u[:, :, 3].max(axis=1) - u[:, :, 2].max(axis=1)) / 1e3)
                self.last_exchange_us = float(wait[len(wait) // 2])
        if tail_xy is not None and drop_at is None:
            # the trailing short batch, as DataLoader(drop_last=False) yields it
            eng.step(tail_xy[0].to(self.device), tail_xy[1].to(self.device))
        self.steps_this_epoch = run_steps
        if drop_at is not None:
            eng.leave()
            self.dropped = True
        ctl = eng.read_ctl()
        pinned = None
        if ctl.status & 1:
            self._log(f"[rank {self.rank}] peers timed out during the epoch: mask 0x{(ctl.status >> 8) & 0xffff:x}")
            self.saw_timeout = True
        return float(ctl.loss_sum), int(ctl.loss_steps)

    def _train_epoch_cpu(self, shard, B, local_sgd, drop_at) -> Tuple[float, int]:
        model, cfg = self.model, self.cfg
        model.train()
        total, nb = 0.0, 0
        exch = torch.zeros(L.FLAT_NUMEL + 1)
        for i, (x, y) in enumerate(D.batches(shard, B, shuffle=True, generator=self.gen)):
            gone = drop_at is not None and i >= drop_at
            if not gone:
                model.zero_grad()
                loss = F.cross_entropy(model(x), y)
                loss.backward()
                total += loss.item()
                nb += 1
            if not local_sgd:
                # gloo stand-in for the peer-memory exchange: the extra element counts who is present,
                # so a dropped rank (zeros, count 0) changes the divisor exactly like a leave marker
                exch[:-1] = 0.0 if gone else model.flat_grads
                exch[-1] = 0.0 if gone else 1.0
                self.comm.allreduce_sum_(exch)
                if gone:
                    continue
                g = exch[:-1] / exch[-1].clamp_min(1.0)
            else:
                if gone:
                    break
                g = model.flat_grads
            Fn.sgd_momentum_(model.flat_params.data, g, self.momentum, cfg.lr, cfg.momentum, self.first_step)
            self.first_step = False
        self.steps_this_epoch = nb
        if drop_at is not None:
            self.dropped = True
        return total, nb

    def _align_sequence(self) -> None:
        """Ranks that took a different number of local steps (or none: the parent) carry different
        exchange sequence numbers; agree on the largest one before the next collective exchange."""
        if self.engine is None or self.world == 1:
            return
        steps = self.comm.allgather(int(self.engine.read_ctl().step))
        self.engine.rejoin(max(steps))
        self._sync()
        self.comm.barrier()

    # ------------------------------------------------------------------ parameter averaging (R1)
    @torch.no_grad()
    def average_parameters(self) -> None:
        """mean over the trainers of the flat parameter vector; everyone (parent included) gets it."""
        if self.world == 1:
            return
        n_tr = len(self.membership.active)
        if self.engine is not None:
            self._align_sequence()
            contrib = self.model.flat_params if (self.is_trainer and self.rank in self.membership.active) \
                else torch.zeros_like(self.model.flat_params)
            full = (1 << self.world) - 1
            self.engine.set_membership(full, int(self.cfg.peer_timeout_s * 1e9))
            self.engine.set_divisor(n_tr)
            out = torch.empty_like(self.model.flat_params)
            self.engine.allreduce(contrib.contiguous(), out, mode=0, two_shot=self._standalone_variant())
            self.model.flat_params.copy_(out)
            self.engine.set_divisor(0)
            self._sync()
        else:
            buf = self.model.flat_params.detach().clone() if (self.is_trainer and self.rank in self.membership.active) \
                else torch.zeros(L.FLAT_NUMEL)
            self.comm.allreduce_sum_(buf)
            self.model.flat_params.copy_(buf / float(n_tr))

    # ------------------------------------------------------------------ evaluation
    @torch.no_grad()
    def evaluate(self, test: D.TensorDataset) -> Tuple[float, float]:
        """(mean of per-batch mean losses, accuracy %) -- reference eval :157-183."""
        bs = self.cfg.eval_batch_size or self.cfg.batch_size
        if self.engine is not None:
            loss, hits = self.engine.evaluate(test.images, test.labels, bs)
            return loss, 100.0 * hits / len(test)
        self.model.eval()
        losses, corrects, total = [], 0, 0
        for x, y in D.batches(test, bs, shuffle=False):
            out = self.model(x)
            losses.append(F.cross_entropy(out, y).item())
            corrects += (torch.argmax(out, dim=1) == y).sum().item()
            total += y.size(0)
        return float(np.mean(losses)), 100.0 * corrects / total

    # ------------------------------------------------------------------ membership at epoch boundaries
    def _epoch_boundary(self) -> None:
        """Membership epoch: agree on who is present; returning ranks adopt a survivor's state."""
        if self.world == 1:
            return
        status = self.comm.allgather({"rank": self.rank, "dropped": self.dropped, "timeout": self.saw_timeout,
                                      "steps": self.steps_this_epoch, "trainer": self.is_trainer})
        if any(s["timeout"] for s in status) and self.cfg.sync_mode == "step-allreduce":
            # A peer was given up on mid-epoch (unannounced stall): different ranks may have summed different sets
            # for some slices, so the replicas can differ in the last bits. Re-establish identical replicas by
            # averaging parameters and momentum over the trainers, and clear the sticky status.
            self._log(f"[rank {self.rank}] re-synchronising replicas after a peer timeout")
            self._resync_state()
            self.saw_timeout = False
        dropped = [s["rank"] for s in status if s["dropped"]]
        if not dropped:
            return
        # whose state survives: a trainer that stayed all epoch, else (everybody dropped) the one that trained longest
        donor = self.membership.donor({s["rank"]: (s["dropped"], s["steps"]) for s in status if s["trainer"]})
        if donor in dropped:
            self._log(f"[rank {self.rank}] every trainer dropped out this epoch; rank {donor} trained longest and "
                      f"becomes the donor")
        came_back = [r for r in dropped if r != donor]
        for r in came_back:
            self.membership.leave(r)
        state = None
        if self.rank == donor:
            step = self.engine.read_ctl().step if self.engine is not None else 0
            state = (self.model.flat_params.detach().cpu(), self.momentum.detach().cpu(), int(step))
        state = self.comm.bcast(state, root=donor)
        if self.dropped and self.rank != donor:
            self.model.flat_params.copy_(state[0].to(self.device))
            self.momentum.copy_(state[1].to(self.device))
            if self.engine is not None:
                self.engine.rejoin(state[2])
                self._sync()
        if self.dropped and self.rank == donor and self.engine is not None:
            self.engine.rejoin(state[2])   # clears the donor's own leave marker; its step counter is the reference
            self._sync()
        self.dropped = False
        for r in came_back:
            self.membership.join(r)
        self.comm.barrier()  # nobody steps before every marker is cleared
        self._apply_membership()

    @torch.no_grad()
    def _resync_state(self) -> None:
        for buf in (self.model.flat_params, self.momentum):
            if self.engine is not None:
                self._align_sequence()
                full = (1 << self.world) - 1
                self.engine.set_membership(full, int(self.cfg.peer_timeout_s * 1e9))
                self.engine.set_divisor(len(self.trainers))
                contrib = buf.clone() if self.is_trainer else torch.zeros_like(buf)
                out = torch.empty_like(buf)
                self.engine.allreduce(contrib, out, mode=0, two_shot=False)
                buf.copy_(out)
                self.engine.set_divisor(0)
                self._sync()
            else:
                t = buf.detach().clone() if self.is_trainer else torch.zeros_like(buf)
                self.comm.allreduce_sum_(t)
                buf.copy_(t / float(len(self.trainers)))
        self._apply_membership()

    # ------------------------------------------------------------------ the whole run
    def fit(self, shard: Optional[D.TensorDataset] = None, test: Optional[D.TensorDataset] = None) -> List[Dict]:
        """Trains ``cfg.epochs`` epochs. A rank that is really gone (process dead, not the simulated failure) ends the
        job on the survivors with ``bootstrap.PeerLost`` instead of the reference's endless recv
        (data_parallelism_train.py:118,135); with ``cfg.save`` set each survivor first writes what it has to
        ``<save>.peerlost.rank<r>`` (epoch = last completed one), which ``--resume`` accepts."""
        try:
            return self._fit(shard, test)
        except bootstrap.PeerLost as e:
            self._log(f"Process {self.rank}: {e}")
            if self.cfg.save:
                done = self.history[-1]["epoch"] if self.history else self.start_epoch - 1
                path = f"{self.cfg.save}.peerlost.rank{self.rank}"
                try:
                    self._sync()
                    ckpt.save(path, self.model, self.momentum, epoch=done, membership_epoch=self.membership.epoch)
                    self._log(f"Process {self.rank}: state after epoch {done} saved to {path}")
                except Exception as save_err:  # noqa: BLE001 -- the PeerLost is what the caller must see
                    self._log(f"Process {self.rank}: could not save {path}: {save_err}")
            raise

    def _fit(self, shard: Optional[D.TensorDataset] = None, test: Optional[D.TensorDataset] = None) -> List[Dict]:
        cfg = self.cfg
        if shard is None and test is None:
            shard, test = self.load_data()
        if self.rank == 0:
            self.run["parameters"] = {"learning_rate": cfg.lr, "optimizer": "SGD", "model_name": "distmodel",
                                      "epochs": cfg.epochs, "sync_mode": cfg.sync_mode, "world": self.world}
        steps_per_epoch = (len(shard) + cfg.batch_size - 1) // cfg.batch_size if shard is not None else 0
        for epoch in range(self.start_epoch, cfg.epochs):
            if self.rank == 0:
                self._log("Starting epoch ", epoch)
            with nvtx_range(f"membership/epoch {epoch}"):
                self._epoch_boundary()
            ev = self.failure.draw(steps_per_epoch) if cfg.failure_probability > 0 else None
            drop_at = None
            if ev is not None and cfg.failure_mode == "sleep":
                # parent: inside the communication timer, children: untimed (reference :116-120,141)
                if self.parent_only:
                    with self.timers.add("mpi_communication_time_parent"):
                        self.failure.sleep(ev)
                else:
                    self.failure.sleep(ev)
            elif ev is not None and not (self.is_trainer and cfg.sync_mode == "step-allreduce"
                                         and len(self.membership.active) > 1):
                self.failure.sleep(ev)  # nothing to drop out of: behave like the reference
            elif ev is not None:
                drop_at = ev.at_step
                self._log(f"Process {self.rank} failed! Leaving the group at step {drop_at} for {ev.duration} seconds.")

            total_loss, n_batches = 0.0, 0
            t_dev = None
            if self.is_trainer:
                ev0 = ev1 = None
                if self.device.type == "cuda":
                    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                with self.timers.add("training_time", self._sync), nvtx_range(f"train/epoch {epoch}"):
                    if ev0 is not None:
                        ev0.record()
                    total_loss, n_batches = self.train_epoch(shard, drop_at)
                    if ev1 is not None:
                        ev1.record()
                if ev0 is not None:
                    self._sync()
                    t_dev = ev0.elapsed_time(ev1) / 1e3
                if drop_at is not None:
                    time.sleep(ev.duration)
                    self._log(f"Process {self.rank} woke up!")

            # parity modes average once per epoch; with a non-training parent the parent also needs the
            # trainers' (identical) weights in step-allreduce mode before it can evaluate
            if (cfg.sync_mode in ("epoch-average", "replication") or
                    (cfg.sync_mode == "step-allreduce" and cfg.reference_topology)) and self.world > 1:
                if self.rank == 0:
                    self._log("* Averaging models...")
                tname = "mpi_communication_time_parent" if self.rank == 0 else "mpi_communication_time_children"
                with self.timers.add(tname, self._sync), nvtx_range(f"average/epoch {epoch}"):
                    self.average_parameters()

            # loss bookkeeping: true mean over all batches of all trainers
            stats = self.comm.allgather((total_loss, n_batches, t_dev, len(shard) if shard is not None else 0,
                                         self.last_exchange_us))
            sum_loss = sum(s[0] for s in stats)
            sum_batches = sum(s[1] for s in stats)
            mean_loss = sum_loss / max(1, sum_batches)
            rec: Dict = {"epoch": epoch, "train_loss": mean_loss, "membership_epoch": self.membership.epoch,
                         "active": sorted(self.membership.active)}
            tds = [s[2] for s in stats if s[2] is not None]
            if tds:
                t_max = max(tds)                      # device-timed, max over ranks
                imgs = sum(s[3] for s in stats)
                rec.update(train_seconds_device_max=t_max, images_per_s=imgs / t_max if t_max > 0 else None,
                           step_us=1e6 * t_max / max(1, max(s[1] for s in stats)))
            xus = [s[4] for s in stats if s[4] is not None]
            if xus and cfg.sync_mode == "step-allreduce":
                # exposed gradient exchange per step (max over ranks of the per-rank median) and what fraction of the
                # NVLink roofline that is: bytes per direction at 900 GB/s over the measured time
                variant = self._exchange_variant()
                roof_us = exchange_bytes_per_direction(variant, len(self.membership.active)) / NVLINK_BYTES_PER_S * 1e6
                rec.update(allreduce_us=max(xus), exchange_variant=[k for k, v in native.EXCHANGE_VARIANTS.items()
                                                                    if v == variant][0],
                           exchange_roofline_frac=roof_us / max(xus) if max(xus) > 0 else None)
            if self.rank == 0:
                self._log(f"Global Average Training Loss: {mean_loss}")
                self.run["train/loss"].append(mean_loss)
                self._log("evaluating model")
                with self.timers.add("evaluation_time", self._sync), nvtx_range(f"eval/epoch {epoch}"):
                    val_loss, val_acc = self.evaluate(test)
                self.run["val/loss"].append(val_loss)
                self.run["val/acc"].append(val_acc)
                self._log("Validation loss of updated master model: ", val_loss)
                self._log("Validation Accuracy: %.2f %%" % val_acc)
                rec.update(val_loss=val_loss, val_acc=val_acc)
                self.jsonl.write(**rec)
            self.history.append(rec)
            if cfg.save and self.rank == 0:
                ckpt.save(cfg.save, self.model, self.momentum, epoch=epoch, membership_epoch=self.membership.epoch)
            self.comm.barrier()  # children wait for the parent's evaluation, as they do in recv (:135)
        self._epoch_boundary()   # ranks that dropped out in the last epoch still end with the final model
        self._write_logs()
        return self.history

    def _write_logs(self) -> None:
        cfg = self.cfg
        if self.rank == 0:
            logs.print_parent_summary(self.timers)
            logs.write_parent_log(self.timers, cfg.batch_size, cfg.epochs, cfg.nb_proc, cfg.log_dir)
        # the reference writes the children file from rank 2 only (:143); fall back to the last trainer
        writer = 2 if 2 in self.trainers else self.trainers[-1]
        if self.rank == writer and (self.world > 1):
            logs.print_children_summary(self.timers)
            logs.write_children_log(self.timers, cfg.batch_size, cfg.epochs, cfg.nb_proc, cfg.log_dir)

    def close(self) -> None:
        if self.engine is not None:
            self._sync()
            self.engine.close()
            self.engine = None
        if self.group is not None:
            self.comm.barrier()                 # nobody stores into a peer's (or the multicast) mapping any more
            if getattr(self, "nvls_buf", None) is not None:
                self.nvls_buf.close()
                self.nvls_buf = None
            self.group.close()
            self.group = None
        try:
            self.run.stop()
        except Exception:
            pass
