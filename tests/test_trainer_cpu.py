"""Trainer on CPU: single process, and world_size > 1 over gloo through the bundled launcher
(SURVEY §4 'collective (fake backend)' and 'end-to-end' rows; BASELINE.json configs 1 and 2)."""
import copy
import json
import os
import subprocess
import sys
from pathlib import Path

import pytest
import torch
import torch.nn.functional as F

from distributed_neural_network_b200 import Network
from distributed_neural_network_b200.parallel import bootstrap
from distributed_neural_network_b200.parallel.membership import FailureSimulator, Membership
from distributed_neural_network_b200.trainer import TrainConfig, Trainer, reference_global_avg_loss
from distributed_neural_network_b200.utils import data as D

ROOT = Path(__file__).resolve().parent.parent


def _launch(n, out_dir, mode, **kw):
    from distributed_neural_network_b200.parallel import launch
    env = {"OMP_NUM_THREADS": "1"}
    rc = launch.spawn(n, [sys.executable, str(ROOT / "tests" / "_mp_worker.py"), str(out_dir), mode, json.dumps(kw)],
                      env_extra=env, gpus=0)
    assert rc == 0
    return [torch.load(f"{out_dir}/rank{r}.pt", weights_only=False) for r in range(n)]


def test_single_process_matches_plain_pytorch_loop():
    """Same seed, same batches: the Trainer's CPU path is the reference's loop (single_proc_train.py:57-75)."""
    cfg = TrainConfig(device="cpu", synthetic=True, train_size=96, test_size=32, epochs=2, batch_size=16, lr=0.05,
                      sync_mode="none", seed=5, quiet=True, log_dir="/tmp/_dnn_test_log")
    tr = Trainer(cfg, comm=bootstrap.Comm())
    ref = copy.deepcopy(tr.model)
    opt = torch.optim.SGD(ref.parameters(), lr=0.05, momentum=0.9)
    shard, test = tr.load_data()
    gen = copy.deepcopy(tr.gen)
    hist = tr.fit(shard, test)
    for _ in range(2):
        for x, y in D.batches(shard, 16, shuffle=True, generator=gen):
            ref.zero_grad()
            F.cross_entropy(ref(x), y).backward()
            opt.step()
    assert torch.allclose(tr.model.flat_params, ref.flat_params, atol=1e-6)
    assert len(hist) == 2 and 0 <= hist[-1]["val_acc"] <= 100


def test_training_learns_on_synthetic_data():
    cfg = TrainConfig(device="cpu", synthetic=True, train_size=512, test_size=128, epochs=4, batch_size=16, lr=0.05,
                      sync_mode="none", seed=0, quiet=True, log_dir="/tmp/_dnn_test_log")
    tr = Trainer(cfg, comm=bootstrap.Comm())
    hist = tr.fit()
    assert hist[-1]["train_loss"] < hist[0]["train_loss"] - 0.3
    assert hist[-1]["val_acc"] > 50


def test_step_allreduce_world2_equals_big_batch(tmp_path):
    res = _launch(2, tmp_path, "step-allreduce", train_size=128, test_size=32, epochs=1, batch_size=8, lr=0.05, seed=11)
    assert torch.equal(res[0]["flat"], res[1]["flat"]), "replicas must stay bit-identical"
    # oracle: one process, batch 16 = the two ranks' batches concatenated, same shuffles
    cfg = TrainConfig(device="cpu", synthetic=True, train_size=128, test_size=32, epochs=1, batch_size=8, lr=0.05,
                      seed=11, quiet=True, sync_mode="none", log_dir=str(tmp_path / "l"))
    torch.manual_seed(11)          # rank 0's seed: its initial weights are broadcast to everyone
    model = Network()
    full = D.synthetic(128, seed=0)
    shards = [D.partition_symmetric(full, r, 2).materialize() for r in range(2)]
    gens = [torch.Generator().manual_seed(11 + 1000 * r) for r in range(2)]
    opt = torch.optim.SGD(model.parameters(), lr=0.05, momentum=0.9)
    its = [D.batches(shards[r], 8, True, gens[r]) for r in range(2)]
    for (x0, y0), (x1, y1) in zip(*its):
        model.zero_grad()
        F.cross_entropy(model(torch.cat([x0, x1])), torch.cat([y0, y1])).backward()
        opt.step()
    assert torch.allclose(res[0]["flat"], model.flat_params, atol=2e-6)


def test_epoch_average_reference_topology_world3(tmp_path):
    """The reference's algorithm: rank 0 parent, children train contiguous shards with a fresh optimizer,
    parameters are averaged once per epoch (data_parallelism_train.py:114-121,134-142,185-244)."""
    res = _launch(3, tmp_path, "epoch-average", reference_topology=True, train_size=96, test_size=32, epochs=2,
                  batch_size=8, lr=0.05, seed=21, nb_proc=3)
    for r in (1, 2):
        assert torch.equal(res[0]["flat"], res[r]["flat"])
    torch.manual_seed(21)
    master = Network()
    full = D.synthetic(96, seed=0)
    gens = {r: torch.Generator().manual_seed(21 + 1000 * r) for r in (1, 2)}
    for _ in range(2):
        kids = []
        for r in (1, 2):
            child = copy.deepcopy(master)
            opt = torch.optim.SGD(child.parameters(), lr=0.05, momentum=0.9)   # new optimizer each epoch (:187)
            shard = D.partition_dataset(full, r, 3).materialize()
            for x, y in D.batches(shard, 8, True, gens[r]):
                child.zero_grad()
                F.cross_entropy(child(x), y).backward()
                opt.step()
            kids.append(child.flat_params.clone())
        master.flat_params.copy_(sum(kids) / float(len(kids)))                  # :238-240
    assert torch.allclose(res[0]["flat"], master.flat_params, atol=2e-6)
    # timing logs: parent file from rank 0, children file from rank 2 (:126-129,143-152)
    names = sorted(os.listdir(tmp_path / "log"))
    assert names == ["bs8_log_epochs2_proc3_children.txt", "bs8_log_epochs2_proc3_parent.txt"]
    assert res[0]["timers"]["training_time"] == 0.0 and res[2]["timers"]["training_time"] > 0.0


def test_replication_mode_world3(tmp_path):
    res = _launch(3, tmp_path, "replication", reference_topology=True, train_size=64, test_size=32, epochs=1,
                  batch_size=16, lr=0.05, seed=4)
    assert torch.equal(res[0]["flat"], res[1]["flat"]) and torch.equal(res[1]["flat"], res[2]["flat"])
    assert res[0]["history"][0]["train_loss"] > 0


def test_drop_mode_survivors_continue_world4(tmp_path):
    """Ranks leave mid-epoch with high probability; nobody hangs, everybody agrees afterwards."""
    res = _launch(4, tmp_path, "step-allreduce", train_size=256, test_size=32, epochs=3, batch_size=8, lr=0.02,
                  seed=9, failure_probability=0.6, failure_duration=0.05, failure_mode="drop")
    h = res[0]["history"]
    assert len(h) == 3 and h[-1]["membership_epoch"] > 0      # membership actually changed
    for r in range(1, 4):
        assert torch.equal(res[0]["flat"], res[r]["flat"])    # returning ranks adopted the survivors' state
    assert all(torch.isfinite(r["flat"]).all() for r in res)


def test_membership_and_failure_simulator():
    m = Membership(4)
    assert m.mask == 0b1111 and m.divisor == 4 and m.epoch == 0
    assert m.leave(2) and m.mask == 0b1011 and m.divisor == 3 and m.epoch == 1
    assert not m.leave(2) and m.donor() == 0
    assert m.join(2) and m.epoch == 2 and m.divisor == 4
    for r in (0, 1, 3):
        m.leave(r)
    assert not m.leave(2), "the last member cannot leave"
    sim = FailureSimulator(1.0, 0.0, "drop", seed=1, rank=3)
    ev = sim.draw(10)
    assert ev is not None and ev.rank == 3 and 0 <= ev.at_step < 10
    assert FailureSimulator(0.0, 1.0, "sleep", seed=1).draw() is None
    with pytest.raises(ValueError):
        FailureSimulator(0.1, 1.0, "explode")


def test_rank_discovery_from_launcher_environments(monkeypatch):
    for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK", "OMPI_COMM_WORLD_RANK", "OMPI_COMM_WORLD_SIZE", "PMI_RANK",
              "PMI_SIZE", "SLURM_PROCID", "SLURM_NTASKS"):
        monkeypatch.delenv(k, raising=False)
    assert bootstrap.discover() == (0, 1, 0)
    monkeypatch.setenv("OMPI_COMM_WORLD_RANK", "2")
    monkeypatch.setenv("OMPI_COMM_WORLD_SIZE", "4")
    monkeypatch.setenv("OMPI_COMM_WORLD_LOCAL_RANK", "1")
    assert bootstrap.discover() == (2, 4, 1)
    monkeypatch.setenv("RANK", "3")
    monkeypatch.setenv("WORLD_SIZE", "8")
    monkeypatch.setenv("LOCAL_RANK", "3")
    assert bootstrap.discover() == (3, 8, 3)


def test_reference_loss_formula_kept_for_parity():
    assert reference_global_avg_loss(60.0, 3) == 2.0     # sum of batch losses / (10 keys * children), quirk Q4


def test_mpiexec_front_end_runs_the_documented_command_line(tmp_path):
    out = subprocess.run([str(ROOT / "bin" / "mpiexec"), "-n", "2", sys.executable,
                          str(ROOT / "data_parallelism_train.py"), "--nb-proc", "2", "--batch-size", "8", "--epochs", "1",
                          "--synthetic", "--train-size", "64", "--test-size", "16", "--device", "cpu", "--log-dir",
                          str(tmp_path / "log"), "--save", str(tmp_path / "model.pt")],
                         capture_output=True, text=True, cwd=ROOT, timeout=300)
    assert out.returncode == 0, out.stderr[-2000:]
    assert "Starting epoch  0" in out.stdout and "Validation loss of updated master model" in out.stdout
    assert (tmp_path / "log" / "bs8_log_epochs1_proc2_parent.txt").exists()
    sd = torch.load(tmp_path / "model.pt", weights_only=True)
    assert len(sd) == 10


def test_single_proc_and_model_replication_entry_points(tmp_path):
    """The other two reference command lines (single_proc_train.py:16-107, model_replication_train.py:153-160)."""
    common = ["--synthetic", "--train-size", "64", "--test-size", "16", "--device", "cpu", "--epochs", "1",
              "--log-dir", str(tmp_path / "log")]
    out = subprocess.run([sys.executable, str(ROOT / "single_proc_train.py"), "--batch-size", "4", *common,
                          "--jsonl", str(tmp_path / "single.jsonl")], capture_output=True, text=True, cwd=ROOT,
                         timeout=300)
    assert out.returncode == 0, out.stderr[-2000:]
    rec = json.loads((tmp_path / "single.jsonl").read_text().splitlines()[-1])
    assert rec["epoch"] == 0 and rec["train_loss"] > 0 and 0.0 <= rec["val_acc"] <= 100.0
    out = subprocess.run([str(ROOT / "bin" / "mpiexec"), "-n", "3", sys.executable,
                          str(ROOT / "model_replication_train.py"), "--batch-size", "8", *common],
                         capture_output=True, text=True, cwd=ROOT, timeout=300)
    assert out.returncode == 0, out.stderr[-2000:]
    assert "Validation loss of updated master model" in out.stdout


def test_run_training_sh_issues_the_reference_sweep(tmp_path):
    """run_training.sh:1-4 -- batch sizes 1..64 at 4 processes (and, new, 8) -- checked with a recording mpiexec."""
    fake = tmp_path / "mpiexec"
    fake.write_text("#!/bin/sh\nprintf '%%s\\n' \"$*\" >> \"%s\"\n" % (tmp_path / "calls.txt"))
    fake.chmod(0o755)
    env = dict(os.environ, PATH=f"{tmp_path}:{os.environ['PATH']}")
    out = subprocess.run(["sh", str(ROOT / "run_training.sh"), "--synthetic"], capture_output=True, text=True, cwd=ROOT,
                         env=env, timeout=60)
    assert out.returncode == 0, out.stderr
    calls = (tmp_path / "calls.txt").read_text().splitlines()
    want = [f"-n {n} python data_parallelism_train.py --nb-proc {n} --batch-size {bs} --synthetic"
            for n in (4, 8) for bs in (1, 2, 4, 8, 16, 32, 64)]
    assert calls == want


def test_mpi4py_control_plane_world2_equals_gloo(tmp_path):
    """The mpi4py provider of the control plane (bootstrap._MpiComm: COMM_WORLD, Barrier, allgather, bcast,
    Allreduce -- what the reference bootstraps with, data_parallelism_train.py:60-62). The image has no MPI, so a
    socket-backed test double of mpi4py (baseline/shims) stands in; launched the way mpiexec would launch it
    (OMPI_COMM_WORLD_* in the environment), the same run must give the same bits as over torch.distributed/gloo."""
    import socket

    from distributed_neural_network_b200.parallel import launch
    kw = dict(train_size=128, test_size=32, epochs=1, batch_size=8, lr=0.05, seed=11)
    gloo = _launch(2, tmp_path / "gloo", "step-allreduce", **kw)
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    out = tmp_path / "mpi"
    out.mkdir()
    probe = out / "provider"
    env = {"OMP_NUM_THREADS": "1", "OMPI_COMM_WORLD_SIZE": "2", "REFSHIM_PORT": str(port),
           "PYTHONPATH": str(ROOT / "baseline" / "shims") + os.pathsep + os.environ.get("PYTHONPATH", ""),
           "DNNB200_PROVIDER_PROBE": str(probe)}
    rc = launch.spawn(2, [sys.executable, str(ROOT / "tests" / "_mp_worker.py"), str(out), "step-allreduce",
                          json.dumps(dict(kw, device="cpu"))], env_extra=env, gpus=0)
    assert rc == 0
    mpi = [torch.load(out / f"rank{r}.pt", weights_only=False) for r in range(2)]
    assert (probe.parent / "provider.0").read_text() == "mpi4py" and (probe.parent / "provider.1").read_text() == "mpi4py"
    assert torch.equal(mpi[0]["flat"], mpi[1]["flat"])
    assert torch.equal(mpi[0]["flat"], gloo[0]["flat"]), "the control plane must not change the numbers"


def test_buckets_cover_the_flat_vector_and_leave_in_backward_order():
    """The three gradient buckets tile the flat vector, and autograd finishes them in the order fc2+fc3, fc1, conv --
    so the first two exchanges can run while the convolution backward is still computing."""
    from distributed_neural_network_b200.ops import flat as L
    from distributed_neural_network_b200.parallel.bucketed import BUCKETS, BucketedGradientAverager

    spans = sorted((lo, hi) for _, lo, hi, _ in BUCKETS)
    assert spans[0][0] == 0 and spans[-1][1] == L.FLAT_NUMEL
    assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
    assert sorted(n for _, _, _, names in BUCKETS for n in names) == sorted(L.PARAM_SHAPES)
    torch.manual_seed(0)
    model = Network()
    seen = []
    avg = BucketedGradientAverager(model, world=2, allreduce_fn=lambda t: seen.append(t.numel()) or t.mul_(2.0))
    x, y = torch.randn(8, 3, 32, 32), torch.randint(0, 10, (8,))
    ref = copy.deepcopy(model)
    ref.zero_grad()
    F.cross_entropy(ref(x), y).backward()
    model.zero_grad()
    avg.arm()
    F.cross_entropy(model(x), y).backward()
    assert avg.launched == ["fc2+fc3", "fc1", "conv"]          # all three left DURING backward, in this order
    avg.finish()
    assert seen == [hi - lo for _, lo, hi, _ in BUCKETS]
    assert torch.allclose(model.flat_grads, ref.flat_grads)    # "sum" of two identical ranks / 2 = the gradient
    avg.close()


def test_autograd_step_path_world2_equals_fused_semantics(tmp_path):
    """--step-path autograd (layer ops through autograd, per-bucket exchange) trains the same model as the
    step-allreduce reference semantics: world 2 x batch 8 == the CPU run of the default path."""
    kw = dict(train_size=128, test_size=32, epochs=1, batch_size=8, lr=0.05, seed=11)
    a = _launch(2, tmp_path / "fused", "step-allreduce", **kw)
    b = _launch(2, tmp_path / "autograd", "step-allreduce", step_path="autograd", **kw)
    assert torch.equal(b[0]["flat"], b[1]["flat"])
    assert torch.allclose(a[0]["flat"], b[0]["flat"], atol=2e-6)


def test_share_fd_hands_a_descriptor_from_rank0_to_every_process(tmp_path):
    """The rendezvous under the NVLS group (csrc/nvls_group.cu): SCM_RIGHTS over an abstract Unix socket. The clients
    start BEFORE the server here, so the connect retry is exercised too; a missing rank must end in a timeout error."""
    import time as _time
    from distributed_neural_network_b200.ops import native
    if not native.available():
        pytest.skip("native runtime not built")
    secret = tmp_path / "secret.txt"
    secret.write_text("handed over by the kernel")
    name = f"t-{os.getpid()}-{_time.time_ns():x}"
    worker = str(Path(__file__).parent / "_mp_sharefd_worker.py")
    procs = [subprocess.Popen([sys.executable, worker, str(r), "3", name, str(secret), str(tmp_path)]) for r in (2, 1)]
    _time.sleep(0.3)
    procs.append(subprocess.Popen([sys.executable, worker, "0", "3", name, str(secret), str(tmp_path)]))
    assert [p.wait(timeout=120) for p in procs] == [0, 0, 0]
    for r in range(3):
        assert (tmp_path / f"rank{r}.txt").read_text() == "c0ffee True handed over by the kernel"
    with pytest.raises(RuntimeError, match="did not get the descriptor"):
        native.share_fd(1, 2, name + "-nobody", timeout_s=0.2)


def test_a_rank_that_really_dies_ends_in_peer_lost_not_in_a_hang(tmp_path):
    """the reference blocks for ever in recv when a process is gone. Here the control plane's collectives
    carry a timeout (DNNB200_COMM_TIMEOUT_S) and surface PeerLost on the survivors."""
    from distributed_neural_network_b200.parallel import launch
    out = tmp_path / "verdict.txt"
    rc = launch.spawn(2, [sys.executable, str(Path(__file__).parent / "_mp_peerlost_worker.py"), str(out)],
                      env_extra={"DNNB200_COMM_TIMEOUT_S": "5", "CUDA_VISIBLE_DEVICES": ""}, gpus=0)
    assert rc == 0
    verdict = out.read_text()
    assert verdict.startswith("PeerLost after"), verdict
    assert float(verdict.split()[2].rstrip("s:")) < 30.0, verdict


def test_fit_saves_what_it_has_when_a_peer_is_lost(tmp_path):
    """Trainer.fit on PeerLost: emergency checkpoint named after the rank, epoch = last completed, error re-raised;
    the file resumes like any other checkpoint."""
    save = str(tmp_path / "ck.pt")
    cfg = TrainConfig(device="cpu", synthetic=True, train_size=64, test_size=32, epochs=3, batch_size=16, quiet=True,
                      sync_mode="none", seed=2, save=save, log_dir=str(tmp_path / "log"))
    tr = Trainer(cfg)
    calls = {"n": 0}
    real = tr.comm.allgather

    def flaky(obj):
        calls["n"] += 1
        if len(tr.history) == 1 and isinstance(obj, tuple):      # the loss bookkeeping of the SECOND epoch
            raise bootstrap.PeerLost("control plane: all-gather failed (test)")
        return real(obj)

    tr.comm.allgather = flaky
    with pytest.raises(bootstrap.PeerLost):
        tr.fit()
    lost = Path(save + ".peerlost.rank0")
    assert lost.exists() and len(tr.history) == 1
    after = tr.model.flat_params.detach().clone()
    tr.close()
    tr2 = Trainer(TrainConfig(device="cpu", synthetic=True, train_size=64, test_size=32, epochs=3, batch_size=16,
                              quiet=True, sync_mode="none", seed=2, resume=str(lost), log_dir=str(tmp_path / "log2")))
    assert tr2.start_epoch == 1 and torch.equal(tr2.model.flat_params.detach(), after)
    tr2.close()


class _ThreadComm:
    """Control plane for ranks that are threads of this process (tests of collective choreography)."""

    def __init__(self, rank, world, shared):
        self.rank, self.world, self.local_rank, self._s = rank, world, rank, shared

    def allgather(self, obj):
        s = self._s
        s["slots"][self.rank] = obj
        s["bar"].wait(30)
        out = list(s["slots"])
        s["bar"].wait(30)
        return out

    def bcast(self, obj, root=0):
        return self.allgather(obj)[root]

    def barrier(self):
        self.allgather(None)


def _run_nvls_choreography(monkeypatch, world, hosts, fail_open=(), fail_map=()):
    import threading

    from distributed_neural_network_b200.ops import native
    from distributed_neural_network_b200.parallel import nvls
    events, lock = [], threading.Lock()

    class FakeGroup:
        def __init__(self, rank, world_, idx, nbytes, name, timeout_s):
            if rank in fail_open:
                raise RuntimeError(f"open failed on {rank}")
            self.rank, self.name, self.multicast_ptr, self.nbytes = rank, name, 0, nbytes
            with lock:
                events.append(("open", rank))

        def map(self):
            if self.rank in fail_map:
                raise RuntimeError(f"map failed on {self.rank}")
            with lock:
                events.append(("map", self.rank))
            self.multicast_ptr = 0xABC000
            return 0x123000, self.multicast_ptr

        def close(self):
            with lock:
                events.append(("close", self.rank))

    monkeypatch.setattr(native, "NvlsGroup", FakeGroup)
    monkeypatch.setattr(native, "nvls_supported", lambda idx: True)
    shared = {"slots": [None] * world, "bar": threading.Barrier(world)}
    names = iter(hosts)
    host_of = {r: next(names) for r in range(world)}
    monkeypatch.setattr(nvls.socket, "gethostname", lambda: host_of[int(threading.current_thread().name)])
    out = {}

    def run(r):
        out[r] = nvls.try_create(torch.device("cuda", r), _ThreadComm(r, world, shared))

    ts = [threading.Thread(target=run, args=(r,), name=str(r)) for r in range(world)]
    for t in ts:
        t.start()
    for t in ts:
        t.join(60)
    return out, events


def test_nvls_group_is_made_by_all_ranks_or_by_nobody(monkeypatch):
    """parallel/nvls.py: every rank opens, THEN every rank maps (the driver refuses to bind before all devices are in);
    one failure anywhere and every rank closes what it has and reports 'no NVLS'; ranks on two hosts do not even try."""
    out, ev = _run_nvls_choreography(monkeypatch, 4, ["a"] * 4)
    assert all(o is not None and o[1] == 0xABC000 for o in out.values())
    assert len({o[0].name for o in out.values()}) == 1                         # one job-unique name, from rank 0
    assert max(i for i, e in enumerate(ev) if e[0] == "open") < min(i for i, e in enumerate(ev) if e[0] == "map")
    out, ev = _run_nvls_choreography(monkeypatch, 4, ["a"] * 4, fail_open={2})
    assert all(o is None for o in out.values())
    assert sorted(e for e in ev if e[0] == "close") == [("close", 0), ("close", 1), ("close", 3)] and not [e for e in ev if e[0] == "map"]
    out, ev = _run_nvls_choreography(monkeypatch, 3, ["a"] * 3, fail_map={0})
    assert all(o is None for o in out.values()) and sorted(e for e in ev if e[0] == "close") == [("close", r) for r in range(3)]
    out, ev = _run_nvls_choreography(monkeypatch, 2, ["a", "b"])
    assert all(o is None for o in out.values()) and ev == []


def test_exchange_variants_offered_depend_on_probes_and_on_the_failure_mode():
    """Trainer._usable_variants / _exchange_variant / _probe_ll128 without a GPU: the 128-byte lines need a passed probe
    on EVERY rank, two-shot and NVLS a membership that cannot change, a forced variant that is not usable falls back to
    the one that always is."""
    from types import SimpleNamespace

    from distributed_neural_network_b200.ops import native
    from distributed_neural_network_b200.parallel.membership import Membership

    def make(world=4, ll128=True, nvls=True, mode="sleep", p=0.0, variant="auto", choice=None, active=None):
        t = Trainer.__new__(Trainer)
        t.world, t.rank, t.ll128_ok, t.nvls_ok, t._nvls_probed = world, 0, ll128, nvls, True
        t.cfg = TrainConfig(failure_mode=mode, failure_probability=p, allreduce_variant=variant)
        t.membership = Membership(world, active=set(active if active is not None else range(world)))
        t.exchange_choice = choice
        return t

    assert make()._usable_variants() == ["one-shot", "one-shot-128", "two-shot", "nvls"]
    assert make(ll128=False)._usable_variants() == ["one-shot", "two-shot", "nvls"]
    assert make(nvls=False)._usable_variants() == ["one-shot", "one-shot-128", "two-shot"]
    assert make(active=[1, 2, 3])._usable_variants() == ["one-shot", "one-shot-128", "two-shot"]     # a non-stepping parent
    assert make(mode="drop", p=0.1)._usable_variants() == ["one-shot", "one-shot-128"]                # ranks may leave
    assert make(world=9)._usable_variants()[:2] == ["one-shot", "two-shot"]
    assert make(choice="two-shot")._exchange_variant() == native.variant_id("two-shot")
    assert make(choice="two-shot", mode="drop", p=0.1)._exchange_variant() == native.variant_id("one-shot")
    assert make(variant="one-shot-128", ll128=False)._exchange_variant() == native.variant_id("one-shot")
    assert make(variant="one-shot-128")._exchange_variant() == 3
    with pytest.raises(RuntimeError, match="multicast is not available"):
        make(variant="nvls", nvls=False)._exchange_variant()

    # the probe: pairs 0-1, 2-3; the odd rank out has no partner and passes; one torn line anywhere fails everybody
    class Comm:
        def __init__(self, verdicts):
            self.verdicts = verdicts

        def barrier(self):
            pass

        def allgather(self, ok):
            return [ok] + self.verdicts

    def probe(rank, world, result, others):
        t = Trainer.__new__(Trainer)
        t.world, t.rank, t.comm = world, rank, Comm(others)
        calls = []

        def ll128_probe(peer, iters, n_ctas):
            calls.append(peer)
            if isinstance(result, Exception):
                raise result
            return result
        t.engine = SimpleNamespace(ll128_probe=ll128_probe)
        return t._probe_ll128(), calls

    assert probe(2, 4, (0, 3.1, False), [True] * 3) == (True, [3])
    assert probe(0, 4, (5, 3.1, False), [True] * 3) == (False, [1])            # torn lines
    assert probe(0, 4, (0, 3.1, True), [True] * 3) == (False, [1])             # the pair did not finish
    assert probe(0, 4, RuntimeError("launch failed"), [True] * 3) == (False, [1])
    assert probe(2, 3, (0, 3.1, False), [True] * 2) == (True, [])              # rank 2 of 3 has no partner
    assert probe(0, 2, (0, 3.1, False), [False]) == (False, [1])               # the OTHER pair saw a torn line
    assert probe(0, 16, (0, 3.1, False), [True] * 15) == (False, [])           # the variant parks at most 7 peers' lines


def test_mpiexec_front_end_accepts_the_usual_open_mpi_and_mpich_options():
    """Command lines written for a real mpiexec keep working: placement options are accepted and ignored, `-x` and
    `-genv` reach the processes' environment, every process gets its rank."""
    code = "import os; print('rank', os.environ['RANK'], os.environ['WORLD_SIZE'], os.environ['FOO'], os.environ['BAZ'], os.environ['HOME_COPY'])"
    env = dict(os.environ, HOME_COPY="from-parent")
    out = subprocess.run([str(ROOT / "bin" / "mpiexec"), "--oversubscribe", "--allow-run-as-root", "--bind-to", "none",
                          "--map-by", "slot", "--mca", "btl", "self,vader", "-x", "FOO=bar", "-x", "HOME_COPY",
                          "-genv", "BAZ", "7", "-np", "3", sys.executable, "-c", code],
                         capture_output=True, text=True, cwd=ROOT, timeout=120, env=env)
    assert out.returncode == 0, out.stderr[-2000:]
    assert sorted(out.stdout.split("\n")[:3]) == [f"rank {r} 3 bar 7 from-parent" for r in range(3)]
    bad = subprocess.run([str(ROOT / "bin" / "mpiexec"), "-n", "2"], capture_output=True, text=True, cwd=ROOT, timeout=60)
    assert bad.returncode != 0 and "no command given" in bad.stderr
