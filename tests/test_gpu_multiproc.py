"""Real multi-process, multi-GPU path: CUDA-IPC peer group + fused exchange, one process per GPU.
Needs at least 2 GPUs (`gpurun --gpus 2`); ranks sharing ONE GPU must never spin on each other."""
import json
import os
import sys
from pathlib import Path

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = Path(__file__).resolve().parent.parent

needs2 = pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs >= 2 GPUs")


def _launch(n, out_dir, mode, **kw):
    from distributed_neural_network_b200.parallel import launch
    kw["device"] = "cuda"
    rc = launch.spawn(n, [sys.executable, str(ROOT / "tests" / "_mp_worker.py"), str(out_dir), mode, json.dumps(kw)])
    assert rc == 0
    return [torch.load(f"{out_dir}/rank{r}.pt", weights_only=False) for r in range(n)]


@needs2
def test_ipc_step_allreduce_replicas_identical_and_match_cpu_run(tmp_path):
    n = min(torch.cuda.device_count(), 8)
    kw = dict(train_size=64 * n, test_size=32, epochs=2, batch_size=8, lr=0.05, seed=11)
    gpu = _launch(n, tmp_path / "gpu", "step-allreduce", **kw)
    for r in range(1, n):
        assert torch.equal(gpu[0]["flat"], gpu[r]["flat"]), f"rank {r} is not bit-identical to rank 0"
    # the same run on CPU over gloo (same seeds => same shuffles and initial weights)
    from distributed_neural_network_b200.parallel import launch
    (tmp_path / "cpu").mkdir()
    kw_cpu = dict(kw, device="cpu")
    rc = launch.spawn(n, [sys.executable, str(ROOT / "tests" / "_mp_worker.py"), str(tmp_path / "cpu"),
                          "step-allreduce", json.dumps(kw_cpu)], env_extra={"OMP_NUM_THREADS": "1"}, gpus=0)
    assert rc == 0
    cpu = torch.load(tmp_path / "cpu" / "rank0.pt", weights_only=False)
    rel = ((gpu[0]["flat"] - cpu["flat"]).abs().max() / cpu["flat"].abs().max()).item()
    assert rel < 1e-4, rel


@needs2
def test_ipc_epoch_average_reference_topology(tmp_path):
    n = min(torch.cuda.device_count(), 4)
    res = _launch(n, tmp_path, "epoch-average", reference_topology=True, train_size=96, test_size=32, epochs=2,
                  batch_size=8, lr=0.05, seed=21, nb_proc=n)
    for r in range(1, n):
        assert torch.equal(res[0]["flat"], res[r]["flat"])
    assert res[0]["timers"]["training_time"] == 0.0     # the parent never trains


@needs2
def test_ipc_drop_mode_survivors_continue(tmp_path):
    n = min(torch.cuda.device_count(), 8)
    res = _launch(n, tmp_path, "step-allreduce", train_size=128 * n, test_size=32, epochs=3, batch_size=8, lr=0.02,
                  seed=9, failure_probability=0.5, failure_duration=0.05, failure_mode="drop", peer_timeout_s=20.0)
    assert res[0]["history"][-1]["membership_epoch"] > 0
    for r in range(1, n):
        assert torch.equal(res[0]["flat"], res[r]["flat"])


@needs2
def test_nvls_in_switch_allreduce_matches_peer_memory_variant(tmp_path):
    """multimem.ld_reduce / multimem.st over an NVLink multicast mapping vs the one-shot push variant."""
    from distributed_neural_network_b200.parallel import launch
    n = min(torch.cuda.device_count(), 8)
    rc = launch.spawn(n, [sys.executable, str(ROOT / "tests" / "_mp_nvls_worker.py"), str(tmp_path)])
    assert rc == 0
    res = [torch.load(f"{tmp_path}/rank{r}.pt", weights_only=False) for r in range(n)]
    if not all(r["nvls"] for r in res):
        pytest.skip("NVLink multicast is not exposed on this host: " + "; ".join(sorted({r["why_not"] for r in res})))
    assert all(r["status"] == 0 for r in res)
    for key in res[0]["outs"]:
        nelem = int(key.split("_")[0])
        want = torch.stack([r["outs"][key][0][:nelem].double() for r in res]).mean(0)
        for r in res:
            x, y_one, y_nvls = r["outs"][key]
            assert torch.equal(y_nvls, res[0]["outs"][key][2]), "NVLS result differs between ranks"
            assert (y_nvls[:nelem].double() - want).abs().max() < 1e-6
            assert (y_one[:nelem].double() - want).abs().max() < 1e-6


@needs2
def test_ipc_fused_step_with_nvls_exchange_matches_one_shot(tmp_path):
    """Same seeds, same data: the step with the in-switch (NVLS) exchange must train the same model as with one-shot."""
    n = min(torch.cuda.device_count(), 8)
    kw = dict(train_size=64 * n, test_size=32, epochs=2, batch_size=8, lr=0.05, seed=11)
    a = _launch(n, tmp_path / "one", "step-allreduce", allreduce_variant="one-shot", **kw)
    b = _launch(n, tmp_path / "nvls", "step-allreduce", allreduce_variant="nvls", **kw)
    for r in range(1, n):
        assert torch.equal(b[0]["flat"], b[r]["flat"]), "NVLS replicas differ between ranks"
    rel = ((a[0]["flat"] - b[0]["flat"]).abs().max() / a[0]["flat"].abs().max()).item()
    assert rel < 1e-5, rel     # summation order inside the switch differs from rank order: last-bit differences only


@needs2
def test_ll128_lines_are_never_torn(tmp_path):
    """The one-shot-128 exchange assumes that a warp-wide 128-byte store arrives in the peer's memory as one unit
    (a load that returns this step's flag word also returns this step's 31 values). Hammer exactly that over
    NVLink: 122 CTAs x 8 warps x 3000 rounds of ping-pong per pair of GPUs, every arriving line checked."""
    from distributed_neural_network_b200.parallel import launch
    n = min(torch.cuda.device_count(), 8) & ~1
    rc = launch.spawn(n, [sys.executable, str(ROOT / "tests" / "_mp_probe_worker.py"), str(tmp_path), "3000", "122"])
    assert rc == 0
    res = [torch.load(f"{tmp_path}/rank{r}.pt", weights_only=False) for r in range(n)]
    for r in res:
        torn, us_round_trip, gave_up = r["probe"]
        assert not gave_up, f"rank {r['rank']}: a line never arrived"
        assert torn == 0, f"rank {r['rank']}: {torn} words arrived under the wrong flag"
        assert 0.5 < us_round_trip < 200.0


@needs2
def test_ipc_one_shot_128_trains_the_same_bits_as_one_shot(tmp_path):
    """Both one-shot forms add the contributions in rank order, so they must produce the SAME model bit for bit."""
    n = min(torch.cuda.device_count(), 8)
    kw = dict(train_size=64 * n, test_size=32, epochs=2, batch_size=8, lr=0.05, seed=11)
    a = _launch(n, tmp_path / "ll16", "step-allreduce", allreduce_variant="one-shot", **kw)
    b = _launch(n, tmp_path / "ll128", "step-allreduce", allreduce_variant="one-shot-128", **kw)
    for r in range(n):
        assert torch.equal(a[0]["flat"], b[r]["flat"])
        # ... and the second run really used the 128-byte lines: the start-up probe found them whole on these links
        assert b[r]["ll128_ok"] and b[r]["variant"] == 3 and a[r]["variant"] == 0


@needs2
def test_exchange_variant_is_chosen_by_measurement(tmp_path):
    n = min(torch.cuda.device_count(), 8)
    res = _launch(n, tmp_path, "step-allreduce", train_size=64 * n, test_size=32, epochs=1, batch_size=16, lr=0.05,
                  seed=5, calibrate_steps=100)
    table = res[0]["exchange_table"]
    assert table["world"] == n and set(table["us_per_step_max_over_ranks"]) >= {"one-shot", "one-shot-128", "two-shot"}
    assert table["chosen"] == min(table["us_per_step_max_over_ranks"], key=table["us_per_step_max_over_ranks"].get)
    assert all(r["exchange_table"]["chosen"] == table["chosen"] for r in res)
    for r in range(1, n):
        assert torch.equal(res[0]["flat"], res[r]["flat"])
    assert res[0]["history"][-1].get("allreduce_us", 0) > 0


@needs2
def test_second_device_in_one_process():
    """Function attributes and the no-allocation scratch are per DEVICE: everything that worked on cuda:0 must
    work on cuda:1 afterwards in the same process (launches with > 48 KB of dynamic shared memory included)."""
    import ctypes as C

    from distributed_neural_network_b200 import Network
    from distributed_neural_network_b200.ops import flat as L
    from distributed_neural_network_b200.ops import native

    lib = native.require()
    fn = lib.dnnb200_batch_grad_noalloc
    fn.restype = C.c_int
    fn.argtypes = [C.c_void_p] * 3 + [C.c_int] + [C.c_void_p] * 3
    torch.manual_seed(5)
    base = Network()
    x, y = torch.randn(16, 3, 32, 32), torch.randint(0, 10, (16,))
    got = []
    for d in (0, 1):
        dev = torch.device("cuda", d)
        with torch.cuda.device(dev):
            import copy
            m = copy.deepcopy(base).to(dev)
            eng = native.Engine(m.flat_params, torch.zeros(L.FLAT_NUMEL, device=dev), None, 64, 0.01, 0.9)
            xd, yd = x.to(dev), y.to(dev)
            g_eng = torch.zeros(L.FLAT_NUMEL, device=dev)
            eng.grad_only(xd, yd, g_eng)                           # clustered kernel (> 48 KB smem)
            g_na = torch.full((L.FLAT_NUMEL,), float("nan"), device=dev)
            loss = torch.zeros(1, device=dev)
            rc = fn(m.flat_params.data_ptr(), xd.data_ptr(), yd.data_ptr(), 16, g_na.data_ptr(), loss.data_ptr(),
                    torch.cuda.current_stream().cuda_stream)      # one-CTA kernel + the per-device scratch address
            assert rc == 0, lib.dnnb200_last_error().decode()
            ev_loss, hits = eng.evaluate(xd, yd, 8)                # eval kernel (> 48 KB smem)
            eng.step(xd, yd)
            torch.cuda.synchronize()
            rel = ((g_na - g_eng).abs().max() / g_eng.abs().max()).item()
            assert rel < 5e-5, (d, rel)
            got.append((g_eng.cpu(), ev_loss, hits, m.flat_params.detach().cpu()))
            eng.close()
    assert torch.allclose(got[0][0], got[1][0], rtol=0, atol=0) and got[0][1:3] == got[1][1:3]
    assert torch.equal(got[0][3], got[1][3])


@needs2
def test_ipc_autograd_step_path_with_bucketed_overlap(tmp_path):
    """--step-path autograd on GPUs: layer kernels through autograd, the gradient averaged bucket by bucket over peer
    memory on a side stream while backward runs. Replicas must stay bit-identical, with and without the overlap, and
    the two runs must agree (the exchange is the same arithmetic in the same order either way)."""
    n = min(torch.cuda.device_count(), 8)
    kw = dict(train_size=256 * n, test_size=32, epochs=2, batch_size=64, lr=0.05, seed=13, step_path="autograd")
    a = _launch(n, tmp_path / "overlap", "step-allreduce", overlap_exchange=True, **kw)
    b = _launch(n, tmp_path / "after", "step-allreduce", overlap_exchange=False, **kw)
    for r in range(1, n):
        assert torch.equal(a[0]["flat"], a[r]["flat"]) and torch.equal(b[0]["flat"], b[r]["flat"])
    # (large-batch layer kernels add their batch chunks with atomics, so two RUNS agree to round-off, not to the bit)
    rel = ((a[0]["flat"] - b[0]["flat"]).norm() / b[0]["flat"].norm()).item()
    assert rel < 1e-4, rel
    assert a[0]["history"][-1]["train_loss"] < a[0]["history"][0]["train_loss"] + 0.05


@needs2
def test_native_nvls_group_two_devices_in_one_process():
    """csrc/nvls_group.cu directly: two threads stand in for two ranks (devices 0 and 1), rank 0 creates the multicast
    object and hands it over as a file descriptor, both bind and map; then the in-switch all-reduce through that mapping
    must equal the one-shot push variant and the mean computed on the host."""
    import threading
    import time

    from distributed_neural_network_b200.ops import flat as L
    from distributed_neural_network_b200.ops import native

    world = 2
    if not all(native.nvls_supported(d) for d in range(world)):
        pytest.skip("NVLink multicast is not exposed on this host")
    groups, engines = [], []
    for d in range(world):
        with torch.cuda.device(d):
            groups.append(native.Group(d, world))
    native.Group.connect_local(groups)
    for d in range(world):
        with torch.cuda.device(d):
            p = torch.zeros(L.FLAT_NUMEL, device=f"cuda:{d}")
            e = native.Engine(p, torch.zeros_like(p), None, 16, 0.01, 0.9, precision="fp32")
            e.attach_group(groups[d])
            e.set_membership((1 << world) - 1, timeout_ns=5_000_000_000)
            torch.cuda.synchronize(d)
            engines.append(e)
    name = f"test-{os.getpid()}-{time.time_ns():x}"
    gen = torch.Generator().manual_seed(77)
    xs = [torch.randn(L.FLAT_NUMEL, generator=gen) for _ in range(world)]
    bar = threading.Barrier(world)
    out, errs, made = {}, [], {}

    def run(r):
        try:
            with torch.cuda.device(r):
                g = native.NvlsGroup(r, world, r, native.NVLS_FLOATS * 4, name, 20.0)
                made[r] = g
                bar.wait(30)                       # every device has been added
                g.map()
                bar.wait(30)                       # every rank has bound its memory
                assert g.nbytes >= native.NVLS_FLOATS * 4 and g.local_ptr and g.multicast_ptr
                engines[r].attach_nvls(g, g.multicast_ptr)
                x = xs[r].cuda(r)
                y_one, y_nvls = torch.zeros_like(x), torch.zeros_like(x)
                for _ in range(3):                 # both parities, flag reuse
                    engines[r].allreduce(x, y_one, mode=0, two_shot="one-shot")
                    engines[r].allreduce(x, y_nvls, mode=0, two_shot="nvls")
                torch.cuda.synchronize(r)
                out[r] = (y_one.cpu(), y_nvls.cpu(), int(engines[r].read_ctl().status))
                bar.wait(30)
        except Exception as e:                     # noqa: BLE001 -- reported below, the other thread must not wait
            errs.append((r, repr(e)))
            bar.abort()

    threads = [threading.Thread(target=run, args=(r,)) for r in range(world)]
    for t in threads:
        t.start()
    for t in threads:
        t.join(120)
    for e in engines:
        e.close()
    for g in made.values():
        g.close()
    for g in groups:
        g.close()
    assert not errs, errs
    want = torch.stack([x.double() for x in xs]).mean(0)
    for r in range(world):
        y_one, y_nvls, status = out[r]
        assert status == 0
        assert (y_one.double() - want).abs().max() < 1e-6 and (y_nvls.double() - want).abs().max() < 1e-6
    assert torch.equal(out[0][1], out[1][1]), "NVLS result differs between ranks"
