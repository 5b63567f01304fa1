"""The allocation-free batch-gradient entry point, called straight through the C ABI.

Who needs it: callers that are not allowed to call cudaMalloc -- a process whose GPU memory is already
handed out in full to another allocator, or a latency-critical path where cudaMalloc's driver locks and
device-wide synchronisation are unacceptable. The entry point keeps its scratch in a fixed array inside
the library, so this test checks the three things such a caller relies on: the numbers equal those of the
ordinary (allocating) entry point, a second call is not disturbed by what the first one left in the shared
array, and free device memory does not move across calls. Oversized batches must be refused, not truncated.

This file sorts last on purpose: if it fails under `-x`, every other test has already run.
"""
import ctypes as C

import pytest
import torch

from distributed_neural_network_b200 import Network
from distributed_neural_network_b200.ops import flat as L
from distributed_neural_network_b200.ops import native

pytestmark = pytest.mark.gpu

MAX_BATCH = 64           # DNNB200_NOALLOC_MAX_BATCH
SCRATCH_BYTES = 967168   # DNNB200_NOALLOC_SCRATCH_BYTES


def _entry():
    lib = native.require()
    fn = lib.dnnb200_batch_grad_noalloc
    fn.restype = C.c_int
    fn.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
    return lib, fn


def _call(fn, params, x, y, grads, loss):
    stream = torch.cuda.current_stream().cuda_stream
    return fn(params.data_ptr(), x.data_ptr(), y.data_ptr(), x.shape[0], grads.data_ptr(), loss.data_ptr(), stream)


def _rel(a, b):
    return ((a.double() - b.double()).abs().max() / b.double().abs().max().clamp_min(1e-12)).item()


def test_no_alloc_entry_matches_the_allocating_one_twice_in_a_row():
    lib, fn = _entry()
    torch.manual_seed(2024)
    model = Network().cuda()
    eng = native.Engine(model.flat_params, torch.zeros(L.FLAT_NUMEL, device="cuda"), None, MAX_BATCH, 0.0, 0.0)
    # everything the calls touch is created up front: the entry point itself must not need memory
    cases = []
    for B in (16, 5, MAX_BATCH):   # the 2nd call is SMALLER than the 1st: stale records of images 5..15 stay in
        x = torch.randn(B, 3, 32, 32, device="cuda")          # the shared array and must not leak into the result
        y = torch.randint(0, 10, (B,), device="cuda")
        cases.append((x, y, torch.full((L.FLAT_NUMEL,), float("nan"), device="cuda"),
                      torch.zeros(L.FLAT_NUMEL, device="cuda"), torch.zeros(1, device="cuda")))
    torch.cuda.synchronize()
    for n, (x, y, g_new, g_old, loss_new) in enumerate(cases):
        free_before = torch.cuda.mem_get_info()[0]
        rc = _call(fn, model.flat_params, x, y, g_new, loss_new)
        assert rc == 0, lib.dnnb200_last_error().decode()
        torch.cuda.synchronize()
        if n > 0:   # the very first call on a device may load the module that holds the array
            assert torch.cuda.mem_get_info()[0] == free_before, "the call changed the amount of free device memory"
        loss_old = eng.grad_only(x, y, g_old)                # the existing entry point (engine-owned scratch)
        assert torch.isfinite(g_new).all()
        assert _rel(g_new, g_old) < 5e-5, (n, _rel(g_new, g_old))
        assert abs(loss_new.item() - loss_old) < 1e-5 * max(1.0, abs(loss_old))
        assert g_new[~L.real_mask("cuda")].abs().sum().item() == 0.0   # alignment gaps are written as zeros
    eng.close()


def test_no_alloc_entry_refuses_batches_beyond_its_scratch():
    lib, fn = _entry()
    model = Network().cuda()
    B = MAX_BATCH + 1
    x = torch.randn(B, 3, 32, 32, device="cuda")
    y = torch.randint(0, 10, (B,), device="cuda")
    g = torch.full((L.FLAT_NUMEL,), 7.0, device="cuda")
    loss = torch.zeros(1, device="cuda")
    rc = _call(fn, model.flat_params, x, y, g, loss)
    assert rc == 2                                            # DNNB200_ERR_ARG
    msg = lib.dnnb200_last_error().decode()
    assert "exceeds" in msg and str(MAX_BATCH) in msg and str(SCRATCH_BYTES) in msg
    torch.cuda.synchronize()
    assert (g == 7.0).all(), "a refused call must not launch anything"
    assert _call(fn, model.flat_params, x[:0], y[:0], g, loss) == 2   # empty batch is refused too


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs >= 2 GPUs")
def test_no_alloc_entry_on_a_second_device_of_the_same_process():
    """Every GPU that loads the library holds its own copy of the built-in array, and the entry point looks the
    array's address up per device: after cuda:0 has been served, cuda:1 must get its own address (and its own
    kernel attributes), not a pointer into cuda:0's memory."""
    lib, fn = _entry()
    torch.manual_seed(77)
    base = Network()
    x, y = torch.randn(24, 3, 32, 32), torch.randint(0, 10, (24,))
    results = []
    for d in (0, 1, 0):
        dev = torch.device("cuda", d)
        with torch.cuda.device(dev):
            params = base.flat_params.detach().to(dev)
            xd, yd = x.to(dev), y.to(dev)
            g = torch.full((L.FLAT_NUMEL,), float("nan"), device=dev)
            loss = torch.zeros(1, device=dev)
            rc = _call(fn, params, xd, yd, g, loss)
            assert rc == 0, lib.dnnb200_last_error().decode()
            torch.cuda.synchronize(dev)
            assert torch.isfinite(g).all()
            results.append((g.cpu(), loss.item()))
    for g, loss in results[1:]:
        assert torch.equal(g, results[0][0]) and loss == results[0][1]
