"""The fused step engine vs the CPU oracle (the reference's own code path: Network + CE + SGD)."""
import copy

import pytest
import torch
import torch.nn.functional as F

from distributed_neural_network_b200 import Network
from distributed_neural_network_b200.ops import flat as L
from distributed_neural_network_b200.ops import native

pytestmark = pytest.mark.gpu


def _rel(a, b):
    a, b = a.detach().cpu().double(), b.detach().cpu().double()
    return ((a - b).abs().max() / b.abs().max().clamp_min(1e-9)).item()


def _oracle_grads(model_cpu, x, y):
    model_cpu.zero_grad()
    loss = F.cross_entropy(model_cpu(x), y)
    loss.backward()
    return loss.item(), model_cpu.flat_grads.clone()


@pytest.mark.parametrize("cs", [0, 1, 2, 3, 4, 5, 6, 7, 8])
@pytest.mark.parametrize("B", [1, 4, 16, 64])
def test_grad_only_matches_autograd(B, cs):
    """Every cluster size of the forward/backward kernel (cs CTAs per image, DSMEM all-gathers
    between layers) must give the autograd gradient; B=64 with cs=8 also exercises the
    several-images-per-cluster loop."""
    torch.manual_seed(10 + B)
    ref = Network()
    gpu = copy.deepcopy(ref).cuda()
    x, y = torch.randn(B, 3, 32, 32), torch.randint(0, 10, (B,))
    loss_ref, g_ref = _oracle_grads(ref, x, y)
    mom = torch.zeros(L.FLAT_NUMEL, device="cuda")
    eng = native.Engine(gpu.flat_params, mom, None, max_batch=64, lr=0.001, mu=0.9)
    eng.set_cluster_size(cs)
    g = torch.zeros(L.FLAT_NUMEL, device="cuda")
    p_before = gpu.flat_params.clone()
    loss = eng.grad_only(x.cuda(), y.cuda(), g)
    assert abs(loss - loss_ref) < 1e-5 * max(1.0, abs(loss_ref))
    views_ref, views = L.views(g_ref), L.views(g.cpu())
    for name in L.PARAM_SHAPES:
        assert _rel(views[name], views_ref[name]) < 5e-5, name
    assert torch.equal(gpu.flat_params, p_before), "grad_only must not touch the parameters"
    assert torch.count_nonzero(mom) == 0
    eng.close()


def test_steps_match_torch_sgd():
    torch.manual_seed(42)
    ref = Network()
    gpu = copy.deepcopy(ref).cuda()
    opt = torch.optim.SGD(ref.parameters(), lr=0.01, momentum=0.9)
    mom = torch.zeros(L.FLAT_NUMEL, device="cuda")
    eng = native.Engine(gpu.flat_params, mom, None, max_batch=16, lr=0.01, mu=0.9)
    losses = []
    for step in range(6):
        x, y = torch.randn(16, 3, 32, 32), torch.randint(0, 10, (16,))
        ref.zero_grad()
        loss = F.cross_entropy(ref(x), y)
        loss.backward()
        opt.step()
        losses.append(loss.item())
        eng.step(x.cuda(), y.cuda())
    ctl = eng.read_ctl()
    assert ctl.step == 6 and ctl.loss_steps == 6 and ctl.status == 0
    assert abs(ctl.loss_sum - sum(losses)) < 1e-4 * sum(losses)
    for (name, p_ref), (_, p_gpu) in zip(ref.named_flat_params(), gpu.named_flat_params()):
        assert _rel(p_gpu, p_ref) < 2e-5, name
    # the state_dict round-trips into the oracle model
    ref2 = Network()
    ref2.load_state_dict({k: v.cpu() for k, v in gpu.state_dict().items()})
    assert eng.launches == 12
    eng.close()


def test_train_host_pipeline_equals_step_loop():
    torch.manual_seed(7)
    base = Network()
    a, b = copy.deepcopy(base).cuda(), copy.deepcopy(base).cuda()
    steps, B = 9, 16
    hx = torch.randn(steps, B, 3, 32, 32).pin_memory()
    hy = torch.randint(0, 10, (steps, B)).pin_memory()
    hl = torch.zeros(steps).pin_memory()
    ea = native.Engine(a.flat_params, torch.zeros(L.FLAT_NUMEL, device="cuda"), None, 16, 0.01, 0.9)
    eb = native.Engine(b.flat_params, torch.zeros(L.FLAT_NUMEL, device="cuda"), None, 16, 0.01, 0.9)
    ea.train_host(hx, hy, hl)
    torch.cuda.synchronize()
    for t in range(steps):
        eb.step(hx[t].cuda(), hy[t].cuda())
    torch.cuda.synchronize()
    assert torch.equal(a.flat_params, b.flat_params)
    ctl = eb.read_ctl()
    assert abs(hl.sum().item() - ctl.loss_sum) < 1e-4
    assert (hl > 0).all()
    ea.close(); eb.close()


def test_chained_steps_equal_single_steps_and_respect_foreign_writes():
    """Inside one train_device call the forward/backward kernel of step t+1 waits on the update counter instead
    of on the completion of step t's update grid; the first step of every call must still honour whatever the
    stream did in between (here: a PyTorch kernel that rescales the parameters)."""
    torch.manual_seed(11)
    base = Network()
    a, b = copy.deepcopy(base).cuda(), copy.deepcopy(base).cuda()
    steps, B = 40, 16
    x = torch.randn(steps, B, 3, 32, 32, device="cuda")
    y = torch.randint(0, 10, (steps, B), device="cuda")
    ea = native.Engine(a.flat_params, torch.zeros(L.FLAT_NUMEL, device="cuda"), None, 16, 0.01, 0.9)
    eb = native.Engine(b.flat_params, torch.zeros(L.FLAT_NUMEL, device="cuda"), None, 16, 0.01, 0.9)
    for rnd in range(3):
        ea.train_device(x, y)                       # 40 chained steps
        a.flat_params.mul_(0.97)                    # foreign kernel on the same stream, no synchronisation
        for t in range(steps):
            eb.step(x[t], y[t])                     # 40 stand-alone steps (grid-completion dependency)
        b.flat_params.mul_(0.97)
    torch.cuda.synchronize()
    assert torch.equal(a.flat_params, b.flat_params)
    ca, cb = ea.read_ctl(), eb.read_ctl()
    assert ca.step == cb.step == 3 * steps
    assert ca.updates_done == cb.updates_done == 3 * steps
    assert ca.loss_sum == cb.loss_sum
    ea.close(); eb.close()


@pytest.mark.parametrize("N,bs", [(16, 16), (1000, 16), (37, 8)])
def test_eval_matches_reference_semantics(N, bs):
    torch.manual_seed(N)
    ref = Network()
    gpu = copy.deepcopy(ref).cuda()
    x, y = torch.randn(N, 3, 32, 32), torch.randint(0, 10, (N,))
    # reference: mean of per-batch mean losses, and correct count (data_parallelism_train.py:157-183)
    losses, correct = [], 0
    with torch.no_grad():
        for i in range(0, N, bs):
            out = ref(x[i:i + bs])
            losses.append(F.cross_entropy(out, y[i:i + bs]).item())
            correct += (out.argmax(1) == y[i:i + bs]).sum().item()
    eng = native.Engine(gpu.flat_params, torch.zeros(L.FLAT_NUMEL, device="cuda"), None, 16, 0.01, 0.9)
    logits = torch.empty(N, 10, device="cuda")
    loss, hits = eng.evaluate(x.cuda(), y.cuda(), bs, logits)
    assert abs(loss - sum(losses) / len(losses)) < 1e-4
    assert hits == correct
    with torch.no_grad():
        assert _rel(logits, ref(x)) < 2e-5
    eng.close()


def test_network_autograd_path_on_gpu():
    torch.manual_seed(5)
    ref = Network()
    gpu = copy.deepcopy(ref).cuda()
    x, y = torch.randn(16, 3, 32, 32), torch.randint(0, 10, (16,))
    _, g_ref = _oracle_grads(ref, x, y)
    from distributed_neural_network_b200.ops import functional as Fn
    Fn.set_conv_precision("fp32")
    try:
        gpu.zero_grad()
        loss = Fn.cross_entropy(gpu(x.cuda()), y.cuda())
        loss.backward()
    finally:
        Fn.set_conv_precision("tf32")
    assert _rel(gpu.flat_grads, g_ref) < 5e-5
    gpu.zero_grad()                                  # default: TF32 tensor-core conv forward
    Fn.cross_entropy(gpu(x.cuda()), y.cuda()).backward()
    assert _rel(gpu.flat_grads, g_ref) < 2e-2


def test_train_host_long_runs_reuse_the_ring_without_events():
    """The host-fed path of the clustered kernel uses no stream event: the copy stream writes a sequence number behind
    each group's inputs, the kernel that reads the group waits for it, and a ring region is reused once the update
    kernel of its last reader has reported in through a pinned progress word. 150 steps cross the two regions many
    times; several calls in a row, a call with another batch size (other packing of the region), a batch that runs
    the one-CTA kernel (event variant) in between, and the host buffer being overwritten right after each call's
    synchronize: the weights must be exactly those of the device-resident path."""
    torch.manual_seed(17)
    base = Network()
    a, b = copy.deepcopy(base).cuda(), copy.deepcopy(base).cuda()
    ea = native.Engine(a.flat_params, torch.zeros(L.FLAT_NUMEL, device="cuda"), None, 256, 0.01, 0.9)
    eb = native.Engine(b.flat_params, torch.zeros(L.FLAT_NUMEL, device="cuda"), None, 256, 0.01, 0.9)
    for steps, B in ((150, 16), (3, 16), (37, 8), (2, 256), (41, 16)):
        hx = torch.randn(steps, B, 3, 32, 32).pin_memory()
        hy = torch.randint(0, 10, (steps, B)).pin_memory()
        hl = torch.full((steps,), float("nan")).pin_memory()
        dx, dy = hx.cuda(), hy.cuda()
        ea.train_host(hx, hy, hl)
        eb.train_device(dx, dy)
        torch.cuda.synchronize()
        hx.fill_(123.0)                       # the call is over: its host buffers belong to the caller again
        assert torch.isfinite(hl).all() and (hl > 0).all()
        assert torch.equal(a.flat_params, b.flat_params), (steps, B)
    assert ea.read_ctl().step == eb.read_ctl().step == 150 + 3 + 37 + 2 + 41
    ea.close(); eb.close()
