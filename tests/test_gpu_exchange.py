"""Peer-memory exchange protocol with the ranks emulated on ONE device.

Kernels that wait on each other must not run as separate launches on one GPU, so the
W "ranks" here are the y-dimension of a single cooperative launch; the code path inside
the kernel (push, flags, parity, rank-ordered sum, leave markers) is the production one.
The real multi-process / multi-GPU path is exercised by tests/test_gpu_multiproc.py.
"""
import copy

import pytest
import torch
import torch.nn.functional as F

from distributed_neural_network_b200 import Network
from distributed_neural_network_b200.ops import flat as L
from distributed_neural_network_b200.ops import native

pytestmark = pytest.mark.gpu


def _make(world, lr=0.01, mu=0.9, B=16, seed=0):
    torch.manual_seed(seed)
    base = Network()
    models = [copy.deepcopy(base).cuda() for _ in range(world)]
    groups = [native.Group(r, world) for r in range(world)]
    native.Group.connect_local(groups)
    engines = []
    for r in range(world):
        e = native.Engine(models[r].flat_params, torch.zeros(L.FLAT_NUMEL, device="cuda"), None, B, lr, mu)
        e.attach_group(groups[r])
        e.set_membership((1 << world) - 1, timeout_ns=2_000_000_000)
        engines.append(e)
    return base, models, groups, engines


def _oracle_step(ref, opt, xs, ys):
    """Averaged gradient over the participating ranks == one big batch (equal local batch sizes)."""
    x, y = torch.cat(xs), torch.cat(ys)
    ref.zero_grad()
    F.cross_entropy(ref(x), y).backward()
    opt.step()


def _rel(a, b):
    a, b = a.detach().cpu().double(), b.detach().cpu().double()
    return ((a - b).abs().max() / b.abs().max().clamp_min(1e-9)).item()


VARIANTS = ["one-shot", "two-shot", "one-shot-128"]


@pytest.mark.parametrize("variant", VARIANTS)
@pytest.mark.parametrize("world", [2, 4, 8])
def test_vr_step_replicas_identical_and_match_oracle(world, variant):
    B = 16
    base, models, groups, engines = _make(world, B=B, seed=world)
    for e in engines:
        e.set_exchange_variant(variant)
    ref = copy.deepcopy(base)
    opt = torch.optim.SGD(ref.parameters(), lr=0.01, momentum=0.9)
    for step in range(4):
        xs = [torch.randn(B, 3, 32, 32) for _ in range(world)]
        ys = [torch.randint(0, 10, (B,)) for _ in range(world)]
        _oracle_step(ref, opt, xs, ys)
        native.vr_step(engines, [x.cuda() for x in xs], [y.cuda() for y in ys], B)
    torch.cuda.synchronize()
    for r in range(1, world):
        assert torch.equal(models[0].flat_params, models[r].flat_params), f"rank {r} drifted from rank 0"
    assert _rel(models[0].flat_params, ref.flat_params) < 3e-5
    for e in engines:
        c = e.read_ctl()
        assert c.step == 4 and c.status == 0
    for e in engines: e.close()
    for g in groups: g.close()


@pytest.mark.parametrize("variant", ["one-shot", "one-shot-128"])
def test_membership_epoch_leave_and_rejoin(variant):
    world, B = 4, 8
    base, models, groups, engines = _make(world, B=B, seed=99)
    for e in engines:
        e.set_exchange_variant(variant)
    ref = copy.deepcopy(base)
    opt = torch.optim.SGD(ref.parameters(), lr=0.01, momentum=0.9)

    def run(active):
        xs = [torch.randn(B, 3, 32, 32) if r in active else None for r in range(world)]
        ys = [torch.randint(0, 10, (B,)) if r in active else None for r in range(world)]
        _oracle_step(ref, opt, [x for x in xs if x is not None], [y for y in ys if y is not None])
        native.vr_step(engines, [x.cuda() if x is not None else None for x in xs],
                       [y.cuda() if y is not None else None for y in ys], B)

    run({0, 1, 2, 3})
    run({0, 1, 2, 3})
    engines[2].leave()            # rank 2 drops out mid-epoch
    torch.cuda.synchronize()
    run({0, 1, 3})                # survivors keep training, divisor 3
    run({0, 1, 3})
    torch.cuda.synchronize()
    for r in (1, 3):
        assert torch.equal(models[0].flat_params, models[r].flat_params)
    assert _rel(models[0].flat_params, ref.flat_params) < 3e-5
    assert not torch.equal(models[0].flat_params, models[2].flat_params)
    assert all(engines[r].read_ctl().status == 0 for r in (0, 1, 3))
    # rejoin at a boundary: copy state from a survivor, adopt its sequence number, clear the marker
    models[2].flat_params.copy_(models[0].flat_params)
    engines[2].momentum.copy_(engines[0].momentum)
    engines[2].rejoin(engines[0].read_ctl().step)
    torch.cuda.synchronize()
    run({0, 1, 2, 3})
    torch.cuda.synchronize()
    for r in (1, 2, 3):
        assert torch.equal(models[0].flat_params, models[r].flat_params)
    assert _rel(models[0].flat_params, ref.flat_params) < 3e-5
    for e in engines: e.close()
    for g in groups: g.close()


@pytest.mark.parametrize("two_shot", VARIANTS)
@pytest.mark.parametrize("world", [2, 8])
def test_vr_allreduce_mean(world, two_shot):
    base, models, groups, engines = _make(world, seed=5)
    torch.manual_seed(123)
    for n in (L.FLAT_NUMEL, 4096, 512 + 64):
        ins = [torch.randn(n, device="cuda") for _ in range(world)]
        outs = [torch.empty(n, device="cuda") for _ in range(world)]
        native.vr_allreduce(engines, ins, outs, n, mode=0, two_shot=two_shot)
        torch.cuda.synchronize()
        want = torch.stack([t.double() for t in ins]).mean(0)
        for r in range(world):
            assert torch.equal(outs[0], outs[r])
        assert (outs[0].double() - want).abs().max().item() < 1e-6
    assert all(e.read_ctl().status == 0 for e in engines)
    for e in engines: e.close()
    for g in groups: g.close()


def test_vr_allreduce_fused_sgd_equals_unfused():
    world = 4
    base, models, groups, engines = _make(world, lr=0.05, mu=0.9, seed=8)
    p_ref = base.flat_params.clone().cuda()
    buf_ref = torch.zeros_like(p_ref)
    for step in range(3):
        grads = [torch.randn(L.FLAT_NUMEL, device="cuda") for _ in range(world)]
        native.vr_allreduce(engines, grads, [None] * world, L.FLAT_NUMEL, mode=1)
        g = grads[0].clone()
        for t in grads[1:]:
            g = g + t                      # same rank order as the kernel
        g = g * (1.0 / world)
        buf_ref = 0.9 * buf_ref + g
        p_ref = p_ref - 0.05 * buf_ref
    torch.cuda.synchronize()
    for r in range(world):
        assert torch.equal(models[0].flat_params, models[r].flat_params)
    assert _rel(models[0].flat_params, p_ref) < 1e-6
    for e in engines: e.close()
    for g in groups: g.close()


@pytest.mark.parametrize("variant", VARIANTS)
def test_timeout_marks_status_and_applies_a_clean_step(variant):
    """A peer that neither steps nor announces that it left: the waiting rank must give up (not hang), say so in
    its status word, and apply a step that is a true mean over the ranks that DID arrive -- here itself alone.
    For two-shot that covers both halves: the slices rank 0 owns (missing contribution left out, sum rescaled)
    and the slices the silent rank owns (rank 0 keeps its own gradient)."""
    world, B = 2, 4
    base, models, groups, engines = _make(world, B=B, seed=3)
    for e in engines:
        e.set_membership(0b11, timeout_ns=50_000_000)  # 50 ms
        e.set_exchange_variant(variant)
    xc, yc = torch.randn(B, 3, 32, 32), torch.randint(0, 10, (B,))
    ref = copy.deepcopy(base)
    opt = torch.optim.SGD(ref.parameters(), lr=0.01, momentum=0.9)
    _oracle_step(ref, opt, [xc], [yc])
    native.vr_step(engines, [xc.cuda(), None], [yc.cuda(), None], B)
    torch.cuda.synchronize()
    c = engines[0].read_ctl()
    assert c.status & 1 and (c.status >> 8) & 0b10
    assert c.step == 1
    assert _rel(models[0].flat_params, ref.flat_params) < 3e-5
    for e in engines: e.close()
    for g in groups: g.close()


def test_vr_allreduce_short_vectors_with_128_byte_lines():
    """Vector lengths that end inside a line, inside a slice, and exactly on their borders."""
    world = 4
    base, models, groups, engines = _make(world, seed=6)
    torch.manual_seed(7)
    for n in (4, 28, 32, 60, 64, 124, 508, 512, 516, 1024 + 36, L.FLAT_NUMEL - 4):
        ins = [torch.randn(n, device="cuda") for _ in range(world)]
        outs = [torch.full((n,), float("nan"), device="cuda") for _ in range(world)]
        native.vr_allreduce(engines, ins, outs, n, mode=0, two_shot="one-shot-128")
        torch.cuda.synchronize()
        want = torch.stack([t.double() for t in ins]).mean(0)
        for r in range(world):
            assert torch.equal(outs[0], outs[r])
        assert (outs[0].double() - want).abs().max().item() < 1e-6, n
    for e in engines: e.close()
    for g in groups: g.close()
