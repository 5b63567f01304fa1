"""The tensor-core instantiation of the fused step kernel (conv1 / conv2 forward and the conv2 data gradient as
tcgen05 implicit GEMMs inside fwd_bwd_cluster_kernel, TF32 operands, fp32 accumulation) against the fp32 FFMA
instantiation of the same kernel and against PyTorch fp32.

What TF32 costs: operands keep 10 mantissa bits, so activations differ by ~1e-3 relative, and a max-pool argmax can
flip at a near-tie, which reroutes one pixel's gradient. The bounds are therefore on the direction and the size of
the whole gradient (and on a 100-step training run), not on its largest element.
reference: models/model.py:20-27 forward, autograd of data_parallelism_train.py:195-198.
"""
import copy

import pytest
import torch
import torch.nn.functional as F

from distributed_neural_network_b200 import Network
from distributed_neural_network_b200.ops import flat as L
from distributed_neural_network_b200.ops import native

pytestmark = pytest.mark.gpu


def _engine(model, B, precision, lr=0.0, mu=0.0, cs=0):
    eng = native.Engine(model.flat_params, torch.zeros(L.FLAT_NUMEL, device="cuda"), None, B, lr, mu, precision=precision)
    if cs:
        eng.set_cluster_size(cs)
    return eng


def _cos_err(a, b):
    a, b = a.double().flatten(), b.double().flatten()
    return (torch.dot(a, b) / (a.norm() * b.norm())).item(), ((a - b).norm() / b.norm()).item()


@pytest.mark.parametrize("cs", [2, 3, 4, 5, 6, 7, 8])
@pytest.mark.parametrize("B", [1, 5, 16, 40])
def test_tc_gradient_close_to_fp32_for_every_cluster_size(cs, B):
    torch.manual_seed(100 * cs + B)
    model = Network().cuda()
    x = torch.randn(B, 3, 32, 32, device="cuda")
    y = torch.randint(0, 10, (B,), device="cuda")
    e32, etc = _engine(model, B, "fp32", cs=cs), _engine(model, B, "tf32", cs=cs)
    g32 = torch.zeros(L.FLAT_NUMEL, device="cuda")
    gtc = torch.full((L.FLAT_NUMEL,), float("nan"), device="cuda")
    l32 = e32.grad_only(x, y, g32)
    ltc = etc.grad_only(x, y, gtc)
    assert torch.isfinite(gtc).all()
    assert abs(l32 - ltc) < 3e-3 * max(1.0, abs(l32)), (l32, ltc)
    cos, err = _cos_err(gtc, g32)
    assert cos > 0.9995 and err < 3e-2, (cs, B, cos, err)
    if B >= 16:   # every tensor on its own (a wrong layout in one small tensor would drown in the norm of fc1)
        vt, v3 = L.views(gtc), L.views(g32)
        for name in L.PARAM_SHAPES:
            c, e = _cos_err(vt[name], v3[name])
            # (conv1.weight sits at the end of the backward chain and collects every rounding and every flipped argmax)
            assert c > (0.99 if name.startswith("conv1") else 0.998), (name, c, e)
    assert gtc[~L.real_mask("cuda")].abs().sum().item() == 0.0
    e32.close(); etc.close()


def test_tc_conv1_is_exact_when_its_operands_are_tf32_representable():
    """If the image and conv1.weight survive TF32 rounding unchanged, the tensor core adds exactly the products the
    FFMA path adds (in another order), so conv1 no longer contributes any error: what is left is the rounding of p1
    in front of conv2 and of the gradient planes in front of the data gradient. The bound tightens accordingly."""
    def tf32(t):  # keep the 10 explicit mantissa bits TF32 has
        return (t.contiguous().view(torch.int32) & ~0x1fff).view(torch.float32)

    torch.manual_seed(7)
    model = Network()
    with torch.no_grad():
        model.conv1.weight.copy_(tf32(model.conv1.weight.clone()))
    model = model.cuda()
    B = 16
    x = tf32(torch.randn(B, 3, 32, 32)).cuda()
    y = torch.randint(0, 10, (B,), device="cuda")
    e32, etc = _engine(model, B, "fp32"), _engine(model, B, "tf32")
    g32, gtc = torch.zeros(L.FLAT_NUMEL, device="cuda"), torch.zeros(L.FLAT_NUMEL, device="cuda")
    l32, ltc = e32.grad_only(x, y, g32), etc.grad_only(x, y, gtc)
    cos, err = _cos_err(gtc, g32)
    assert abs(l32 - ltc) < 1.5e-3 and cos > 0.9997 and err < 2.5e-2, (l32, ltc, cos, err)
    e32.close(); etc.close()


def test_tc_training_run_tracks_fp32_and_pytorch():
    """100 optimizer steps from the same weights on the same batches: TF32 conv stages vs fp32 FFMA vs PyTorch."""
    torch.manual_seed(3)
    base = Network()
    B, steps, lr, mu = 16, 100, 0.01, 0.9
    xs = torch.randn(steps, B, 3, 32, 32)
    ys = torch.randint(0, 10, (steps, B))
    ref = copy.deepcopy(base)
    opt = torch.optim.SGD(ref.parameters(), lr=lr, momentum=mu)
    for t in range(steps):
        ref.zero_grad()
        F.cross_entropy(ref(xs[t]), ys[t]).backward()
        opt.step()
    out = {}
    for prec in ("fp32", "tf32"):
        m = copy.deepcopy(base).cuda()
        eng = _engine(m, B, prec, lr=lr, mu=mu)
        eng.train_device(xs.cuda(), ys.cuda())
        torch.cuda.synchronize()
        ctl = eng.read_ctl()
        assert ctl.step == steps
        out[prec] = (m.flat_params.detach().cpu(), ctl.loss_sum / steps)
        eng.close()
    delta = (ref.flat_params - base.flat_params).norm().item()          # how far training moved the weights
    d32 = (out["fp32"][0] - ref.flat_params).norm().item()
    dtc = (out["tf32"][0] - ref.flat_params).norm().item()
    assert d32 < 1e-4 * max(delta, 1e-6) + 1e-5, (d32, delta)
    assert dtc < 2e-2 * delta, (dtc, delta)                             # TF32: within 2% of the distance travelled
    rel_w = ((out["tf32"][0] - out["fp32"][0]).norm() / out["fp32"][0].norm()).item()
    assert rel_w < 1e-3, rel_w                                          # weights after 100 steps within 1e-3 relative
    assert abs(out["tf32"][1] - out["fp32"][1]) < 5e-3


def test_tc_multi_image_loop_and_chained_steps():
    """More images than resident clusters (the kernel loops, re-zeroing its gradient planes) and steps chained on the
    update counter: the tensor-memory accumulators and operand buffers must be reusable image after image."""
    torch.manual_seed(11)
    model = Network().cuda()
    B = 64
    x = torch.randn(B, 3, 32, 32, device="cuda")
    y = torch.randint(0, 10, (B,), device="cuda")
    for cs in (8, 6, 2):
        e32, etc = _engine(model, B, "fp32", cs=cs), _engine(model, B, "tf32", cs=cs)
        g32, gtc = torch.zeros(L.FLAT_NUMEL, device="cuda"), torch.zeros(L.FLAT_NUMEL, device="cuda")
        e32.grad_only(x, y, g32)
        etc.grad_only(x, y, gtc)
        etc.grad_only(x, y, gtc)          # a second launch reuses nothing from the first
        cos, err = _cos_err(gtc, g32)
        assert cos > 0.9995 and err < 3e-2, (cs, cos, err)
        e32.close(); etc.close()
