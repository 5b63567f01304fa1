"""Layer kernels vs a plain PyTorch fp32 reference of the same op (unit level)."""
import pytest
import torch
import torch.nn.functional as F

from distributed_neural_network_b200.ops import functional as Fn
from distributed_neural_network_b200.ops import native

pytestmark = pytest.mark.gpu


@pytest.fixture(autouse=True)
def _fp32_convs():
    """These tests compare at fp32 tolerances: use the FFMA conv forward (tests/test_gpu_tc.py covers TF32)."""
    Fn.set_conv_precision("fp32")
    yield
    Fn.set_conv_precision("tf32")


def _close(a, b, tol=2e-5):
    a, b = a.detach().cpu().double(), b.detach().cpu().double()
    denom = b.abs().max().clamp_min(1e-6)
    assert ((a - b).abs().max() / denom).item() < tol, ((a - b).abs().max().item(), denom.item())


@pytest.mark.parametrize("N,Ci,H,Co", [(16, 3, 32, 6), (16, 6, 14, 16), (3, 3, 32, 6), (65, 6, 14, 16), (520, 3, 32, 6),
                                       (1000, 6, 14, 16)])
def test_conv_relu_pool_fwd_bwd(N, Ci, H, Co):
    torch.manual_seed(0)
    x = torch.randn(N, Ci, H, H)
    w = torch.randn(Co, Ci, 5, 5) * 0.2
    b = torch.randn(Co) * 0.1
    xr, wr, br = (t.clone().requires_grad_(True) for t in (x, w, b))
    yr = F.max_pool2d(F.relu(F.conv2d(xr, wr, br)), 2, 2)
    gy = torch.randn_like(yr)
    yr.backward(gy)

    xg, wg, bg = (t.cuda().requires_grad_(True) for t in (x, w, b))
    yg = Fn.conv_relu_pool(xg, wg, bg)
    yg.backward(gy.cuda())
    tol = 2e-5 if N <= 100 else 1e-4     # fp32 sums over tens of thousands of terms, in a different order
    _close(yg, yr)
    _close(wg.grad, wr.grad, tol)
    _close(bg.grad, br.grad, tol)
    _close(xg.grad, xr.grad)


@pytest.mark.parametrize("N,K,M,relu", [(16, 400, 120, True), (16, 120, 84, True), (16, 84, 10, False), (5, 400, 120, True),
                                        (2000, 84, 10, False), (700, 120, 84, True)])
def test_linear_fwd_bwd(N, K, M, relu):
    torch.manual_seed(1)
    x, w, b = torch.randn(N, K), torch.randn(M, K) * 0.1, torch.randn(M) * 0.1
    xr, wr, br = (t.clone().requires_grad_(True) for t in (x, w, b))
    yr = F.linear(xr, wr, br)
    yr = F.relu(yr) if relu else yr
    gy = torch.randn_like(yr)
    yr.backward(gy)
    xg, wg, bg = (t.cuda().requires_grad_(True) for t in (x, w, b))
    yg = Fn.linear(xg, wg, bg, relu=relu)
    yg.backward(gy.cuda())
    tol = 2e-5 if N <= 100 else 1e-4
    _close(yg, yr)
    _close(xg.grad, xr.grad)
    _close(wg.grad, wr.grad, tol)
    _close(bg.grad, br.grad, tol)


@pytest.mark.parametrize("N", [1, 16, 300, 5000])
def test_softmax_xent(N):
    torch.manual_seed(2)
    z = torch.randn(N, 10) * 3
    t = torch.randint(0, 10, (N,))
    zr = z.clone().requires_grad_(True)
    lr_ = F.cross_entropy(zr, t)
    lr_.backward()
    zg = z.cuda().requires_grad_(True)
    lg = Fn.cross_entropy(zg, t.cuda())
    lg.backward()
    _close(lg, lr_, 1e-5)
    _close(zg.grad, zr.grad, 1e-5)


def test_sgd_momentum_matches_torch_optim():
    torch.manual_seed(3)
    n = 62144
    p0 = torch.randn(n)
    pr = p0.clone().requires_grad_(True)
    opt = torch.optim.SGD([pr], lr=0.01, momentum=0.9)
    pg, buf = p0.cuda(), torch.zeros(n, device="cuda")
    for step in range(4):
        g = torch.randn(n)
        pr.grad = g.clone()
        opt.step()
        Fn.sgd_momentum_(pg, g.cuda(), buf, 0.01, 0.9, first_step=(step == 0))
    _close(pg, pr, 1e-6)


def test_missing_library_is_loud(monkeypatch):
    monkeypatch.setattr(native, "_lib", None)
    monkeypatch.setattr(native, "_load_error", "simulated")
    with pytest.raises(RuntimeError, match="never routed to an eager fallback"):
        Fn.linear(torch.zeros(1, 4, device="cuda"), torch.zeros(2, 4, device="cuda"), torch.zeros(2, device="cuda"))
