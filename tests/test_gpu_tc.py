"""Tensor-core (tcgen05 / TMEM) kernels vs plain PyTorch fp32 references of the same op.
TF32 operands (10-bit mantissa) with fp32 accumulation: tolerances are relative to the output scale."""
import pytest
import torch
import torch.nn.functional as F

from distributed_neural_network_b200.ops import functional as Fn
from distributed_neural_network_b200.ops import native

pytestmark = pytest.mark.gpu


def _rel(a, b):
    a, b = a.detach().cpu().double(), b.detach().cpu().double()
    return ((a - b).abs().max() / b.abs().max().clamp_min(1e-9)).item()


@pytest.mark.parametrize("K", [8, 16, 104, 200, 256])
def test_tcgen05_descriptor_probe_gemm(K):
    torch.manual_seed(K)
    A, B = torch.randn(128, K), torch.randn(16, K)
    D = torch.full((128, 16), float("nan"), device="cuda")
    native.tc_gemm_probe(A.cuda(), B.cuda(), D)
    torch.cuda.synchronize()
    want = A.double() @ B.double().t()
    assert torch.isfinite(D).all()
    assert _rel(D, want) < 3e-3, _rel(D, want)


@pytest.mark.parametrize("N,Ci,H,Co", [(1, 3, 32, 6), (16, 3, 32, 6), (300, 3, 32, 6), (1, 6, 14, 16), (16, 6, 14, 16),
                                       (700, 6, 14, 16)])
def test_tc_conv_relu_pool_forward(N, Ci, H, Co):
    torch.manual_seed(N + Ci)
    x = torch.randn(N, Ci, H, H)
    w = torch.randn(Co, Ci, 5, 5) * (Ci * 25) ** -0.5
    b = torch.randn(Co) * 0.1
    conv = F.relu(F.conv2d(x, w, b))
    want, want_idx = F.max_pool2d(conv, 2, 2, return_indices=True)
    Hp = (H - 4) // 2
    y = torch.full((N, Co, Hp, Hp), float("nan"), device="cuda")
    idx = torch.full((N, Co, Hp, Hp), 255, dtype=torch.uint8, device="cuda")
    native.tc_conv_relu_pool_fwd(x.cuda(), w.cuda(), b.cuda(), y, idx)
    torch.cuda.synchronize()
    assert torch.isfinite(y).all()
    assert _rel(y, want) < 4e-3, _rel(y, want)
    # argmax: torch gives flat indices into the conv map; ours is 2 bits (dy*2+dx) within the window
    Wc = H - 4
    oh, ow = want_idx // Wc, want_idx % Wc
    want2 = ((oh % 2) * 2 + (ow % 2)).to(torch.uint8)
    agree = (idx.cpu() == want2) | (want == 0)          # all-zero windows have no meaningful argmax
    assert agree.float().mean().item() > 0.995          # TF32 rounding may flip near-ties


@pytest.mark.parametrize("N,K,M,relu", [(16, 400, 120, True), (16, 120, 84, True), (16, 84, 10, False),
                                        (1, 400, 120, True), (300, 400, 120, True), (1000, 120, 84, False)])
def test_tc_linear_forward_tma_tcgen05(N, K, M, relu):
    torch.manual_seed(N + K)
    x, w, b = torch.randn(N, K), torch.randn(M, K) * K ** -0.5, torch.randn(M) * 0.1
    want = F.linear(x, w, b)
    want = F.relu(want) if relu else want
    y = torch.full((N, M), float("nan"), device="cuda")
    native.tc_linear_fwd(x.cuda(), w.cuda(), b.cuda(), y, relu)
    torch.cuda.synchronize()
    assert torch.isfinite(y).all()
    assert _rel(y, want) < 4e-3, _rel(y, want)


@pytest.mark.parametrize("N", [1, 16, 333])
def test_tc_conv2_dgrad(N):
    torch.manual_seed(N)
    x = torch.randn(N, 6, 14, 14, requires_grad=True)
    w = (torch.randn(16, 6, 5, 5) * 150 ** -0.5).requires_grad_(True)
    b = (torch.randn(16) * 0.1).requires_grad_(True)
    out = F.max_pool2d(F.relu(F.conv2d(x, w, b)), 2, 2)
    gy = torch.randn_like(out)
    out.backward(gy)
    xg = x.detach().cuda().requires_grad_(True)
    wg, bg = w.detach().cuda().requires_grad_(True), b.detach().cuda().requires_grad_(True)
    Fn.set_conv_precision("fp32")                      # exact forward (argmax identical to the oracle) ...
    try:
        yg = Fn.conv_relu_pool(xg, wg, bg)
    finally:
        Fn.set_conv_precision("tf32")
    yg.backward(gy.cuda())                             # ... tensor-core data gradient, FFMA weight gradient
    assert _rel(xg.grad, x.grad) < 4e-3, _rel(xg.grad, x.grad)
    assert _rel(wg.grad, w.grad) < 1e-4 and _rel(bg.grad, b.grad) < 1e-4


def test_network_forward_uses_tensor_cores_by_default_and_fp32_on_request():
    from distributed_neural_network_b200 import Network
    import copy
    torch.manual_seed(3)
    ref = Network()
    gpu = copy.deepcopy(ref).cuda()
    x = torch.randn(64, 3, 32, 32)
    want = ref(x)
    assert Fn.get_conv_precision() == "tf32"
    got_tf32 = gpu(x.cuda())
    Fn.set_conv_precision("fp32")
    try:
        got_fp32 = gpu(x.cuda())
    finally:
        Fn.set_conv_precision("tf32")
    assert _rel(got_fp32, want) < 2e-5
    assert 1e-7 < _rel(got_tf32, want) < 5e-3           # close, but visibly not the fp32 path


@pytest.mark.parametrize("N,K,M,relu", [(16, 400, 120, True), (100, 400, 120, True), (1000, 400, 120, False),
                                        (4096, 400, 120, True), (64, 120, 84, True), (777, 120, 84, False),
                                        (5000, 120, 84, True)])
def test_tc_linear_backward_mn_major_operands(N, K, M, relu):
    """dX and dW as tcgen05 GEMMs fed by TMA boxes of the row-major matrices (MN-major descriptors)."""
    torch.manual_seed(N + M)
    x, w, b = torch.randn(N, K), torch.randn(M, K) * K ** -0.5, torch.randn(M) * 0.1
    dy = torch.randn(N, M)
    pre = F.linear(x, w, b)
    y = F.relu(pre) if relu else pre
    d = dy * (pre > 0) if relu else dy
    want_dx, want_dw, want_db = d @ w, d.t() @ x, d.sum(0)
    dx = torch.full((N, K), float("nan"), device="cuda")
    dw = torch.full((M, K), float("nan"), device="cuda")
    db = torch.full((M,), float("nan"), device="cuda")
    native.tc_linear_bwd(x.cuda(), w.cuda(), y.cuda(), dy.cuda(), dx, dw, db, relu)
    torch.cuda.synchronize()
    assert torch.isfinite(dx).all() and torch.isfinite(dw).all() and torch.isfinite(db).all()
    assert _rel(dx, want_dx) < 4e-3, _rel(dx, want_dx)
    assert _rel(dw, want_dw) < 4e-3, _rel(dw, want_dw)
    assert _rel(db, want_db) < 1e-4, _rel(db, want_db)


def test_autograd_linear_uses_tensor_core_backward_for_large_batches():
    torch.manual_seed(3)
    x = torch.randn(512, 400)
    w = torch.randn(120, 400) * 0.05
    b = torch.randn(120) * 0.1
    xr, wr, br = x.clone().requires_grad_(True), w.clone().requires_grad_(True), b.clone().requires_grad_(True)
    F.relu(F.linear(xr, wr, br)).square().sum().backward()
    xg, wg, bg = (t.cuda().requires_grad_(True) for t in (x, w, b))
    assert Fn.get_conv_precision() == "tf32"
    Fn.linear(xg, wg, bg, relu=True).square().sum().backward()
    assert _rel(xg.grad, xr.grad) < 6e-3
    assert _rel(wg.grad, wr.grad) < 6e-3
    assert _rel(bg.grad, br.grad) < 6e-3


def test_mn_major_tf32_operand_address_function():
    """The only MN-major layout tcgen05.mma reads for 4-byte operands (128-byte swizzle, 32-byte atoms), mapped
    with a ramp image against an identity: D is the operand exactly as the tensor core sees it (csrc/tc_lab.cu)."""
    def desc(layout_type, lbo, sbo):
        return ((lbo >> 4) & 0x3FFF) << 16 | ((sbo >> 4) & 0x3FFF) << 32 | 1 << 46 | layout_type << 61

    def idesc(M, N, a_mn, b_mn):
        return (1 << 4) | (2 << 7) | (2 << 10) | (a_mn << 15) | (b_mn << 16) | ((N >> 3) << 17) | ((M >> 4) << 24)

    ident = torch.zeros(2 * 2 * 32)                    # B[16][8] = identity, canonical no-swizzle K-major
    for k in range(8):
        ident[((k >> 2) * 2 + (k >> 3)) * 32 + (k & 7) * 4 + (k & 3)] = 1.0
    words = torch.arange(4096)
    lbo, sbo = 4096, 512
    m = torch.arange(128).view(-1, 1)
    k = torch.arange(8).view(1, -1)
    want = ((m // 32) * lbo + (k // 4) * sbo + (k % 4) * 128 + ((((m % 32) // 8) ^ (k % 4)) * 32) + (m % 8) * 4) // 4
    got = torch.zeros(128, 8, dtype=torch.long)
    for shift, mask in ((0, 1023), (10, 1023)):        # two runs: TF32 keeps 10 mantissa bits exactly
        img = ((words >> shift) & mask).float().cuda()
        D = native.tc_lab(img, ident.cuda(), desc(1, lbo, sbo), desc(0, 256, 128), idesc(128, 16, 1, 0), 16)
        torch.cuda.synchronize()
        got += D[:, :8].cpu().long() << shift
    assert torch.equal(got, want)
    # the ordinary 128-byte swizzle (layout type 2) is NOT accepted for an MN-major fp32 operand: it reads zeros
    D = native.tc_lab((words & 1023).float().cuda() + 1.0, ident.cuda(), desc(2, 4096, 1024), desc(0, 256, 128),
                      idesc(128, 16, 1, 0), 16)
    torch.cuda.synchronize()
    assert float(D[:, :8].abs().max()) == 0.0


@pytest.mark.parametrize("shape", [(3, 32, 6), (6, 14, 16)])
@pytest.mark.parametrize("N", [1, 16, 300, 1111])
def test_tc_conv_weight_gradient(N, shape):
    """Tensor-core weight/bias gradient of conv+ReLU+pool (transposed im2col gathered into shared memory, reduction
    over pixels on tcgen05, accumulators kept in tensor memory over all images of a CTA) against PyTorch autograd.
    Forward and argmax come from the exact fp32 kernel, so the two sides route every gradient to the same pixel and
    only the TF32 rounding of the operands separates them."""
    Ci, H, Co = shape
    torch.manual_seed(17 * N + Ci)
    x = torch.randn(N, Ci, H, H)
    w = (torch.randn(Co, Ci, 5, 5) * (Ci * 25) ** -0.5).requires_grad_(True)
    b = (torch.randn(Co) * 0.1).requires_grad_(True)
    out = F.max_pool2d(F.relu(F.conv2d(x, w, b)), 2, 2)
    gy = torch.randn_like(out)
    out.backward(gy)
    xg, wg, bg = x.cuda(), w.detach().cuda(), b.detach().cuda()
    Hp = (H - 4) // 2
    y = torch.empty(N, Co, Hp, Hp, device="cuda")
    idx = torch.empty(N, Co, Hp, Hp, dtype=torch.uint8, device="cuda")
    native.conv_relu_pool_fwd(xg, wg, bg, y, idx)
    dw = torch.full((Co, Ci, 5, 5), float("nan"), device="cuda")
    db = torch.full((Co,), float("nan"), device="cuda")
    native.tc_conv_wgrad(xg, y, idx, gy.cuda().contiguous(), dw, db)
    torch.cuda.synchronize()
    assert _rel(dw, w.grad) < 3e-3, _rel(dw, w.grad)
    assert _rel(db, b.grad) < 2e-3, _rel(db, b.grad)
    # and it agrees with the FFMA kernel it is the alternative to
    dw2, db2 = torch.empty_like(dw), torch.empty_like(db)
    native.conv_relu_pool_bwd(xg, wg, y, idx, gy.cuda().contiguous(), None, dw2, db2)
    assert _rel(dw, dw2) < 3e-3
