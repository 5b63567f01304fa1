"""Short command for compute-sanitizer (memcheck / racecheck / synccheck): every kernel family once, small shapes."""
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from distributed_neural_network_b200 import Network  # noqa: E402
from distributed_neural_network_b200.ops import flat as L  # noqa: E402
from distributed_neural_network_b200.ops import functional as Fn  # noqa: E402
from distributed_neural_network_b200.ops import native  # noqa: E402


def main():
    torch.manual_seed(0)
    dev = torch.device("cuda", 0)
    # fused engine: clustered and one-CTA forward/backward, update kernel, chained steps, host-fed ring, eval
    m = Network().to(dev)
    eng = native.Engine(m.flat_params, torch.zeros(L.FLAT_NUMEL, device=dev), None, 16, 0.01, 0.9)
    x = torch.randn(4, 16, 3, 32, 32, device=dev)
    y = torch.randint(0, 10, (4, 16), device=dev)
    for cs in (0, 4, 2, 1):
        eng.set_cluster_size(cs)
        eng.step(x[0], y[0])
    eng.set_cluster_size(0)
    eng.train_device(x, y)
    hx, hy, hl = x.cpu().pin_memory(), y.cpu().pin_memory(), torch.zeros(4).pin_memory()
    eng.train_host(hx, hy, hl)
    g = torch.zeros(L.FLAT_NUMEL, device=dev)
    eng.grad_only(x[1], y[1], g)
    eng.evaluate(x.reshape(64, 3, 32, 32), y.reshape(64), 16)
    torch.cuda.synchronize()
    eng.close()
    # layer kernels through autograd: FFMA (fp32) at a small batch, tensor cores (TF32) at 300 rows
    for prec, n in (("fp32", 16), ("tf32", 300)):
        Fn.set_conv_precision(prec)
        net = Network().to(dev)
        xb = torch.randn(n, 3, 32, 32, device=dev, requires_grad=True)
        yb = torch.randint(0, 10, (n,), device=dev)
        Fn.cross_entropy(net(xb), yb).backward()
        torch.cuda.synchronize()
    Fn.set_conv_precision("tf32")
    # stand-alone pieces
    p, gr, buf = (torch.randn(1000, device=dev) for _ in range(3))
    Fn.sgd_momentum_(p, gr, buf, 0.01, 0.9, False)
    A, B = torch.randn(128, 16, device=dev), torch.randn(16, 16, device=dev)
    native.tc_gemm_probe(A, B, torch.empty(128, 16, device=dev))
    torch.cuda.synchronize()
    print("ok")


if __name__ == "__main__":
    main()
