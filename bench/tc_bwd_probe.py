"""Numerics + timing of the tensor-core linear backward against an fp32 PyTorch reference.

``python bench/tc_bwd_probe.py [variant]`` -- one JSON line; run each variant in its own process so that a
faulting encoding cannot poison the next one.
"""
import json
import sys

import torch

sys.path.insert(0, ".")
from distributed_neural_network_b200.ops import native  # noqa: E402


def t_us(fn, iters=50):
    for _ in range(5):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(iters):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) * 1000.0 / iters


def main():
    variant = int(sys.argv[1]) if len(sys.argv) > 1 else 0
    torch.manual_seed(0)
    out = {"variant": variant}
    for name, K, M in (("fc1", 400, 120), ("fc2", 120, 84)):
        for N in (16, 100, 1000, 4096):
            for relu in (True, False):
                x = torch.randn(N, K)
                w = torch.randn(M, K) * 0.1
                b = torch.randn(M) * 0.1
                dy = torch.randn(N, M)
                y = x @ w.t() + b
                yy = torch.relu(y) if relu else y
                d = dy * (y > 0) if relu else dy
                ref_dx, ref_dw, ref_db = d @ w, d.t() @ x, d.sum(0)
                xc, wc, yc, dyc = x.cuda(), w.cuda(), yy.cuda(), dy.cuda()
                dx, dw, db = torch.full_like(xc, 7.0), torch.full_like(wc, 7.0), torch.full((M,), 7.0, device="cuda")
                native.tc_linear_bwd(xc, wc, yc, dyc, dx, dw, db, relu)
                torch.cuda.synchronize()

                def rel(a, r):
                    return float((a.cpu() - r).abs().max() / r.abs().max())

                out[f"{name}_N{N}_relu{int(relu)}"] = [round(rel(dx, ref_dx), 5), round(rel(dw, ref_dw), 5),
                                                      round(rel(db, ref_db), 6)]
    # timing
    for name, K, M in (("fc1", 400, 120), ("fc2", 120, 84)):
        for N in (16, 64, 256, 1024, 10000):
            x, w = torch.randn(N, K, device="cuda"), torch.randn(M, K, device="cuda")
            y, dy = torch.randn(N, M, device="cuda"), torch.randn(N, M, device="cuda")
            dx, dw, db = torch.empty_like(x), torch.empty_like(w), torch.empty(M, device="cuda")
            out[f"{name}_bwd_N{N}_tcgen05_us"] = round(t_us(lambda: native.tc_linear_bwd(x, w, y, dy, dx, dw, db, True)), 2)
            out[f"{name}_bwd_N{N}_ffma_us"] = round(t_us(lambda: native.linear_bwd(x, w, y, dy, dx, dw, db, True)), 2)

            def torch_bwd():
                d = dy * (y > 0)
                return d @ w, d.t() @ x, d.sum(0)

            out[f"{name}_bwd_N{N}_torch_cublas_us"] = round(t_us(torch_bwd), 2)
    print(json.dumps(out))


if __name__ == "__main__":
    main()
