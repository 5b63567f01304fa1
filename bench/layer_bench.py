"""Layer kernels: tcgen05 (TF32) vs FFMA (fp32) vs PyTorch/cuDNN/cuBLAS, and eval throughput. Device-timed."""
import json
import sys
from pathlib import Path

import torch
import torch.nn.functional as F

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from distributed_neural_network_b200 import Network  # noqa: E402
from distributed_neural_network_b200.ops import flat as L  # noqa: E402
from distributed_neural_network_b200.ops import functional as Fn  # noqa: E402
from distributed_neural_network_b200.ops import native  # noqa: E402


def t_us(fn, iters=50, warm=10):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(iters):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) * 1e3 / iters


def main():
    out = {}
    torch.manual_seed(0)
    for N in (16, 1024, 10000):
        for name, (Ci, H, Co) in (("conv1", (3, 32, 6)), ("conv2", (6, 14, 16))):
            x = torch.randn(N, Ci, H, H, device="cuda")
            w = torch.randn(Co, Ci, 5, 5, device="cuda") * 0.1
            b = torch.randn(Co, device="cuda") * 0.1
            Hp = (H - 4) // 2
            y = torch.empty(N, Co, Hp, Hp, device="cuda")
            idx = torch.empty(N, Co, Hp, Hp, dtype=torch.uint8, device="cuda")
            out[f"{name}_fwd_N{N}_tcgen05_us"] = round(t_us(lambda: native.tc_conv_relu_pool_fwd(x, w, b, y, idx)), 2)
            out[f"{name}_fwd_N{N}_ffma_us"] = round(t_us(lambda: native.conv_relu_pool_fwd(x, w, b, y, idx)), 2)
            out[f"{name}_fwd_N{N}_torch_cudnn_us"] = round(t_us(lambda: F.max_pool2d(F.relu(F.conv2d(x, w, b)), 2, 2)), 2)
            # weight + bias gradient (K16 / K17): pooled-grid FFMA kernel, tensor-core kernel, cuDNN (dense gradient given)
            native.conv_relu_pool_fwd(x, w, b, y, idx)
            dyp = torch.randn_like(y)
            dw, db = torch.empty_like(w), torch.empty_like(b)
            out[f"{name}_wgrad_N{N}_ffma_pooled_grid_us"] = round(
                t_us(lambda: native.conv_relu_pool_bwd(x, w, y, idx, dyp, None, dw, db)), 2)
            out[f"{name}_wgrad_N{N}_tcgen05_us"] = round(t_us(lambda: native.tc_conv_wgrad(x, y, idx, dyp, dw, db)), 2)
            dense = torch.zeros(N, Co, H - 4, H - 4, device="cuda")
            out[f"{name}_wgrad_N{N}_torch_cudnn_dense_us"] = round(
                t_us(lambda: torch.nn.grad.conv2d_weight(x, w.shape, dense)), 2)
        for name, (K, M) in (("fc1", (400, 120)), ("fc2", (120, 84))):
            x = torch.randn(N, K, device="cuda")
            w = torch.randn(M, K, device="cuda") * 0.1
            b = torch.randn(M, device="cuda")
            y = torch.empty(N, M, device="cuda")
            out[f"{name}_fwd_N{N}_tcgen05_tma_us"] = round(t_us(lambda: native.tc_linear_fwd(x, w, b, y, True)), 2)
            out[f"{name}_fwd_N{N}_ffma_us"] = round(t_us(lambda: native.linear_fwd(x, w, b, y, True)), 2)
            out[f"{name}_fwd_N{N}_torch_cublas_us"] = round(t_us(lambda: F.relu(F.linear(x, w, b))), 2)
            dy = torch.randn(N, M, device="cuda")
            dx, dw, db = torch.empty_like(x), torch.empty_like(w), torch.empty(M, device="cuda")
            out[f"{name}_bwd_N{N}_tcgen05_tma_us"] = round(t_us(lambda: native.tc_linear_bwd(x, w, y, dy, dx, dw, db, True)), 2)
            out[f"{name}_bwd_N{N}_ffma_us"] = round(t_us(lambda: native.linear_bwd(x, w, y, dy, dx, dw, db, True)), 2)

            def torch_bwd():
                d = dy * (y > 0)
                return d @ w, d.t() @ x, d.sum(0)
            out[f"{name}_bwd_N{N}_torch_cublas_us"] = round(t_us(torch_bwd), 2)
    # one forward+backward of the whole model through autograd at large batches: our layer kernels vs PyTorch ops
    for N in (1024, 8192):
        m = Network().cuda()
        x = torch.randn(N, 3, 32, 32, device="cuda")
        y = torch.randint(0, 10, (N,), device="cuda")

        def ours():
            m.zero_grad()
            Fn.cross_entropy(m(x), y).backward()

        def eager():
            m.zero_grad()
            h = F.max_pool2d(F.relu(F.conv2d(x, m.conv1.weight, m.conv1.bias)), 2, 2)
            h = F.max_pool2d(F.relu(F.conv2d(h, m.conv2.weight, m.conv2.bias)), 2, 2).flatten(1)
            h = F.relu(F.linear(h, m.fc1.weight, m.fc1.bias))
            h = F.relu(F.linear(h, m.fc2.weight, m.fc2.bias))
            F.cross_entropy(F.linear(h, m.fc3.weight, m.fc3.bias), y).backward()
        out[f"autograd_fwd_bwd_N{N}_layer_kernels_us"] = round(t_us(ours, 20, 3), 1)
        out[f"autograd_fwd_bwd_N{N}_torch_eager_us"] = round(t_us(eager, 20, 3), 1)
    # eval of 10 000 test images (reference eval loop semantics), batch 16
    m = Network().cuda()
    eng = native.Engine(m.flat_params, torch.zeros(L.FLAT_NUMEL, device="cuda"), None, 16, 0.001, 0.9)
    x = torch.randn(10000, 3, 32, 32, device="cuda")
    y = torch.randint(0, 10, (10000,), device="cuda")
    out["eval_10000_fused_engine_us"] = round(t_us(lambda: eng.evaluate(x, y, 16), 10, 3), 1)

    def torch_eval():
        with torch.no_grad():
            for i in range(0, 10000, 16):
                o = m(x[i:i + 16])
                F.cross_entropy(o, y[i:i + 16]).item()
    out["eval_10000_layer_kernels_bs16_with_item_us"] = round(t_us(torch_eval, 2, 1), 1)
    print(json.dumps(out, indent=1))
    Path("gpurun_out").mkdir(exist_ok=True)
    Path("gpurun_out/layer_bench.json").write_text(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
