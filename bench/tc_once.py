"""Short command for ncu: the tensor-core layer kernels at eval-sized batches."""
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from distributed_neural_network_b200.ops import native  # noqa: E402

N = 4096
for _ in range(3):
    x = torch.randn(N, 3, 32, 32, device="cuda"); w = torch.randn(6, 3, 5, 5, device="cuda") * 0.1; b = torch.zeros(6, device="cuda")
    y = torch.empty(N, 6, 14, 14, device="cuda"); idx = torch.empty(N, 6, 14, 14, dtype=torch.uint8, device="cuda")
    native.tc_conv_relu_pool_fwd(x, w, b, y, idx)
    w2 = torch.randn(16, 6, 5, 5, device="cuda") * 0.1; b2 = torch.zeros(16, device="cuda")
    y2 = torch.empty(N, 16, 5, 5, device="cuda"); idx2 = torch.empty(N, 16, 5, 5, dtype=torch.uint8, device="cuda")
    native.tc_conv_relu_pool_fwd(y, w2, b2, y2, idx2)
    dx = torch.empty(N, 6, 14, 14, device="cuda"); dw = torch.empty_like(w2); db = torch.empty(16, device="cuda")
    native.conv_relu_pool_bwd(y, w2, y2, idx2, torch.randn_like(y2), dx, dw, db, tensor_cores=True)
    native.tc_conv_wgrad(y, y2, idx2, torch.randn_like(y2), dw, db)               # conv2 weight gradient on tcgen05
    dw1, db1 = torch.empty_like(w), torch.empty(6, device="cuda")
    native.tc_conv_wgrad(x, y, idx, torch.randn_like(y), dw1, db1)                # conv1 weight gradient on tcgen05
    f = y2.reshape(N, 400).contiguous(); w3 = torch.randn(120, 400, device="cuda") * 0.05; b3 = torch.zeros(120, device="cuda")
    h = torch.empty(N, 120, device="cuda")
    native.tc_linear_fwd(f, w3, b3, h, True)
torch.cuda.synchronize()
print("ok")
