"""Short command for ncu: the tensor-core linear backward (fc1 shapes) at an eval-sized batch."""
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from distributed_neural_network_b200.ops import native  # noqa: E402

N, K, M = 8192, 400, 120
x, w = torch.randn(N, K, device="cuda"), torch.randn(M, K, device="cuda") * 0.05
y, dy = torch.randn(N, M, device="cuda"), torch.randn(N, M, device="cuda")
dx, dw, db = torch.empty_like(x), torch.empty_like(w), torch.empty(M, device="cuda")
for _ in range(3):
    native.tc_linear_bwd(x, w, y, dy, dx, dw, db, True)
torch.cuda.synchronize()
print("ok")
