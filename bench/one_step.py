"""A handful of fused steps at B=16 -- the short command that ncu wraps."""
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from distributed_neural_network_b200 import Network  # noqa: E402
from distributed_neural_network_b200.ops import flat as L  # noqa: E402
from distributed_neural_network_b200.ops import native  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 16
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 6
torch.manual_seed(0)
m = Network().cuda()
eng = native.Engine(m.flat_params, torch.zeros(L.FLAT_NUMEL, device="cuda"), None, B, 0.001, 0.9)
x = torch.randn(B, 3, 32, 32, device="cuda")
y = torch.randint(0, 10, (B,), device="cuda")
for _ in range(steps):
    eng.step(x, y)
torch.cuda.synchronize()
print("ok", eng.read_ctl().step)
