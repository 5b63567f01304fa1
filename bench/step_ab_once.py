"""Twelve chained steps of the fused engine at batch 16 in each precision (fp32 FFMA, then TF32 tensor cores): the short
command the ncu capture of both instantiations of fwd_bwd_cluster_kernel is taken from
(ncu -k regex:fwd_bwd_cluster_kernel -s 10 -c 4 -> launches 10, 11 are fp32, 12 is the first TF32 step, which packs
its B operands in shared memory, 13 the first one that fetches them ready-made)."""
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from distributed_neural_network_b200 import Network  # noqa: E402
from distributed_neural_network_b200.ops import flat as L  # noqa: E402
from distributed_neural_network_b200.ops import native  # noqa: E402


def main():
    torch.manual_seed(0)
    B, steps = 16, 12
    x = torch.randn(steps, B, 3, 32, 32, device="cuda")
    y = torch.randint(0, 10, (steps, B), device="cuda")
    for prec in ("fp32", "tf32"):
        m = Network().cuda()
        eng = native.Engine(m.flat_params, torch.zeros(L.FLAT_NUMEL, device="cuda"), None, B, 0.001, 0.9, precision=prec)
        eng.train_device(x, y)
        torch.cuda.synchronize()
        print(prec, "loss", eng.read_ctl().loss_sum / steps)
        eng.close()


if __name__ == "__main__":
    main()
