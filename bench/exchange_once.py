"""Short command for the ncu capture of the kernels that exchange: four ranks emulated on ONE GPU (one cooperative
launch per step, as in tests/test_gpu_exchange.py), three steps of the fused update kernel and three stand-alone
all-reduces per exchange variant. ncu replays a kernel ~40 times, which only a single-GPU launch survives (ranks in
different processes would wait for replays that never come), so what this shows is the instruction mix, the stall
reasons and the memory traffic of the exchange code itself; the bytes that cross NVLink in a real run are counted by
`bench.py` from the NVML NVLink counters (`exchange.nvlink_counters_bytes_per_step_per_gpu`).
   ncu -k regex:"update_kernel|allreduce_kernel" ...: launches come in the order one-shot, two-shot, one-shot-128."""
import copy
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from distributed_neural_network_b200 import Network  # noqa: E402
from distributed_neural_network_b200.ops import flat as L  # noqa: E402
from distributed_neural_network_b200.ops import native  # noqa: E402


def main():
    world, B = 4, 16
    torch.manual_seed(0)
    base = Network()
    models = [copy.deepcopy(base).cuda() for _ in range(world)]
    groups = [native.Group(r, world) for r in range(world)]
    native.Group.connect_local(groups)
    engines = []
    for r in range(world):
        e = native.Engine(models[r].flat_params, torch.zeros(L.FLAT_NUMEL, device="cuda"), None, B, 0.01, 0.9, precision="fp32")
        e.attach_group(groups[r])
        e.set_membership((1 << world) - 1, timeout_ns=2_000_000_000)
        engines.append(e)
    xs = [torch.randn(B, 3, 32, 32, device="cuda") for _ in range(world)]
    ys = [torch.randint(0, 10, (B,), device="cuda") for _ in range(world)]
    for variant in ("one-shot", "two-shot", "one-shot-128"):
        for e in engines:
            e.set_exchange_variant(variant)
        for _ in range(3):
            native.vr_step(engines, xs, ys, B)
        ins = [torch.randn(L.FLAT_NUMEL, device="cuda") for _ in range(world)]
        outs = [torch.empty(L.FLAT_NUMEL, device="cuda") for _ in range(world)]
        for _ in range(3):
            native.vr_allreduce(engines, ins, outs, L.FLAT_NUMEL, mode=0, two_shot=variant)
        torch.cuda.synchronize()
        same = all(torch.equal(models[0].flat_params, m.flat_params) for m in models)
        print(variant, "replicas identical:", same, "status", [e.read_ctl().status for e in engines])
    for e in engines:
        e.close()
    for g in groups:
        g.close()


if __name__ == "__main__":
    main()
