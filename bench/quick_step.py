"""Quick device-timed microbenchmarks of the fused step (development aid, not the contract bench)."""
import json
import sys
import time
from pathlib import Path

import torch
import torch.nn.functional as F

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from distributed_neural_network_b200 import Network  # noqa: E402
from distributed_neural_network_b200.ops import flat as L  # noqa: E402
from distributed_neural_network_b200.ops import native  # noqa: E402


def time_steps(fn, steps, warmup=20):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(steps):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) * 1e3 / steps  # us


def main():
    out = {}
    torch.manual_seed(0)
    for B in (1, 4, 16, 64, 256):
        m = Network().cuda()
        eng = native.Engine(m.flat_params, torch.zeros(L.FLAT_NUMEL, device="cuda"), None, B, 0.001, 0.9)
        x = torch.randn(B, 3, 32, 32, device="cuda")
        y = torch.randint(0, 10, (B,), device="cuda")
        us = time_steps(lambda: eng.step(x, y), 500)
        out[f"fused_step_us_B{B}"] = round(us, 2)
        out[f"fused_img_per_s_B{B}"] = round(B / us * 1e6)
        out[f"cluster_B{B}"] = str(eng.cluster_info(B))
        if B in (16, 64):
            for cs in (1, 2, 3, 4, 5, 6, 7, 8):
                eng.set_cluster_size(cs)
                out[f"fused_step_us_B{B}_cs{cs}"] = round(time_steps(lambda: eng.step(x, y), 300), 2)
            eng.set_cluster_size(0)
        if B == 16:
            steps = 512
            hx = torch.randn(steps, B, 3, 32, 32).pin_memory()
            hy = torch.randint(0, 10, (steps, B)).pin_memory()
            hl = torch.zeros(steps).pin_memory()
            eng.train_host(hx[:32], hy[:32], hl[:32])
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            eng.train_host(hx, hy, hl)
            b.record()
            t1 = time.perf_counter()
            torch.cuda.synchronize()
            t2 = time.perf_counter()
            out["train_host_us_per_step_B16"] = round(a.elapsed_time(b) * 1e3 / steps, 2)
            out["train_host_enqueue_us_per_step"] = round((t1 - t0) * 1e6 / steps, 2)
            out["train_host_wall_us_per_step"] = round((t2 - t0) * 1e6 / steps, 2)
        eng.close()
    # torch eager baseline (cuDNN/cuBLAS) on the same architecture, same batch
    for B in (16, 64):
        m = Network()
        import torch.nn as nn

        class Ref(nn.Module):
            def __init__(s):
                super().__init__()
                s.conv1, s.pool, s.conv2 = nn.Conv2d(3, 6, 5), nn.MaxPool2d(2, 2), nn.Conv2d(6, 16, 5)
                s.fc1, s.fc2, s.fc3 = nn.Linear(400, 120), nn.Linear(120, 84), nn.Linear(84, 10)

            def forward(s, x):
                x = s.pool(F.relu(s.conv1(x))); x = s.pool(F.relu(s.conv2(x))); x = torch.flatten(x, 1)
                return s.fc3(F.relu(s.fc2(F.relu(s.fc1(x)))))

        r = Ref().cuda()
        opt = torch.optim.SGD(r.parameters(), lr=0.001, momentum=0.9)
        x = torch.randn(B, 3, 32, 32, device="cuda")
        y = torch.randint(0, 10, (B,), device="cuda")

        def eager():
            opt.zero_grad()
            F.cross_entropy(r(x), y).backward()
            opt.step()

        us = time_steps(eager, 200)
        out[f"torch_eager_step_us_B{B}"] = round(us, 2)
    print(json.dumps(out, indent=1))
    Path("gpurun_out").mkdir(exist_ok=True)
    Path("gpurun_out/quick_step.json").write_text(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
