"""Where a short host-fed call spends its wall clock: train_host(K) + synchronize against train_device(K) + synchronize
for several K (median of 21 calls each, idle GPU in front of every call, distinct batches). The slope is the step, the
intercept is what a call costs before its first kernel can start and after its last one has finished.
    python bench/e2e_breakdown.py [batch]"""
import json
import statistics
import sys
import time
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from distributed_neural_network_b200.trainer import TrainConfig, Trainer  # noqa: E402


def main():
    B = int(sys.argv[1]) if len(sys.argv) > 1 else 16
    tr = Trainer(TrainConfig(batch_size=B, sync_mode="none", synthetic=True, quiet=True, seed=1, lr=0.001, device="cuda"))
    eng = tr.engine
    pool = 512
    hx = (torch.rand(pool, B, 3, 32, 32) * 2 - 1).pin_memory()
    hy = torch.randint(0, 10, (pool, B)).pin_memory()
    hl = torch.zeros(pool).pin_memory()
    dx, dy = hx.cuda(), hy.cuda()
    for _ in range(3):
        eng.train_device(dx[:256], dy[:256])
        eng.train_host(hx[:256], hy[:256], hl[:256])
    torch.cuda.synchronize()
    out = {"batch": B, "unit": "us wall clock per call incl. the final synchronize, median of 21"}
    cur = 0
    for K in (1, 2, 4, 8, 20, 40, 100):
        for name in ("device", "host", "host_enqueue_only"):
            ts = []
            for _ in range(21):
                if cur + K > pool:
                    cur = 0
                torch.cuda.synchronize()
                t0 = time.perf_counter()
                if name == "device":
                    eng.train_device(dx[cur:cur + K], dy[cur:cur + K])
                else:
                    eng.train_host(hx[cur:cur + K], hy[cur:cur + K], hl[cur:cur + K])
                t1 = time.perf_counter()
                torch.cuda.synchronize()
                t2 = time.perf_counter()
                ts.append(((t1 if name == "host_enqueue_only" else t2) - t0) * 1e6)
                cur += K
            out[f"K{K}_{name}"] = round(statistics.median(ts), 1)
    print(json.dumps(out, indent=1))
    tr.close()


if __name__ == "__main__":
    main()
