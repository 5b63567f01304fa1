#!/usr/bin/env python
"""GPU baselines the framework is measured against, one process per GPU:

  * the reference's own Network class under torch DDP/NCCL (eager, cuDNN/cuBLAS), batch B per GPU
  * NCCL all-reduce of the 62 006-float gradient (248 KB): microseconds per call
  * this framework's stand-alone peer-memory all-reduce, one-shot and two-shot, over a range of sizes
  * this framework's fused step (for the same B), so every number in one file comes from the same box

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 bench/baselines.py \
        --batches 16 [--sweep] --out gpurun_out/baselines_nN.json
All timings: CUDA events on the launching stream after warm-up, synchronize on both sides, max over ranks.
"""
import argparse
import importlib.util
import json
import os
import sys
from pathlib import Path

import torch
import torch.distributed as dist
import torch.nn as nn
import torch.nn.functional as F

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))


def reference_network():
    f = ROOT / "baseline" / "_ref" / "models" / "model.py"
    if f.exists():
        spec = importlib.util.spec_from_file_location("_ref_model", f)
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
        return mod.Network(), "reference models/model.py"

    class Net(nn.Module):  # same architecture, used only if the reference checkout is absent
        def __init__(s):
            super().__init__()
            s.conv1, s.pool, s.conv2 = nn.Conv2d(3, 6, 5), nn.MaxPool2d(2, 2), nn.Conv2d(6, 16, 5)
            s.fc1, s.fc2, s.fc3 = nn.Linear(400, 120), nn.Linear(120, 84), nn.Linear(84, 10)

        def forward(s, x):
            x = s.pool(F.relu(s.conv1(x))); x = s.pool(F.relu(s.conv2(x))); x = torch.flatten(x, 1)
            return s.fc3(F.relu(s.fc2(F.relu(s.fc1(x)))))
    return Net(), "re-stated architecture"


def timed(fn, iters, warm, world):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(iters):
        fn()
    b.record()
    torch.cuda.synchronize()
    t = torch.tensor([a.elapsed_time(b) * 1e3 / iters], device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())  # us per iteration, max over ranks


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batches", default="16")
    ap.add_argument("--iters", type=int, default=300)
    ap.add_argument("--out", default=None)
    ap.add_argument("--skip-ours", action="store_true")
    a = ap.parse_args()
    rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29571")
    if world > 1:
        dist.init_process_group("cpu:gloo,cuda:nccl", rank=rank, world_size=world)
    out = {"n_gpus": world, "torch": torch.__version__, "gpu": torch.cuda.get_device_name()}
    batches = [int(b) for b in a.batches.split(",")]

    # ---- 1. reference Network under eager PyTorch (+ DDP/NCCL when world > 1) ----
    torch.manual_seed(0)
    net, src = reference_network()
    net = net.cuda()
    model = nn.parallel.DistributedDataParallel(net, device_ids=[local]) if world > 1 else net
    opt = torch.optim.SGD(model.parameters(), lr=0.001, momentum=0.9)
    out["torch_model_source"] = src
    for B in batches:
        x = torch.randn(B, 3, 32, 32, device="cuda")
        y = torch.randint(0, 10, (B,), device="cuda")

        def step():
            opt.zero_grad()
            F.cross_entropy(model(x), y).backward()
            opt.step()

        us = timed(step, a.iters, 30, world)
        out[f"torch_{'ddp_nccl' if world > 1 else 'eager'}_step_us_B{B}"] = round(us, 2)
        out[f"torch_{'ddp_nccl' if world > 1 else 'eager'}_img_per_s_B{B}"] = round(world * B / us * 1e6)

    # ---- 2. NCCL all-reduce of the gradient-sized buffer ----
    if world > 1:
        for n in (62006, 4096, 1 << 20):
            t = torch.randn(n, device="cuda")
            out[f"nccl_allreduce_us_{n}f"] = round(timed(lambda: dist.all_reduce(t), a.iters, 30, world), 2)

    # ---- 3. this framework: stand-alone all-reduce variants + fused step ----
    if not a.skip_ours:
        from distributed_neural_network_b200.trainer import TrainConfig, Trainer
        from distributed_neural_network_b200.ops import flat as L
        for B in batches:
            tr = Trainer(TrainConfig(batch_size=B, sync_mode="step-allreduce" if world > 1 else "none", synthetic=True,
                                     quiet=True, seed=1, device="cuda", peer_timeout_s=60.0))
            eng = tr.engine
            pool = 256
            dx = torch.rand(pool, B, 3, 32, 32, device="cuda") * 2 - 1
            dy = torch.randint(0, 10, (pool, B), device="cuda")
            us = timed(lambda: eng.train_device(dx, dy), 8, 2, world) / pool
            out[f"ours_fused_step_us_B{B}"] = round(us, 2)
            out[f"ours_img_per_s_B{B}"] = round(world * B / us * 1e6)
            out[f"ours_cluster_size_B{B}"] = eng.cluster_info(B)[0]
            if world > 1 and B == batches[0]:
                tr._align_sequence()
                for n in (4096, 16384, L.FLAT_NUMEL):
                    src_t = torch.randn(L.FLAT_NUMEL, device="cuda")
                    dst_t = torch.empty(L.FLAT_NUMEL, device="cuda")
                    for two in (False, True):
                        us = timed(lambda: eng.allreduce(src_t[:n], dst_t[:n], mode=0, two_shot=two), a.iters, 30, world)
                        out[f"ours_allreduce_{'two' if two else 'one'}_shot_us_{n}f"] = round(us, 2)
                if tr.nvls_ok:
                    for n in (4096, L.FLAT_NUMEL):
                        us = timed(lambda: eng.allreduce(src_t[:n], dst_t[:n], mode=0, two_shot="nvls"), a.iters, 30, world)
                        out[f"ours_allreduce_nvls_us_{n}f"] = round(us, 2)
                out["nvls_available"] = bool(tr.nvls_ok)
                if world >= 2 and rank < 2:   # one-way latency of a 16-byte line between ranks 0 and 1
                    lat = eng.pingpong(1 - rank, 2000)
                    out["ours_ll_one_way_latency_us"] = round(lat, 3)
                st = eng.read_ctl().status
                out["ours_peer_timeouts"] = int(st)
            tr.close()
    if rank == 0:
        print(json.dumps(out, indent=1))
        if a.out:
            Path(a.out).parent.mkdir(parents=True, exist_ok=True)
            Path(a.out).write_text(json.dumps(out, indent=1))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
