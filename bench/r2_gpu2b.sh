#!/bin/sh
mkdir -p gpurun_out
NG=$(nvidia-smi -L | wc -l)
timeout 900 python -m pytest tests/test_gpu_multiproc.py tests/test_zz_no_alloc.py -v -x > gpurun_out/pytest_gpu_n$NG.log 2>&1; echo "pytest exit $?"; tail -4 gpurun_out/pytest_gpu_n$NG.log
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $NG --master-addr 127.0.0.1 --master-port 29677 bench/autograd_overlap.py 2>&1 | tail -2
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $NG --master-addr 127.0.0.1 --master-port 29655 bench/fault_drill.py 2>&1 | tail -2 | cut -c1-400
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $NG --master-addr 127.0.0.1 --master-port 29602 bench.py --gpus $NG --steps 20 --warmup 5 2>/dev/null | tail -1 > gpurun_out/r2b_bench_n$NG.json
python -c "
import json; d=json.load(open('gpurun_out/r2b_bench_n$NG.json')); print(d['value'], round(d['ms_per_step']*1e3,2), 'e2e', d['e2e']['value'], 'xch', d.get('allreduce_us'), d['exchange']['nvlink_counters_bytes_per_step_per_gpu'], d['exchange']['bytes_per_direction_per_gpu'], d['clocks'])"
