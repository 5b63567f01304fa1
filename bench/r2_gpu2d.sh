#!/bin/sh
# two GPUs: multi-process tests after the start-up probe of the 128-byte lines went into the Trainer, bench.py at N=2
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_multiproc.py -q 2>&1 | tail -4 > gpurun_out/pytest_gpu_n2_r2d.log; cat gpurun_out/pytest_gpu_n2_r2d.log
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29621 \
  bench.py --gpus 2 --steps 20 --warmup 5 2>gpurun_out/r2d_n2_k20.err | tail -1 > gpurun_out/r2d_n2_k20.json
python -c "
import json; d=json.load(open('gpurun_out/r2d_n2_k20.json')); print('N=2 K=20:', d['value'], round(d['ms_per_step']*1e3,2), 'us/step | e2e', d['e2e']['value'], '| xch', d.get('allreduce_us'), '| table', d.get('exchange_table'), d['peer_timeouts'])" || tail -5 gpurun_out/r2d_n2_k20.err
