#!/bin/sh
# bench.py as the driver runs it, at every GPU the box has (final tree)
mkdir -p gpurun_out
NG=$(nvidia-smi -L | wc -l)
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $NG --master-addr 127.0.0.1 --master-port 29631 \
  bench.py --gpus $NG --steps 20 --warmup 5 2>gpurun_out/r2d_n${NG}_k20.err | tail -1 > gpurun_out/r2d_n${NG}_k20.json
python -c "
import json; d=json.load(open('gpurun_out/r2d_n${NG}_k20.json')); print('N=$NG K=20:', d['value'], round(d['ms_per_step']*1e3,2), 'us/step | e2e', d['e2e']['value'], '| xch', d.get('allreduce_us'), '| table', d.get('exchange_table',{}).get('us_per_step_max_over_ranks'), d.get('exchange_table',{}).get('chosen'), d['peer_timeouts'])" || tail -5 gpurun_out/r2d_n${NG}_k20.err
