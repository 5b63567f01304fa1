#!/bin/sh
# final tree at 8 GPUs: the runtime's own NVLS group at world 8, stand-alone all-reduce latency table, bench.py as the driver runs it
mkdir -p gpurun_out
NG=$(nvidia-smi -L | wc -l)
timeout 400 python -m pytest tests/test_gpu_multiproc.py -x -q -k "nvls" 2>&1 | tail -4 > gpurun_out/pytest_nvls_n$NG.log; cat gpurun_out/pytest_nvls_n$NG.log
timeout 200 bin/mpiexec -n $NG python bench/allreduce_latency.py 400 > gpurun_out/allreduce_latency_n$NG.json 2> gpurun_out/allreduce_latency_n$NG.err
cat gpurun_out/allreduce_latency_n$NG.json; tail -3 gpurun_out/allreduce_latency_n$NG.err
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $NG --master-addr 127.0.0.1 --master-port 29611 \
  bench.py --gpus $NG --steps 20 --warmup 5 2>gpurun_out/r2d_n${NG}_k20.err | tail -1 > gpurun_out/r2d_n${NG}_k20.json
python -c "
import json; d=json.load(open('gpurun_out/r2d_n${NG}_k20.json')); print('N=$NG K=20:', d['value'], 'img/s', round(d['ms_per_step']*1e3,2), 'us/step | e2e', d['e2e']['value'], '| xch', d.get('allreduce_us'), '| chosen', d.get('exchange_table',{}).get('chosen'), d.get('exchange_table',{}).get('us_per_step'), '| clk', d['clocks'].get('sm_mhz'), d['peer_timeouts'])" || tail -5 gpurun_out/r2d_n${NG}_k20.err
