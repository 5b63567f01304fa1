#!/bin/sh
# Final eight-GPU run of round 2: bench.py as the driver calls it at N = 8, 4, 2, 1 (+ a long run at 8), the fault drill,
# the overlap A/B of the autograd path, and the whole GPU test suite (multi-process tests included) as a kept log.
mkdir -p gpurun_out
NG=$(nvidia-smi -L | wc -l)
run() {  # N K
  if [ "$1" -eq 1 ]; then
    timeout 200 python bench.py --gpus 1 --steps $2 --warmup 5 2>gpurun_out/r2b_scale_n$1_k$2.err | tail -1 > gpurun_out/r2b_scale_n$1_k$2.json
  else
    timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $1 --master-addr 127.0.0.1 --master-port $((29600 + $1)) \
      bench.py --gpus $1 --steps $2 --warmup 5 2>gpurun_out/r2b_scale_n$1_k$2.err | tail -1 > gpurun_out/r2b_scale_n$1_k$2.json
  fi
  python -c "
import json; d=json.load(open('gpurun_out/r2b_scale_n$1_k$2.json')); print('N=$1 K=$2:', d['value'], 'img/s', round(d['ms_per_step']*1e3,2), 'us/step | e2e', d['e2e']['value'], round(d['e2e']['ms_per_step']*1e3,2), '| xch', d.get('allreduce_us'), d.get('exchange_roofline_frac'), '| table', d.get('exchange_table',{}).get('us_per_step_max_over_ranks'), d.get('exchange_table',{}).get('chosen'), '| nvlink', (d.get('exchange') or {}).get('nvlink_counters_bytes_per_step_per_gpu'), '| clk', d['clocks'].get('sm_mhz'), d['clocks'].get('samples'), d['peer_timeouts'])" || tail -5 gpurun_out/r2b_scale_n$1_k$2.err
}
for N in 8 4 2 1; do
  [ "$N" -gt "$NG" ] && continue
  run $N 20
done
run $NG 5000
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $NG --master-addr 127.0.0.1 --master-port 29655 bench/fault_drill.py 2>&1 | tail -2 | cut -c1-500
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $NG --master-addr 127.0.0.1 --master-port 29677 bench/autograd_overlap.py 2>&1 | tail -1
timeout 600 python -m pytest tests -m gpu -v > gpurun_out/pytest_gpu_n$NG.log 2>&1; echo "pytest exit $?"; tail -4 gpurun_out/pytest_gpu_n$NG.log
