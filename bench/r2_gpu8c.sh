#!/bin/sh
# bench.py as the driver calls it at N = 8, 4, 2, 1, plus a long run at 8 and at 1 (final tree of round 2)
mkdir -p gpurun_out
NG=$(nvidia-smi -L | wc -l)
run() {  # N K
  if [ "$1" -eq 1 ]; then
    timeout 200 python bench.py --gpus 1 --steps $2 --warmup 5 2>gpurun_out/r2c_scale_n$1_k$2.err | tail -1 > gpurun_out/r2c_scale_n$1_k$2.json
  else
    timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $1 --master-addr 127.0.0.1 --master-port $((29600 + $1)) \
      bench.py --gpus $1 --steps $2 --warmup 5 2>gpurun_out/r2c_scale_n$1_k$2.err | tail -1 > gpurun_out/r2c_scale_n$1_k$2.json
  fi
  python -c "
import json; d=json.load(open('gpurun_out/r2c_scale_n$1_k$2.json')); print('N=$1 K=$2:', d['value'], 'img/s', round(d['ms_per_step']*1e3,2), 'us/step | e2e', d['e2e']['value'], round(d['e2e']['ms_per_step']*1e3,2), '| xch', d.get('allreduce_us'), d.get('exchange_roofline_frac'), '| chosen', d.get('exchange_table',{}).get('chosen'), '| clk', d['clocks'].get('sm_mhz'), d['clocks'].get('samples'), d['peer_timeouts'])" || tail -5 gpurun_out/r2c_scale_n$1_k$2.err
}
for N in 8 4 2 1; do
  [ "$N" -gt "$NG" ] && continue
  run $N 20
done
run $NG 5000
run 1 5000
