#!/bin/sh
# Round-2 two-GPU check: multi-process tests (kept as a log), bench at 20 and 5000 steps per exchange variant.
mkdir -p gpurun_out
NG=$(nvidia-smi -L | wc -l)
timeout 900 python -m pytest tests/test_gpu_multiproc.py tests/test_gpu_exchange.py tests/test_zz_no_alloc.py -v -x > gpurun_out/pytest_gpu_n$NG.log 2>&1; echo "pytest exit $?"; tail -4 gpurun_out/pytest_gpu_n$NG.log
run() {  # N K exchange tag
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $1 --master-addr 127.0.0.1 --master-port $((29600 + $1)) \
    bench.py --gpus $1 --steps $2 --warmup 5 --exchange $3 2>gpurun_out/r2_bench_n$1_k$2_$3.err | tail -1 > gpurun_out/r2_bench_n$1_k$2_$3.json
  python -c "
import json; d=json.load(open('gpurun_out/r2_bench_n$1_k$2_$3.json')); print('N=$1 K=$2 $3:', d['value'], 'img/s', round(d['ms_per_step']*1e3,2), 'us/step | e2e', d['e2e']['value'], round(d['e2e']['ms_per_step']*1e3,2), '| xch', d.get('allreduce_us'), d.get('exchange_roofline_frac'), '| table', d.get('exchange_table',{}).get('us_per_step_max_over_ranks'), d.get('exchange_table',{}).get('chosen'), '| clk', d['clocks'].get('sm_mhz'), d['clocks'].get('samples'), d['peer_timeouts'])" || tail -5 gpurun_out/r2_bench_n$1_k$2_$3.err
}
for N in 2 4 8; do
  [ "$N" -gt "$NG" ] && break
  run $N 20 auto
  run $N 5000 auto
  for V in one-shot one-shot-128 two-shot nvls; do run $N 2000 $V; done
done
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $NG --master-addr 127.0.0.1 --master-port 29655 bench/fault_drill.py 2>&1 | tail -6
