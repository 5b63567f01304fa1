#!/bin/sh
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_tc_step.py -x -q > gpurun_out/pytest_tc_step.log 2>&1; echo "pytest tc exit $?"; tail -3 gpurun_out/pytest_tc_step.log
for P in fp32 tf32-dgrad fp32 tf32-dgrad; do
  timeout 300 python bench.py --steps 5000 --warmup 5 --precision $P > gpurun_out/r2_bench_n1_$P.json 2> gpurun_out/r2_bench_n1_$P.err
  python -c "
import json; d=json.load(open('gpurun_out/r2_bench_n1_$P.json')); print('$P', d['value'], round(d['ms_per_step']*1e3,2), 'us/step | e2e', d['e2e']['value'])" || tail -5 gpurun_out/r2_bench_n1_$P.err
done
DNNB200_PRECISION=fp32 python bench/step_jitter.py 2>&1 | grep "rank 0" | head -2
DNNB200_PRECISION=tf32-dgrad python bench/step_jitter.py 2>&1 | grep "rank 0" | head -2
