#!/bin/sh
# one-GPU check of the current tree: all GPU tests, the tensor-core A/B of the step, its stage trace
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_n1.log 2>&1; echo "pytest exit $?"; tail -6 gpurun_out/pytest_gpu_n1.log
for P in fp32 tf32; do
  timeout 300 python bench.py --steps 2000 --warmup 5 --precision $P > gpurun_out/r2_bench_n1_$P.json 2> gpurun_out/r2_bench_n1_$P.err; echo "bench $P exit $?"
  python -c "
import json; d=json.load(open('gpurun_out/r2_bench_n1_$P.json')); print('$P', d['value'], round(d['ms_per_step']*1e3,2), 'us/step | e2e', d['e2e']['value'], d['dtype'][:40])" || tail -5 gpurun_out/r2_bench_n1_$P.err
done
python bench/trace_stages.py 16 6 tf32 | head -24
