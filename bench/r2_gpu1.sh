#!/bin/sh
# Round-2 one-GPU check: GPU tests, smoke, bench at the driver's 20 steps and at 5000 steps (they must agree).
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_n1.log 2>&1; echo "pytest exit $?"; tail -5 gpurun_out/pytest_gpu_n1.log
timeout 200 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
for K in 20 200 5000; do
  timeout 300 python bench.py --steps $K --warmup 5 > gpurun_out/r2_bench_n1_k$K.json 2> gpurun_out/r2_bench_n1_k$K.err; echo "bench K=$K exit $?"
  python -c "
import json; d=json.load(open('gpurun_out/r2_bench_n1_k$K.json')); print('K=$K', d['value'], d['ms_per_step'], 'e2e', d['e2e']['value'], d['e2e']['ms_per_step'], d['clocks'])" || tail -5 gpurun_out/r2_bench_n1_k$K.err
done
timeout 300 python bench.py --steps 20 --warmup 5 --align 5 --prep 5 > gpurun_out/r2_bench_n1_k20_noalign.json 2>/dev/null
python -c "
import json; d=json.load(open('gpurun_out/r2_bench_n1_k20_noalign.json')); print('K=20 align5 prep5', d['value'], d['ms_per_step'], 'e2e', d['e2e']['value'], d['clocks'])"
