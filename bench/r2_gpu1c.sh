#!/bin/sh
# one-GPU check + evidence: all GPU tests, smoke, A/B of the two precisions of the step, launch list, ncu of both step kernels
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_n1.log 2>&1; echo "pytest exit $?"; tail -5 gpurun_out/pytest_gpu_n1.log
timeout 200 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
for P in fp32 tf32; do
  for K in 20 5000; do
    timeout 300 python bench.py --steps $K --warmup 5 --precision $P > gpurun_out/r2_bench_n1_${P}_k$K.json 2> gpurun_out/r2_bench_n1_${P}_k$K.err
    python -c "
import json; d=json.load(open('gpurun_out/r2_bench_n1_${P}_k$K.json')); print('$P K=$K', d['value'], round(d['ms_per_step']*1e3,2), 'us/step | e2e', d['e2e']['value'], round(d['e2e']['ms_per_step']*1e3,2), d['clocks'])" || tail -5 gpurun_out/r2_bench_n1_${P}_k$K.err
  done
done
python bench/tc_mma_rate.py > gpurun_out/tc_mma_rate.txt 2>&1; tail -3 gpurun_out/tc_mma_rate.txt
python bench/step_ab_once.py > gpurun_out/plain_ab.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:fwd_bwd_cluster_kernel -s 10 -c 4 -f -o gpurun_out/prof_step_r2 python bench/step_ab_once.py > gpurun_out/ncu_step_r2.log 2>&1; echo "ncu exit $?"; tail -3 gpurun_out/plain_ab.log
