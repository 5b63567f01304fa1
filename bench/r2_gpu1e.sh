#!/bin/sh
# final one-GPU validation + ncu of the tensor-core layer kernels (conv forward with the lean issue loop, conv2 data
# gradient, the new conv weight gradient, linear forward)
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_n1.log 2>&1; echo "pytest exit $?"; tail -3 gpurun_out/pytest_gpu_n1.log
timeout 200 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -1
timeout 300 python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/bench_n1_k20.json 2> gpurun_out/bench_n1_k20.err; echo "bench exit $?"
python -c "
import json; d=json.load(open('gpurun_out/bench_n1_k20.json')); print(d['value'], round(d['ms_per_step']*1e3,2), 'us/step | e2e', d['e2e']['value'], round(d['e2e']['ms_per_step']*1e3,2), d['clocks']['samples'], d['gpu_launches'])" || tail -5 gpurun_out/bench_n1_k20.err
python bench/tc_once.py > gpurun_out/plain_tc.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:"tc_conv" -s 5 -c 5 -f -o gpurun_out/prof_tc_r2 python bench/tc_once.py > gpurun_out/ncu_tc_r2.log 2>&1; echo "ncu exit $?"
