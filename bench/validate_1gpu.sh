#!/bin/sh
# what is run on a one-GPU box before a commit: GPU tests, smoke(), bench.py as the driver calls it
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_n1.log 2>&1; echo "pytest exit $?"; tail -3 gpurun_out/pytest_gpu_n1.log
timeout 200 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
timeout 300 python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/bench_n1_k20.json 2> gpurun_out/bench_n1_k20.err; echo "bench exit $?"
python -c "
import json; d=json.load(open('gpurun_out/bench_n1_k20.json')); print(d['value'], round(d['ms_per_step']*1e3,2), 'us/step | e2e', d['e2e']['value'], round(d['e2e']['ms_per_step']*1e3,2), d['clocks'], d['gpu_launches'], d['dtype'][:30])" || tail -5 gpurun_out/bench_n1_k20.err
