#!/bin/sh
mkdir -p gpurun_out
python -m pytest tests/test_gpu_tc.py -x -q -k "weight_gradient" 2>&1 | tail -2
python bench/layer_bench.py > gpurun_out/layer_bench.log 2>&1; grep -E "wgrad" gpurun_out/layer_bench.log
