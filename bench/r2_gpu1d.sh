#!/bin/sh
# one-GPU: all GPU tests, layer benchmark (incl. the tensor-core weight gradient), ncu of the exchange kernels (emulated ranks)
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_n1.log 2>&1; echo "pytest exit $?"; tail -5 gpurun_out/pytest_gpu_n1.log
timeout 600 python bench/layer_bench.py > gpurun_out/layer_bench.log 2>&1; echo "layer bench exit $?"; grep -E "wgrad|conv._fwd_N10000" gpurun_out/layer_bench.log
python bench/exchange_once.py > gpurun_out/plain_exchange.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:"update_kernel|allreduce_kernel" -c 18 -f -o gpurun_out/prof_exchange_r2 python bench/exchange_once.py > gpurun_out/ncu_exchange_r2.log 2>&1; echo "ncu exit $?"; cat gpurun_out/plain_exchange.log
