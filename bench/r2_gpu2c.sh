#!/bin/sh
# final tree, two GPUs: every multi-process / multi-device test, log kept
mkdir -p gpurun_out
timeout 800 python -m pytest tests/test_gpu_multiproc.py tests/test_zz_no_alloc.py tests/test_gpu_exchange.py -q -v 2>&1 | tail -70 > gpurun_out/pytest_gpu_n2_r2c.log
tail -5 gpurun_out/pytest_gpu_n2_r2c.log
