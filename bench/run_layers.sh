timeout 300 python -m pytest tests/test_gpu_layers.py tests/test_gpu_tc.py tests/test_gpu_fused_step.py -x -q 2>&1 | tail -4
timeout 300 python bench/layer_bench.py > gpurun_out/layer_bench.log 2>&1; echo "layer bench exit $?"; grep "autograd\|bwd_N10000\|bwd_N1024" gpurun_out/layer_bench.log
