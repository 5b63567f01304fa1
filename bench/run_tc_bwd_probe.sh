timeout 120 python bench/tc_bwd_probe.py 0 > gpurun_out/tc_bwd_probe.json 2> gpurun_out/tc_bwd_probe.err; echo "probe exit $?"; tail -c 600 gpurun_out/tc_bwd_probe.err
timeout 300 python -m pytest tests/test_gpu_tc.py -x -q 2>&1 | tail -5
