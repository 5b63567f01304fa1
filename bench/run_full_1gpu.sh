set -o pipefail
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
timeout 300 python bench.py --steps 20000 --warmup 200 > gpurun_out/bench_n1_v4.json 2> gpurun_out/bench_n1_v4.err; echo "bench exit $?"; cat gpurun_out/bench_n1_v4.json | head -c 1500
timeout 300 python bench/layer_bench.py > gpurun_out/layer_bench.log 2>&1; echo "layer bench exit $?"; tail -3 gpurun_out/layer_bench.log
timeout 120 python bench/tc_once_bwd.py && timeout 300 ncu --set full --clock-control none --import-source on -k regex:tc_linear -o gpurun_out/prof_tc_bwd_v1 -f python bench/tc_once_bwd.py > gpurun_out/ncu_tc_bwd.log 2>&1; echo "ncu exit $?"
