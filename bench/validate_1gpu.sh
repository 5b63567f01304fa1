timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
timeout 200 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
timeout 300 python bench.py --steps 20000 --warmup 200 > gpurun_out/bench_n1_v6.json 2> gpurun_out/bench_n1_v6.err; echo "bench exit $?"; python -c "
import json; d=json.load(open('gpurun_out/bench_n1_v6.json')); print(d['value'], d['ms_per_step'], d['e2e']['value'], d['clocks'])"
