#!/bin/sh
python -m pytest tests/test_gpu_fused_step.py -x -q 2>&1 | tail -3
for K in 20 20 5000; do
  python bench.py --steps $K --warmup 5 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('K=$K', d['value'], round(d['ms_per_step']*1e3,2), 'e2e', d['e2e']['value'], round(d['e2e']['ms_per_step']*1e3,2))"
done
