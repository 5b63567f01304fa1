#!/bin/bash
# A/B of environment settings of bench/step_jitter.py at 1 and N GPUs, e.g. the open question of DESIGN.md 9.3:
#   bench/ab_env.sh 2 "DNN_CLUSTER=6" "DNN_CLUSTER=5 DNN_DEBUG_FLAGS=32"
N=${1:-2}
shift
for cfg in "$@"; do
  echo "== $cfg"
  env $cfg timeout 120 python bench/step_jitter.py 2>&1 | grep "^rank" | head -1
  [ "$N" -gt 1 ] && env $cfg timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench/step_jitter.py 2>&1 | grep "^rank [01]/" | grep -v "first update\|tail of"
done
