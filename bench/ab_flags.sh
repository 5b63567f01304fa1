#!/bin/bash
# A/B of engine debug flags: step time at 1 and N GPUs per flag value (8 = no counter-based dependency)
N=${1:-2}
shift
for fl in "$@"; do
  echo "== DNN_DEBUG_FLAGS=$fl"
  DNN_DEBUG_FLAGS=$fl timeout 120 python bench/step_jitter.py 2>&1 | grep "^rank"
  DNN_DEBUG_FLAGS=$fl timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench/step_jitter.py 2>&1 | grep "^rank"
done
