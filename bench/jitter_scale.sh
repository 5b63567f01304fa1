#!/bin/bash
# step_jitter at several world sizes / exchange variants: bench/jitter_scale.sh "4 8" "one-shot two-shot"
for n in $1; do for v in $2; do
  echo "== N=$n $v"
  DNN_EXCHANGE=$v timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29511 bench/step_jitter.py 2>&1 | grep "^rank [01]/" | grep -v "first update\|tail of"
done; done
