#!/bin/sh
NG=$(nvidia-smi -L | wc -l)
for G in 16 4 2 1; do
  for K in 20 3000; do
  DNN_HOST_GROUP=$G timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $NG --master-addr 127.0.0.1 --master-port 29604 bench.py --gpus $NG --steps $K --warmup 5 --exchange one-shot 2>/dev/null | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('G=$G K=$K', round(d['ms_per_step']*1e3,2), 'e2e', round(d['e2e']['ms_per_step']*1e3,2), d['e2e']['value'])"
  done
done
