#!/bin/sh
# one-shot vs two-shot fused exchange at N = 2, 4, 8 GPUs (device-timed step, max over ranks)
STEPS=${1:-10000}
mkdir -p gpurun_out
NG=$(nvidia-smi -L | wc -l)
for N in 2 4 8; do
  [ "$N" -gt "$NG" ] && break
  for V in one-shot two-shot; do
    timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 \
      --master-port $((29800 + N)) bench.py --gpus $N --steps $STEPS --warmup 200 --exchange $V 2>/dev/null | tail -1 > gpurun_out/variant_n${N}_$V.json
    python -c "
import json; d=json.load(open('gpurun_out/variant_n${N}_$V.json')); print('N=$N $V', d['value'], 'img/s', d['ms_per_step']*1e3, 'us/step  e2e', d['e2e']['value'], d['peer_timeouts'])"
  done
done
