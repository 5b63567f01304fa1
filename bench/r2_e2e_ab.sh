#!/bin/sh
NG=$(nvidia-smi -L | wc -l)
DNN_HOST=1 timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $NG --master-addr 127.0.0.1 --master-port 29511 bench/step_jitter.py 2>&1 | grep "^rank 0/" | cut -c1-700
