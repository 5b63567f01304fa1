#!/bin/bash
# A/B of library build variants (DNN_BUILD_TAG builds): step time at 1 and N GPUs per variant
N=${1:-2}
shift
for tag in "$@"; do
  lib=distributed_neural_network_b200/_native/libdnnb200_$tag.so
  [ "$tag" = default ] && lib=distributed_neural_network_b200/_native/libdnnb200.so
  echo "== $tag"
  DNNB200_LIB=$PWD/$lib timeout 120 python bench/step_jitter.py 2>&1 | grep "^rank"
  DNNB200_LIB=$PWD/$lib timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench/step_jitter.py 2>&1 | grep "^rank"
done
