#!/bin/bash
# 1-GPU A/B of library build variants
for tag in "$@"; do
  lib=distributed_neural_network_b200/_native/libdnnb200_$tag.so
  [ "$tag" = default ] && lib=distributed_neural_network_b200/_native/libdnnb200.so
  echo "== $tag"
  DNNB200_LIB=$PWD/$lib timeout 120 python bench/step_jitter.py 2>&1 | grep "^rank"
done
