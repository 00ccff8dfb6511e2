#!/bin/bash
# GPU box: long-chain solve timing (tools/c4_target.py) for every variant library given: bash tools/c4_variants.sh "sizes" name...
sizes=$1; shift
mkdir -p gpurun_out
for v in "$@"; do
  lib=binless_b200/_variants/lib_$v.so
  [ "$v" = "product" ] && lib=binless_b200/libfastbinless_cuda.so
  for n in $sizes; do
    echo "== $v $n"
    FB_LIB=$lib timeout 300 python tools/c4_target.py $n 2 12 2>&1 | tail -2
  done
done
