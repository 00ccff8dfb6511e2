#!/bin/bash
# GPU box: C2 bench line (no CPU leg, no extra legs) for every variant library given: bash tools/variants_bench.sh tag name...
tag=$1; shift
mkdir -p gpurun_out
for v in "$@"; do
  lib=binless_b200/_variants/lib_$v.so
  [ "$v" = "product" ] && lib=binless_b200/libfastbinless_cuda.so
  FB_LIB=$lib timeout 300 python bench.py --steps 6 --warmup 3 --no-cpu --no-extra > gpurun_out/${tag}_$v.json 2> gpurun_out/${tag}_$v.err
  python - <<PY
import json
try:
    d=json.load(open('gpurun_out/${tag}_$v.json'))
    print('$v', 'ms_per_step %.3f'%d['ms_per_step'], 'us/iter %.1f'%d['roofline']['us_per_launch'], 'frac %.4f'%d['roofline']['frac'], d['config']['admm_iters_per_solve'], 'e2e %.2f'%d['e2e']['ms_per_step'])
except Exception as e: print('$v failed', e)
PY
done
