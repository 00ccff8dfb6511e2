#!/bin/bash
# GPU box: phase statistics + ncu captures of k_prox_aset at an early, a middle and a late ADMM iteration of one solve
# (tools/prox_target.py: 2400 bins; its fb_run_signal runs 17 iterations first). Usage: bash tools/aset_profile.sh <tag> [ncu]
tag=${1:-aset}
mkdir -p gpurun_out
FB_NO_GRAPH=1 FB_ASET_STATS=1 timeout 300 python tools/prox_target.py 2400 19 2> gpurun_out/${tag}_stats.log | tail -2
timeout 300 python tools/prox_target.py 2400 19 | tail -1
timeout 200 python bench.py --steps 5 --warmup 3 --no-cpu --no-extra > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err; python -c "
import json;d=json.load(open('gpurun_out/${tag}_bench.json'));print('bench: ms_per_step',d['ms_per_step'],'us/iter',d['roofline']['us_per_launch'],'frac',d['roofline']['frac'],d['config']['admm_iters_per_solve'])"
if [ "$2" = "ncu" ]; then
for skip in 18 25 33; do
  FB_NO_GRAPH=1 timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_prox --launch-skip $skip --launch-count 1 \
      -o gpurun_out/${tag}_it$skip -f python tools/prox_target.py 2400 19 > gpurun_out/${tag}_ncu_it$skip.log 2>&1
done
fi
ls -la gpurun_out/ | grep $tag
