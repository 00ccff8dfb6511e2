#!/bin/bash
# GPU box: C2 bench line for several switch points between the team kernel and the active set (FB_ASET_FROM)
mkdir -p gpurun_out
for f in 1 5 9 13 17 25; do
  FB_ASET_FROM=$f timeout 300 python bench.py --steps 6 --warmup 3 --no-cpu --no-extra > gpurun_out/asetfrom_$f.json 2>/dev/null
  python - <<PY
import json
d=json.loads([l for l in open('gpurun_out/asetfrom_$f.json') if l.startswith('{')][-1])
print('aset_from $f', 'ms_per_step %.3f'%d['ms_per_step'], 'us/iter %.1f'%d['roofline']['us_per_launch'], d['config']['admm_iters_per_solve'])
PY
done
FB_PROX_ASET=0 timeout 300 python bench.py --steps 6 --warmup 3 --no-cpu --no-extra > gpurun_out/asetfrom_team.json 2>/dev/null
python - <<PY
import json
d=json.loads([l for l in open('gpurun_out/asetfrom_team.json') if l.startswith('{')][-1])
print('team only', 'ms_per_step %.3f'%d['ms_per_step'], 'us/iter %.1f'%d['roofline']['us_per_launch'], d['config']['admm_iters_per_solve'])
PY
