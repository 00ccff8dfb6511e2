#!/bin/bash
# GPU box: does the number of host threads that move a staging chunk matter for the pageable (R-style) call? It does not:
# 4 / 8 / 16 threads give 46.7 / 41.9-43.1 / 43.8 ms at C2 and the same 8.9 ms of ingest: the host's memory bandwidth
# (validation reads 322 MB while 92 MB are staged) is the limit.
for t in 8 16 4 8; do
FB_STAGE_THREADS=$t FB_TIMING=1 python bench.py --steps 2 --warmup 2 --no-cpu --no-extra > gpurun_out/pg_$t.json 2> gpurun_out/pg_$t.err
python - <<PY
import json
d=json.loads([l for l in open("gpurun_out/pg_$t.json") if l.startswith("{")][-1])
print("threads $t pageable %.2f e2e %.2f" % (d["e2e_pageable"]["ms_per_step"], d["e2e"]["ms_per_step"]))
PY
grep "ingest" gpurun_out/pg_$t.err | tail -1
done
