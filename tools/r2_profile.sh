#!/bin/bash
# GPU box, round 2: the launch list of the bench command and ncu --set full captures of one ADMM iteration at an early,
# a middle and a late step of the first (cold-start) solve of a C2 fast_binless call. Usage: bash tools/r2_profile.sh <tag>
tag=${1:-r2}
mkdir -p gpurun_out
python bench.py --steps 2 --warmup 3 --no-cpu --no-extra > gpurun_out/${tag}_bench_plain.json 2> gpurun_out/${tag}_bench_plain.err || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 1200 --csv --log-file gpurun_out/${tag}_launches.csv \
    python tools/profile_target.py > gpurun_out/${tag}_launches.log 2>&1
for it in 2 9 17; do
  skip=$(( 4 * (it - 1) ))
  FB_NO_GRAPH=1 timeout 600 ncu --set full --clock-control none --import-source on -k "regex:k_beta|k_prox|k_ctl" --launch-skip $skip --launch-count 4 \
      -o gpurun_out/${tag}_admm_it$it -f python tools/profile_target.py > gpurun_out/${tag}_ncu_it$it.log 2>&1
done
ls -la gpurun_out | grep $tag
# IRLS leg passes (residual mode) of the second outer step
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_leg_pass --launch-skip 9 --launch-count 3 \
    -o gpurun_out/${tag}_irls -f python tools/profile_target.py > gpurun_out/${tag}_ncu_irls.log 2>&1
ls -la gpurun_out | grep ${tag}_irls
