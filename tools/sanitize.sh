#!/bin/bash
# GPU box: compute-sanitizer memcheck + racecheck over small parity cases (both chain-prox kernels, shared-memory and
# global-scratch modes, trapeze, per-cell binner, patch labels, f2). Usage: bash tools/sanitize.sh <tag>
tag=${1:-san}
mkdir -p gpurun_out
run() {  # name, tool, env..., -- pytest args
  name=$1; tool=$2; shift 2
  envs=()
  while [ "$1" != "--" ]; do envs+=("$1"); shift; done; shift
  env "${envs[@]}" timeout 900 compute-sanitizer --tool $tool --print-limit 20 --error-exitcode 9 \
      python -m pytest -m gpu -x -q "$@" > gpurun_out/${tag}_${tool}_${name}.log 2>&1
  echo "$name $tool rc=$? $(grep -E 'ERROR SUMMARY|RACECHECK SUMMARY' gpurun_out/${tag}_${tool}_${name}.log | tail -1)"
}
SIG='test_fast_binless_matches_oracle and (33-2-3 or 5-2-3 or 64-2-6 or 70-2-8)'
GFL='test_gfl_solve_matches_oracle and (17-1.3 or 100-1.3 or 120-2.0)'
run parity memcheck -- tests/test_gpu_parity.py -k "($SIG) or csv_out or per_cell_binner_equals or cv_score"
run gfl memcheck -- tests/test_gpu_parity.py -k "$GFL"
run perf memcheck -- tests/test_gpu_perf_iteration.py -k "12 or rejects"
run parity racecheck -- tests/test_gpu_parity.py -k "($SIG) or cv_score"
run gfl racecheck -- tests/test_gpu_parity.py -k "$GFL"
run gs racecheck FB_PROX_ASET=0 FB_PROX_GS=1 -- tests/test_gpu_parity.py -k "test_fast_binless_matches_oracle and (33-2-3 or 64-2-6)"
run perf racecheck -- tests/test_gpu_perf_iteration.py -k "12"
