"""One fast_binless step on the bench workload (C2), for ncu: `python tools/profile_target.py [nbins] [ndatasets]`.
Results stay on the device (same call as bench.py's `value` leg)."""
import sys
sys.path.insert(0, '.')
from binless_b200 import cuda_lib
from binless_b200.synth import synth
n = int(sys.argv[1]) if len(sys.argv) > 1 else 2400
D = int(sys.argv[2]) if len(sys.argv) > 2 else 4
mat = synth(n, D, 5000, 2)
slim = {k: (None if k in ("name", "bin1", "bin2") else v) for k, v in mat.items()}
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 1
with cuda_lib().create(slim, n) as h:
    for _ in range(reps):
        o = h.run_signal(0.5, 2.5, 0.0, columns="none")
    d = o["diagnostics"]
    print("outer steps", d["nsteps"], "iters", list(d["solve_iters"]), "ms_total", d["ms_total"], "ms_solver", d["ms_solver"],
          "launches", h.kernel_launches())
