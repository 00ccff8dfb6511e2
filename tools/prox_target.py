"""ncu target: a fixed number of ADMM iterations of one fused-lasso solve (k_beta + k_prox + k_ctl) on a realistic
(phihat, weight) pair. `python tools/prox_target.py [nbins] [maxsteps]`"""
import sys, numpy as np
sys.path.insert(0, '.')
from binless_b200._abi import FastBinlessLib, gfl_solve
from binless_b200.synth import synth
n = int(sys.argv[1]) if len(sys.argv) > 1 else 2400
maxsteps = int(sys.argv[2]) if len(sys.argv) > 2 else 12
gpu = FastBinlessLib('binless_b200/libfastbinless_cuda.so', 'fb_')
mat = synth(n, 1, 5000, 2)
with gpu.create({k: (None if k in ("name", "bin1", "bin2") else v) for k, v in mat.items()}, n) as h:
    out = h.run_signal(0.5, [2.5], 0.0, columns=True)
r = gfl_solve(gpu, n, out['phihat'], out['weight'], 2.5, 1e-12, maxsteps, 0)
print("steps", r["steps"], "ms", r.get("ms"))
