"""Development aid: time the ADMM iteration of a library build (`python tools/variant_time.py lib.so [nbins]`):
fast_binless on the bench workload's first dataset gives a realistic (phihat, weight); then the difference of two
fixed-iteration solves (19 vs 3 steps, cold start) is 16 iterations of k_beta + k_prox + k_ctl."""
import sys, time, numpy as np
sys.path.insert(0, '.')
from binless_b200._abi import FastBinlessLib, gfl_solve
from binless_b200.synth import synth
path = sys.argv[1]; n = int(sys.argv[2]) if len(sys.argv) > 2 else 2400
gpu = FastBinlessLib(path, 'fb_')
mat = synth(n, 1, 5000, 2)
with gpu.create({k: (None if k in ("name", "bin1", "bin2") else v) for k, v in mat.items()}, n) as h:
    out = h.run_signal(0.5, [2.5], 0.0, columns=True)
y, w = out['phihat'], out['weight']
def t(ms):
    best = 1e9
    for _ in range(3):
        t0 = time.perf_counter(); r = gfl_solve(gpu, n, y, w, 2.5, 1e-12, ms, 0); best = min(best, time.perf_counter() - t0)
    return best, r
ta, _ = t(3); tb, r = t(19)
print(f"{path}: {(tb-ta)/16*1e6:.1f} us per ADMM iteration (n={n}); checksum {r['beta'].sum():.12f}")
d = out["diagnostics"]
print(f"   run_signal on the same data: {d['ms_solver']*1e3/max(1,d['admm_iters']):.1f} us per ADMM iteration (device events), iterations {list(d['solve_iters'])}")
