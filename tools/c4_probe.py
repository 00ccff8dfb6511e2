"""Solver-level probe at chromosome scale (BASELINE config 4 is 25 000 bins): a fixed number of ADMM iterations of
fb_gfl_solve on a synthetic (y, w) triangle; prints wall time per iteration and the HBM roofline fraction.
`python tools/c4_probe.py [nbins] [maxsteps]`"""
import ctypes as C, sys, time
import numpy as np
sys.path.insert(0, '.')
from binless_b200._abi import FastBinlessLib, fb_gfl_state, _dp
n = int(sys.argv[1]) if len(sys.argv) > 1 else 25000
maxsteps = int(sys.argv[2]) if len(sys.argv) > 2 else 9
gpu = FastBinlessLib('binless_b200/libfastbinless_cuda.so', 'fb_')
nc = n * (n + 1) // 2
rng = np.random.default_rng(4)
# piecewise-constant signal + noise on the packed triangle (blocks of ~40 cells along rows), weights in (0, 3)
t0 = time.time()
y = np.repeat(rng.normal(0, 1, nc // 40 + 1), 40)[:nc] + rng.normal(0, 0.5, nc)
w = rng.uniform(0, 3, nc)
print(f"n={n} cells={nc} generated in {time.time()-t0:.1f}s", flush=True)
def run(ms):
    beta = np.zeros(nc)
    st = fb_gfl_state(beta=_dp(beta), z=None, u=None, alpha=5.0, steps=0)
    t = time.time()
    rc = gpu._gfl_solve(n, 0, _dp(y), _dp(w), 2.5, 1e-12, ms, 0, C.byref(st), None)
    gpu._check(rc, None)
    return time.time() - t, st.steps, beta
run(2)                         # warm-up: allocations
ta, sa, _ = run(2)             # 1 iteration: upload / download overhead
tb, sb, beta = run(maxsteps)
per = (tb - ta) / max(1, (sb - sa))
print(f"steps {sa}->{sb}: {ta:.3f}s -> {tb:.3f}s  => {per*1e3:.3f} ms per ADMM iteration, "
      f"{104*nc/per/1e9:.0f} GB/s algorithmic ({104*nc/per/6449.4e9:.3f} of HBM peak); beta mean {beta.mean():.6f}")
