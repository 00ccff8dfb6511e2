"""A/B of the two chain-prox kernels (parallel active set k_prox_aset vs taut-string team kernel k_prox) on a realistic
(phihat, weight) pair: ms per ADMM iteration of one cold-start solve at several sizes.
    python tools/prox_ab.py [nbins ...]"""
import os, sys, time, numpy as np
sys.path.insert(0, '.')
from binless_b200._abi import FastBinlessLib, gfl_solve
from binless_b200.synth import synth
gpu = FastBinlessLib('binless_b200/libfastbinless_cuda.so', 'fb_')
for n in [int(a) for a in sys.argv[1:]] or [1200, 2400, 3600]:
    mat = synth(n, 1, 5000, 2)
    with gpu.create({k: (None if k in ("name", "bin1", "bin2") else v) for k, v in mat.items()}, n) as h:
        out = h.run_signal(0.5, [2.5], 0.0, columns=True)
    y, w = out['phihat'], out['weight']
    res = {}
    for name, env in (("aset", {"FB_PROX_ASET": "1"}), ("team", {"FB_PROX_ASET": "0"})):
        os.environ.update(env)
        gfl_solve(gpu, n, y, w, 2.5, 1e-12, 3, 0)
        ms = {}
        for steps in (3, 43):
            best = 1e9
            for rep in range(3):
                t0 = time.perf_counter()
                gfl_solve(gpu, n, y, w, 2.5, 1e-12, steps, 0)
                best = min(best, time.perf_counter() - t0)
            ms[steps] = best * 1e3
        res[name] = (ms[43] - ms[3]) / 40.0
    print(f"nbins {n}: per ADMM iteration (iterations 3..42 of a cold start)  aset {res['aset']*1e3:8.1f} us   team {res['team']*1e3:8.1f} us", flush=True)
