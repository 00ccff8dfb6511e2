"""Development aid: per-chain cycle breakdown of the last k_prox launch of a solve (FB_PROX_TRACE)."""
import os, sys, numpy as np
sys.path.insert(0, '.')
os.environ["FB_PROX_TRACE"] = "gpurun_out/prox_trace.bin"
from binless_b200._abi import FastBinlessLib, gfl_solve
from binless_b200.synth import synth
n = int(sys.argv[1]) if len(sys.argv) > 1 else 2400
maxsteps = int(sys.argv[2]) if len(sys.argv) > 2 else 12
gpu = FastBinlessLib('binless_b200/libfastbinless_cuda.so', 'fb_')
rng = np.random.default_rng(0)
mat = synth(n, 1, 5000, 2)
with gpu.create({k: (None if k in ("name", "bin1", "bin2") else v) for k, v in mat.items()}, n) as h:
    out = h.run_signal(0.5, [2.5], 0.0, columns=True)
r = gfl_solve(gpu, n, out['phihat'], out['weight'], 2.5, 1e-12, maxsteps, 0)
t = np.fromfile("gpurun_out/prox_trace.bin", dtype=np.int64).reshape(-1, 4)
load, taut, outp = t[:, 0], t[:, 1], t[:, 2]
smid = (t[:, 3] >> 48); nr = t[:, 3] & ((1 << 48) - 1)
tot = load + taut + outp
print("chains", len(t), "sum cycles load/taut/out (M):", load.sum() / 1e6, taut.sum() / 1e6, outp.sum() / 1e6)
for q in (0, 1, 2, 3, 100, 101, 1000, 1001, 2400, 2401, 4000, 4797):
    L = n - q // 2
    print(f"q={q} L={L} sm={smid[q]} load={load[q]} taut={taut[q]} out={outp[q]}  rounds={nr[q]} cycles/round={taut[q]/max(nr[q],1):.0f}")
print("max chain total", tot.max(), "at q", tot.argmax())
per_sm = np.bincount(smid, weights=tot)
print("per-SM summed warp-cycles: min/mean/max", per_sm.min(), per_sm.mean(), per_sm.max(), " /11 warps ->", per_sm.max() / 11)
print("total rounds", nr.sum(), "mean cycles/round", taut.sum() / nr.sum())
