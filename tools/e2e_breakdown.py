"""Where the end-to-end call of bench.py's e2e leg spends its time (C2, pinned host buffers): wall clock of fb_create,
fb_run_signal and fb_destroy as Python sees them, next to the library's own phase lines (FB_TIMING=1).
    FB_TIMING=1 python tools/e2e_breakdown.py [reps]"""
import sys
import time

import numpy as np
import torch

sys.path.insert(0, '.')
from bench import CONFIGS, ALPHA, LAM2, LAM1, pinned_like, pinned_empty  # noqa: E402
from binless_b200 import cuda_lib  # noqa: E402
from binless_b200.synth import synth  # noqa: E402

reps = int(sys.argv[1]) if len(sys.argv) > 1 else 5
lib = cuda_lib()
cfg = CONFIGS["C2"]
n, D = cfg["nbins"], cfg["ndatasets"]
mat = synth(n, D, cfg["res"], cfg["seed"])
keep, pin = [], {}
for k in ("pos1", "pos2", "observed", "nobs", "distance"):
    pin[k], t = pinned_like(mat[k]); keep.append(t)
pin.update(name=None, bin1=None, bin2=None)
N = pin["distance"].size
out = {}
for k in ("background", "biasmat", "decaymat", "signal", "binless", "phihat", "weight", "log_signal"):
    out[k], t = pinned_empty(N, torch.float64); keep.append(t)
out["patchno"], t = pinned_empty(N, torch.int32); keep.append(t)
acc = np.zeros(4)
for r in range(reps + 2):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    h = lib.create(pin, n)
    t1 = time.perf_counter()
    o = h.run_signal(ALPHA, LAM2, LAM1, columns=True, buffers=out)
    t2 = time.perf_counter()
    h.close()
    t3 = time.perf_counter()
    if r >= 2:
        acc += [t1 - t0, t2 - t1, t3 - t2, t3 - t0]
print("e2e breakdown (ms): create %.2f  run_signal %.2f  destroy %.2f  total %.2f" % tuple(acc / reps * 1e3))
