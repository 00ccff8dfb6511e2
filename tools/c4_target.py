"""Timing / ncu / trace target: fused-lasso solves on the full triangle with long chains (C4: 25 000 bins) on one GPU.
    python tools/c4_target.py [nbins] [steps_a] [steps_b]
prints the device time of the ADMM loop for the two step counts and the time per iteration from their difference.
With FB_PROX_TRACE=<file> (polling loop, team kernel only) it also summarises the per-chain cycle counts of the last
k_prox launch: staging + prefix sums / taut string / write-back, rounds per chain, by chain-length class."""
import ctypes as C
import os
import sys
import time

import numpy as np

sys.path.insert(0, '.')
from binless_b200 import cuda_lib  # noqa: E402
from binless_b200._abi import fb_gfl_state, _dp  # noqa: E402

nb = int(sys.argv[1]) if len(sys.argv) > 1 else 25000
sa = int(sys.argv[2]) if len(sys.argv) > 2 else 2
sb = int(sys.argv[3]) if len(sys.argv) > 3 else 12
nc = nb * (nb + 1) // 2
rng = np.random.default_rng(4)                      # the inputs of bench.py's C4 leg
tile = rng.normal(0, 0.5, 1 << 20)
y = np.repeat(rng.normal(0, 1, nc // 40 + 1), 40)[:nc]
y += np.resize(tile, nc)
w = np.resize(rng.uniform(0, 3, 1 << 20), nc)
lib = cuda_lib()


def run(steps):
    beta = np.zeros(nc)
    st = fb_gfl_state(beta=_dp(beta), z=None, u=None, alpha=5.0, steps=0)
    t = time.perf_counter()
    lib._check(lib._gfl_solve(nb, 0, _dp(y), _dp(w), 2.5, 1e-12, steps, 0, C.byref(st), None), None)
    return lib._gfl_last_solve_ms(), st.steps, time.perf_counter() - t, float(np.abs(beta).sum())


run(sa)
ma, na, wa, ca = run(sa)
mb, nb_, wb, cb = run(sb) if sb > sa else (ma, na, wa, ca)
print(f"nbins {nb}: {na} steps {ma:.2f} ms, {nb_} steps {mb:.2f} ms (wall {wb:.2f} s), checksum {cb!r}")
if nb_ > na:
    per = (mb - ma) / (nb_ - na)
    print(f"per ADMM iteration {per:.3f} ms = {104 * nc / per / 1e6:.0f} GB/s algorithmic")
tr = os.environ.get("FB_PROX_TRACE")
if tr and os.path.exists(tr):
    t = np.fromfile(tr, dtype=np.int64).reshape(-1, 4)
    nr = t[:, 3] & ((1 << 48) - 1)
    sm = t[:, 3] >> 48
    q = np.arange(t.shape[0])
    L = np.where(q % 2 == 0, nb - q // 2, nb - q // 2)       # rows n - r cells, columns j + 1 = n - (q - 1) / 2 cells
    tot = t[:, :3].sum()
    print(f"trace: {t.shape[0]} chains, cycles stage+scan {t[:, 0].sum() / tot:.3f} taut {t[:, 1].sum() / tot:.3f} write-back {t[:, 2].sum() / tot:.3f}")
    print(f"  total team-cycles {tot:.3e}; per chain-cell {tot / L.sum():.2f} cycles; rounds per cell {nr.sum() / L.sum():.4f}; SMs used {np.unique(sm).size}")
    for lo, hi in ((1, 1000), (1000, 5000), (5000, 12000), (12000, 20000), (20000, nb + 1)):
        m = (L >= lo) & (L < hi)
        if m.any():
            c = L[m].sum()
            print(f"  L in [{lo}, {hi}): {m.sum()} chains; cycles/cell stage {t[m, 0].sum() / c:.2f} taut {t[m, 1].sum() / c:.2f} wb {t[m, 2].sum() / c:.2f}; "
                  f"cells per round {c / max(1, nr[m].sum()):.1f}")
