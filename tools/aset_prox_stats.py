"""Design aid (CPU, test infrastructure only): statistics of the parallel active-set chain prox (oracle/aset_prox.c)
on the iterates of a real fused-lasso solve (hypothesis = previous ADMM iterate), per ADMM iteration: passes, points
and knots tested per cell, knots added / removed.

    python tools/aset_prox_stats.py [nbins] [iterations] [rows|cols|both]
"""
import ctypes as C
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))


class AsetStats(C.Structure):
    _fields_ = [(k, C.c_longlong) for k in ("passes", "point_tests", "knot_tests", "added", "removed", "chains", "fallbacks")] + [("max_passes", C.c_int)]


def aset_lib():
    lib = C.CDLL(os.path.join(ROOT, "oracle", "_build", "libaset_prox.so"))
    dp = C.POINTER(C.c_double)
    lib.aset_prox.argtypes = [C.c_int, dp, C.c_double, dp, dp, C.c_int, C.POINTER(AsetStats)]
    lib.aset_prox.restype = C.c_int
    return lib


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 600
    iters = int(sys.argv[2]) if len(sys.argv) > 2 else 18
    which = sys.argv[3] if len(sys.argv) > 3 else "both"
    from oracle.loader import oracle_lib
    from binless_b200._abi import gfl_solve
    from binless_b200.synth import synth
    lib = aset_lib()
    dp = C.POINTER(C.c_double)
    p = lambda a: a.ctypes.data_as(dp)
    ora = oracle_lib("auto")
    mat = synth(n, 1, 5000, 2)
    with ora.create(mat, n) as h:
        out = h.run_signal(0.5, [2.5], 0.0)
    y, w = out["phihat"], out["weight"]
    nc = n * (n + 1) // 2
    lam2 = 2.5
    state = None
    prev = dict(z=np.zeros(2 * nc - 2), u=np.zeros(2 * nc - 2), alpha=5.0)
    # chains in the reference's slot order: rows 0..n-2 (lengths n..2), then columns 1..n-1 (lengths 2..n)
    lens = [n - i for i in range(n - 1)] + [j + 1 for j in range(1, n)]
    sel = {"rows": range(0, n - 1), "cols": range(n - 1, 2 * n - 2), "both": range(0, 2 * n - 2)}[which]
    offs = np.concatenate([[0], np.cumsum(lens)])
    print(f"nbins {n}, {which} chains")
    print("iter | passes/chain max | point tests/cell knot tests/cell | added removed (per 1000 cells) | fallbacks  max|err|")
    for k in range(1, iters + 1):
        cur = gfl_solve(ora, n, y, w, lam2, 1e-300, 2, 0, state=state)    # exactly one more iteration
        lam = lam2 / (prev["alpha"] / 2.0)
        st = AsetStats()
        err = 0.0; cells = 0
        for ci in sel:
            L = lens[ci]; off = int(offs[ci])
            z_k, u_k = cur["z"][off:off + L], cur["u"][off:off + L]
            v = np.ascontiguousarray(z_k + u_k)         # = beta_k + u_{k-1}: the prox input of this iteration
            z_prev = np.ascontiguousarray(prev["z"][off:off + L])
            got = np.empty(L)
            r = lib.aset_prox(L, p(v), lam, p(z_prev), p(got), 64, C.byref(st))
            if r >= 0:
                err = max(err, float(np.abs(got - z_k).max()))
            cells += L
        print(f"{k:4d} | {st.passes / st.chains:6.2f} {st.max_passes:4d} | {st.point_tests / cells:8.2f} {st.knot_tests / cells:8.3f} | "
              f"{1000.0 * st.added / cells:7.2f} {1000.0 * st.removed / cells:7.2f} | {st.fallbacks:4d}  {err:.1e}")
        assert err < 1e-8
        state = dict(beta=cur["beta"], z=cur["z"], u=cur["u"], alpha=cur["alpha"])
        prev = cur


if __name__ == "__main__":
    main()
