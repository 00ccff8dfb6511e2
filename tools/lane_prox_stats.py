"""Design aid (CPU, test infrastructure only): statistics of the split-and-verify chain prox (oracle/lane_prox.c) on
the iterates of a real fused-lasso solve: piece lengths, how many hypothesised splits survive, lane steps per cell,
lock-step (warp) cost and the longest sequential range, per ADMM iteration.

    python tools/lane_prox_stats.py [nbins] [W] [iterations]
"""
import ctypes as C
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))


class LaneStats(C.Structure):
    _fields_ = [("steps_total", C.c_longlong), ("warp_steps", C.c_longlong), ("solves", C.c_longlong),
                ("splits", C.c_longlong), ("failed", C.c_longlong), ("cert_dirty", C.c_longlong), ("passes", C.c_int), ("max_range", C.c_int)]


def lane_lib():
    lib = C.CDLL(os.path.join(ROOT, "oracle", "_build", "liblane_prox.so"))
    dp = C.POINTER(C.c_double)
    lib.lane_prox.argtypes = [C.c_int, dp, C.c_double, dp, dp, C.c_int, C.POINTER(LaneStats)]
    lib.lane_prox.restype = C.c_int
    return lib


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 600
    W = int(sys.argv[2]) if len(sys.argv) > 2 else 32
    iters = int(sys.argv[3]) if len(sys.argv) > 3 else 18
    from binless_b200 import oracle_lib
    from binless_b200._abi import gfl_solve
    from binless_b200.synth import synth
    lib = lane_lib()
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
    print(f"nbins {n}, W {W}: per iteration over the ROW chains")
    print("iter  pieces/chain  mean_piece  p99_piece max_piece | splits failed(%) passes(max) lane_steps/cell warp_steps*32/cell max_range  dirty_share max|err|")
    for k in range(1, iters + 1):
        cur = gfl_solve(ora, n, y, w, lam2, 1e-300, 2, 0, state=state)    # exactly one more iteration
        lam = lam2 / (prev["alpha"] / 2.0)
        off = 0
        st = LaneStats()
        pmax = 0; err = 0.0; cells = 0
        plens = []
        for i in range(n - 1):
            L = n - i
            z_k, u_k = cur["z"][off:off + L], cur["u"][off:off + L]
            v = np.ascontiguousarray(z_k + u_k)         # = beta_k + u_{k-1}: the prox input of this iteration
            z_prev = np.ascontiguousarray(prev["z"][off:off + L])
            got = np.empty(L)
            ps = lib.lane_prox(L, p(v), lam, p(z_prev), p(got), W, C.byref(st))
            pmax = max(pmax, ps)
            err = max(err, float(np.abs(got - z_k).max()))
            kn = np.flatnonzero(np.diff(z_k) != 0)
            plens.append(np.diff(np.concatenate([[0], kn + 1, [L]])))
            cells += L
            off += L
        pl = np.concatenate(plens)
        print(f"{k:4d}  {pl.size / (n - 1):10.1f}  {pl.mean():9.1f}  {np.percentile(pl, 99):8.0f} {pl.max():8d} | "
              f"{st.splits:6d} {100.0 * st.failed / max(1, st.splits):6.1f}   {pmax:5d}     {st.steps_total / cells:8.2f}      "
              f"{32.0 * st.warp_steps / cells:8.2f}      {st.max_range:6d}   {st.cert_dirty / cells:6.3f}   {err:.1e}")
        assert err < 1e-8
        state = dict(beta=cur["beta"], z=cur["z"], u=cur["u"], alpha=cur["alpha"])
        prev = cur


if __name__ == "__main__":
    main()
