"""CPU: the parallel active-set chain prox of k_prox_aset (scalar model oracle/aset_prox.c, same data structures as the
kernel) is EXACT: whatever knot hypothesis it starts from, it returns the minimiser the reference's tf_dp_weight
(gfl_tf.c:184-321, compiled from /root/reference into oracle/_ref) returns, or reports that it did not settle."""
import ctypes as C
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tools"))


def _aset_lib():
    from aset_prox_stats import aset_lib
    path = os.path.join(ROOT, "oracle", "_build", "libaset_prox.so")
    if not os.path.exists(path):
        from binless_b200 import build as b
        b.build_oracle()
    return aset_lib()


@pytest.mark.parametrize("seed", [0, 1, 2])
def test_aset_model_equals_tf_dp_weight_for_any_hypothesis(oracle_ref, seed):
    from aset_prox_stats import AsetStats
    from certificate_prox_prototype import tf_dp_reference
    lib = _aset_lib()
    ref = tf_dp_reference()
    rng = np.random.default_rng(11 + seed)
    dp = C.POINTER(C.c_double)
    p = lambda a: a.ctypes.data_as(dp)
    st = AsetStats()
    for trial in range(400):
        L = int(rng.integers(1, 700)); lam = float(rng.choice([0.0, 0.2, 1.0, 4.0]))
        v = rng.normal(size=L) + np.repeat(rng.normal(0, 1.5, L // 6 + 1), 6)[:L]
        if trial % 5 == 4:
            v = np.round(v)                       # exact ties between the polygon and the tube
        want = ref(v, 1.0, lam) if (L > 1 and lam > 0) else v.copy()
        kind = trial % 4
        hyp = [want.copy(), ref(v + rng.normal(0, 0.05, L), 1.0, max(lam, 1e-3)), rng.normal(size=L).round(), np.zeros(L)][kind]
        hyp = np.ascontiguousarray(hyp, dtype=np.float64)
        out = np.empty(L)
        passes = lib.aset_prox(L, p(np.ascontiguousarray(v)), lam, p(hyp), p(out), 64, C.byref(st))
        if passes < 0:
            continue                               # reported: the kernel falls back to the sequential taut string
        assert float(np.abs(out - want).max()) < 1e-10, (trial, kind, L, lam)
        if kind == 0 and L > 1 and lam > 0 and trial % 5 != 4:
            assert passes == 1                     # the right knots: the first pass only confirms
    assert st.fallbacks <= 2


def test_aset_model_matches_on_real_iterates(oracle_ref):
    """The iterates of a real solve (hypothesis = previous ADMM iterate), 96 bins, rows and columns."""
    from aset_prox_stats import AsetStats
    from binless_b200._abi import gfl_solve
    from binless_b200.synth import synth
    lib = _aset_lib()
    ora = oracle_ref
    n = 96
    mat = synth(n, 1, 5000, 2)
    with ora.create(mat, n) as h:
        out = h.run_signal(0.5, [2.5], 0.0)
    y, w = out["phihat"], out["weight"]
    nc = n * (n + 1) // 2
    dp = C.POINTER(C.c_double)
    p = lambda a: a.ctypes.data_as(dp)
    lens = [n - i for i in range(n - 1)] + [j + 1 for j in range(1, n)]
    offs = np.concatenate([[0], np.cumsum(lens)])
    state = None
    prev = dict(z=np.zeros(2 * nc - 2), alpha=5.0)
    st = AsetStats()
    for k in range(1, 9):
        cur = gfl_solve(ora, n, y, w, 2.5, 1e-300, 2, 0, state=state)
        lam = 2.5 / (prev["alpha"] / 2.0)
        for ci, L in enumerate(lens):
            off = int(offs[ci])
            z_k, u_k = cur["z"][off:off + L], cur["u"][off:off + L]
            v = np.ascontiguousarray(z_k + u_k)
            got = np.empty(L)
            r = lib.aset_prox(L, p(v), lam, p(np.ascontiguousarray(prev["z"][off:off + L])), p(got), 64, C.byref(st))
            assert r >= 0
            assert float(np.abs(got - z_k).max()) < 1e-9
        state = dict(beta=cur["beta"], z=cur["z"], u=cur["u"], alpha=cur["alpha"])
        prev = cur
    assert st.fallbacks == 0 and st.max_passes <= 16
