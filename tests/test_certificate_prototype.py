"""CPU: the certificate-first chain prox planned for the next round (tools/certificate_prox_prototype.py, DESIGN.md
section 8) is EXACT: whatever knot hypothesis it starts from, it returns the minimiser the reference's tf_dp_weight
(gfl_tf.c:184-321, compiled from /root/reference into oracle/_ref) returns."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tools"))


def test_certified_prox_equals_tf_dp_weight_for_any_hypothesis(oracle_ref):
    from certificate_prox_prototype import certified_prox, tf_dp_reference
    ref = tf_dp_reference()
    rng = np.random.default_rng(1)
    searched = total = 0
    for trial in range(120):
        L = int(rng.integers(2, 150)); lam = float(rng.choice([0.2, 1.0, 4.0]))
        v = rng.normal(size=L) + np.repeat(rng.normal(0, 1.5, L // 6 + 1), 6)[:L]
        want = ref(v, 1.0, lam)
        kind = trial % 4
        if kind == 0:
            hyp = want.copy()                                   # the right knots: certificate passes, nothing searched
        elif kind == 1:
            hyp = ref(v + rng.normal(0, 0.05, L), 1.0, lam)     # a neighbouring problem (an ADMM step away)
        elif kind == 2:
            hyp = rng.normal(size=L).round()                    # nonsense
        else:
            hyp = np.zeros(L)                                   # no knots at all
        got, s, passes = certified_prox(v, lam, hyp)
        assert float(np.abs(got - want).max()) < 1e-10, (trial, kind)
        if kind == 0:
            assert s == 0 and passes == 1
        if kind == 1:
            searched += s; total += L
    assert searched < total                                      # a near hypothesis saves sequential work


def test_c_model_with_bit_masks_equals_tf_dp_weight(oracle_ref, oracle_port):
    """oracle/cert_prox.c: the same scheme with the data structures planned for the kernel (knot / side / bad-piece bit
    masks, per-cell certificate by bit scans, exact search between verified knots)."""
    import ctypes as C
    from certificate_prox_prototype import certified_prox, tf_dp_reference
    path = os.path.join(ROOT, "oracle", "_build", "libcert_prox.so")
    if not os.path.exists(path):
        from binless_b200 import build as b
        b.build_oracle()
    lib = C.CDLL(path)
    dp = C.POINTER(C.c_double)
    lib.cert_prox.argtypes = [C.c_int, dp, C.c_double, dp, dp, C.POINTER(C.c_int)]
    lib.cert_prox.restype = C.c_int
    ref = tf_dp_reference()
    rng = np.random.default_rng(2)
    p = lambda a: a.ctypes.data_as(dp)
    near_searched = near_total = 0
    for trial in range(300):
        L = int(rng.integers(1, 400)); lam = float(rng.choice([0.0, 0.2, 1.0, 4.0]))
        v = rng.normal(size=L) + np.repeat(rng.normal(0, 1.5, L // 6 + 1), 6)[:L]
        want = ref(v, 1.0, lam) if (L > 1 and lam > 0) else v.copy()
        kind = trial % 4
        hyp = [want.copy(), ref(v + rng.normal(0, 0.05, L), 1.0, max(lam, 1e-3)), rng.normal(size=L).round(), np.zeros(L)][kind]
        hyp = np.ascontiguousarray(hyp, dtype=np.float64)
        out = np.empty(L); passes = C.c_int(0)
        s = lib.cert_prox(L, p(np.ascontiguousarray(v)), lam, p(hyp), p(out), C.byref(passes))
        assert float(np.abs(out - want).max()) < 1e-10, (trial, kind, L, lam)
        if L > 1 and lam > 0:
            if kind == 0:
                assert s == 0 and passes.value == 1
            got_py, s_py, _ = certified_prox(v, lam, hyp)
            assert float(np.abs(got_py - out).max()) < 1e-10
            if kind == 1:
                near_searched += s; near_total += L
    assert near_searched < 0.8 * near_total
