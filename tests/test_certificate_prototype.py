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
