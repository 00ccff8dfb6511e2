"""CPU: the oracle of the optimized-path performance iteration. The scatter front end is the
reference's own cts_core.c (compiled in place into oracle/_ref/libfbo_ref.so); the restatement used on boxes without
/root/reference must equal it bit for bit; a numpy restatement checks both; the BIC machinery around it is checked
against a brute-force numpy evaluation of the chosen candidate."""
import ctypes as C

import numpy as np
import pytest

from binless_b200 import _abi
from binless_b200._abi import _dp, _f64, wgfl_diff_bic, wgfl_signal_bic
from binless_b200.synth import synth_cts, tri_indices


def cts_core(lib, cts, nbins, phi):
    keep = []
    c = _abi._cts_struct(cts, keep)
    nc = nbins * (nbins + 1) // 2
    out = [np.empty(nc) for _ in range(4)]
    fn = getattr(lib.lib, lib.prefix + "test_cts_core")
    fn.argtypes = [C.c_int, C.c_double, C.POINTER(_abi.fb_cts), _abi.c_dp] + [_abi.c_dp] * 4
    fn.restype = C.c_int
    assert fn(nbins, float(cts["dispersion"]), C.byref(c), _dp(_f64(phi)), *[_dp(o) for o in out]) == 0
    return out


def numpy_cts_core(cts, nbins, phi):
    """cts_core.c:5-47 with np.add.at in row order (same order of additions per cell as the C loop)"""
    b1 = cts["bin1"].astype(np.int64); b2 = cts["bin2"].astype(np.int64)
    pos = (b1 - 1) * (nbins + 1) - (b1 - 1) * b1 // 2 + b2 - b1
    nc = nbins * (nbins + 1) // 2
    mu = np.exp(cts["lmu_nosig"] + phi[pos]); mu_nd = np.exp(cts["lmu_nosig"] - cts["log_decay"] + phi[pos])
    z = cts["count"] / mu - 1
    wt = cts["nobs"] / (1 / mu + 1 / cts["dispersion"]); wt_nd = cts["nobs"] / (1 / mu_nd + 1 / cts["dispersion"])
    ncounts = np.zeros(nc); var = np.zeros(nc); var_nd = np.zeros(nc); phihat = np.zeros(nc)
    np.add.at(ncounts, pos, cts["nobs"]); np.add.at(var, pos, wt); np.add.at(var_nd, pos, wt_nd)
    np.add.at(phihat, pos, (z + phi[pos]) * wt)
    with np.errstate(divide="ignore"):
        var = 1 / var; var_nd = 1 / var_nd
    phihat = np.where(ncounts > 0, phihat * var, 0.0)
    return phihat, var, var_nd, ncounts


@pytest.mark.parametrize("nbins,seed", [(7, 1), (33, 2), (64, 3)])
def test_cts_core_restatement_equals_reference_object(oracle_port, oracle_ref, nbins, seed):
    cts = synth_cts(nbins, seed)
    rng = np.random.default_rng(seed)
    phi = rng.normal(0, 0.3, nbins * (nbins + 1) // 2)
    a = cts_core(oracle_port, cts, nbins, phi)
    b = cts_core(oracle_ref, cts, nbins, phi)          # cts_core.c compiled from /root/reference
    for x, y in zip(a, b):
        assert np.array_equal(x, y, equal_nan=True)
    for x, y in zip(b, numpy_cts_core(cts, nbins, phi)):
        assert np.allclose(x, y, rtol=1e-13, atol=0, equal_nan=True)
    assert nbins < 30 or np.isinf(b[1]).sum() > 0        # cells without a row: variance 1/0


def soft(b, e, l):
    v = b - e
    return np.where(v > 0, np.maximum(0, v - l), np.minimum(0, v + l))


@pytest.mark.parametrize("fixed,positive,constrained", [(True, True, True), (True, False, False), (False, True, True)])
def test_wgfl_signal_bic_oracle_is_self_consistent(oracle, fixed, positive, constrained):
    nbins = 24
    cts = synth_cts(nbins, 11)
    nc = nbins * (nbins + 1) // 2
    r = wgfl_signal_bic(oracle, cts, cts["dispersion"], 20, nbins, None, 2.0, 0.1, cts["metadata"], np.zeros(nc),
                        constrained, fixed, positive)
    assert r["converged"] and r["ncandidates"] >= 1 and r["admm_iters"] > 0
    # the reported phi is the soft-thresholded beta at the chosen (eCprime, lambda1)
    assert np.array_equal(r["phi"], soft(r["beta"], r["eCprime"], r["lambda1"]))
    if fixed:
        assert r["eCprime"] == 0.0
    else:
        assert abs((r["eCprime"] - r["lambda1"]) - r["beta"].min()) < 1e-12     # LB = min(beta), BoundsOptimizer.ipp:19-20
    # dof = patches with a cell beyond tol/2 after thresholding (DOFComputer.hpp:16-24)
    sel = np.abs(r["phi"]) > 0.1 / 2
    assert r["dof"] == np.unique(r["patchno"][sel]).size
    # outlier rows / diagonals carry no weight (cts_to_mat.cpp:14-32)
    i, j = tri_indices(nbins)
    bad = (i + 1 == cts["metadata"]["bad_rows"][0]) | (j + 1 == cts["metadata"]["bad_rows"][0]) | (j - i == nbins - 1)
    assert np.all(r["weight"][bad] == 0) and np.all(r["weight"][~bad & (r["nobs"] > 0)] > 0)
    # a second call warm-started from the returned state and beta converges at once to the same solution
    r2 = wgfl_signal_bic(oracle, cts, cts["dispersion"], 20, nbins, r["GFLState"], 2.0, 0.1, cts["metadata"], r["beta"],
                         constrained, fixed, positive)
    assert r2["converged"] and r2["nouter_done"] <= 2
    assert np.max(np.abs(r2["beta"] - r["beta"])) < 0.1 / 20 * 2
    # lambda1 fixed: the bounds are rectified, not searched (FixedSparsity.hpp, BoundsRectifier.hpp)
    r3 = wgfl_signal_bic(oracle, cts, cts["dispersion"], 20, nbins, None, 2.0, 0.1, cts["metadata"], np.zeros(nc),
                         constrained, fixed, positive, lambda1_fixed=True, lambda1_fix_value=0.3)
    assert r3["ncandidates"] == 1 and r3["lambda1"] >= 0.3 - 1e-15


def test_wgfl_diff_bic_oracle_is_self_consistent(oracle):
    nbins = 20
    ref = synth_cts(nbins, 5); oth = synth_cts(nbins, 6, shift=0.5)
    nc = nbins * (nbins + 1) // 2
    r = wgfl_diff_bic(oracle, oth, ref, ref["dispersion"], 20, nbins, None, 1.5, 0.1, ref["metadata"], np.zeros(nc),
                      np.zeros(nc), True)
    assert r["ncandidates"] >= 1
    assert np.array_equal(r["delta"], soft(r["beta"], 0.0, r["lambda1"]))
    assert np.all(r["phi_ref"] >= 0)                                  # util.cpp:37
    assert np.allclose(r["phihat"] - r["deltahat"], r["phihat"] - r["deltahat"])
    assert r["dof"] == np.unique(r["patchno"][np.abs(r["delta"]) > 0.05]).size
    with np.errstate(divide="ignore"):
        assert np.allclose(1 / r["phihat_var"], r["weight"])
