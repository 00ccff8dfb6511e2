"""CPU: the restated third-party pieces of the oracle (Eigen / QuadProg++ / Boost stand-ins) against independent
implementations (scipy), plus the oracle's driver-level invariants and error behaviour."""
import ctypes as C
import os

import numpy as np
import pytest
from scipy.interpolate import BSpline
from scipy.optimize import minimize
from scipy.sparse import coo_matrix
from scipy.sparse.csgraph import connected_components

from binless_b200._abi import FastBinlessError
from binless_b200.synth import synth

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
c_dp = C.POINTER(C.c_double)


@pytest.fixture(scope="module")
def port_c(oracle_port):
    return C.CDLL(os.path.join(ROOT, "oracle", "_build", "libfbo_port.so"))


def test_spline_basis_matches_scipy(port_c):
    K = 50
    x = np.sort(np.random.default_rng(0).uniform(7.8, 16.3, size=300))
    xmin, xmax = x[0], x[-1]
    out = np.empty((x.size, K))
    port_c.fbo_test_spline(C.c_int(x.size), x.ctypes.data_as(c_dp), C.c_double(xmin), C.c_double(xmax), C.c_int(K),
                           out.ctypes.data_as(c_dp))
    dx = 1.01 * (xmax - xmin) / (K - 3)
    lower = xmin - dx * 0.01
    knots = np.linspace(lower - 3 * dx, lower + K * dx, K + 4)   # spline.cpp:19-26
    ref = np.stack([BSpline.basis_element(knots[j:j + 5], extrapolate=False)(x) for j in range(K)], axis=1)
    ref = np.nan_to_num(ref)
    assert np.allclose(out, ref, atol=1e-12)
    assert np.allclose(out.sum(axis=1), 1.0, atol=1e-12)         # partition of unity inside the support


@pytest.mark.parametrize("seed,free_first", [(0, 10), (1, 0), (2, 30), (3, 48)])
def test_qp_matches_scipy(port_c, seed, free_first):
    """min 1/2 x'Gx + g0'x s.t. x_i - x_{i+1} >= 0 for i >= free_first (Goldfarb-Idnani restatement vs SLSQP)."""
    n = 50
    rng = np.random.default_rng(seed)
    A = rng.normal(size=(n + 20, n))
    G = A.T @ A + 1e-3 * np.eye(n)
    g0 = rng.normal(size=n) * 5
    m = n - 1 - free_first
    CI = np.zeros((n, m))
    for c in range(m):
        CI[free_first + c, c] = 1; CI[free_first + c + 1, c] = -1
    x = np.empty(n)
    rc = port_c.fbo_test_qp(C.c_int(n), C.c_int(m), G.ctypes.data_as(c_dp), g0.ctypes.data_as(c_dp),
                            np.ascontiguousarray(CI).ctypes.data_as(c_dp), x.ctypes.data_as(c_dp))
    assert rc == 0
    assert np.all(CI.T @ x >= -1e-9)
    f = lambda v: 0.5 * v @ G @ v + g0 @ v
    res = minimize(f, x0=np.zeros(n), jac=lambda v: G @ v + g0, method="SLSQP",
                   constraints=[dict(type="ineq", fun=lambda v: CI.T @ v, jac=lambda v: CI.T)],
                   options=dict(maxiter=500, ftol=1e-14))
    assert f(x) <= f(res.x) + 1e-7 * (1 + abs(f(res.x)))
    assert np.allclose(x, res.x, atol=1e-4)


@pytest.mark.parametrize("n,tol,seed", [(2, 0.2, 0), (7, 0.2, 1), (40, 0.05, 2), (64, 0.5, 3)])
def test_patch_numbers_match_scipy_components(port_c, n, tol, seed):
    rng = np.random.default_rng(seed)
    nc = n * (n + 1) // 2
    val = np.round(rng.normal(size=nc), 1) * 0.3
    out = np.empty(nc, dtype=np.int32)
    port_c.fbo_test_patchnos(C.c_int(n), C.c_double(tol), val.ctypes.data_as(c_dp), out.ctypes.data_as(C.POINTER(C.c_int32)))
    rm = lambda i, j: i * n - i * (i - 1) // 2 + (j - i)
    ea, eb = [], []
    for i in range(n):
        for j in range(i, n):
            v = rm(i, j)
            if j + 1 < n and abs(val[v] - val[v + 1]) < tol: ea.append(v); eb.append(v + 1)
            if j >= i + 1 and abs(val[v] - val[rm(i + 1, j)]) < tol: ea.append(v); eb.append(rm(i + 1, j))
    g = coo_matrix((np.ones(len(ea)), (ea, eb)), shape=(nc, nc))
    ncomp, lab = connected_components(g, directed=False)
    # relabel scipy's components by lowest vertex, as Boost's DFS from vertex 0 numbers them
    first = {}
    relab = np.empty(nc, dtype=np.int32)
    for c in range(nc):
        relab[c] = first.setdefault(lab[c], len(first))
    assert out.max() + 1 == ncomp
    assert np.array_equal(out, relab)


def test_oracle_driver_invariants(oracle):
    n, D = 48, 2
    mat = synth(n, D, 5000, 4)
    with oracle.create(mat, n) as h:
        out = h.run_signal(0.49, [2.53, 4.51], [0.13, 0.10])
    nc = n * (n + 1) // 2
    d = out["diagnostics"]
    assert 1 <= d["nsteps"] <= 25 and d["n_solves"] == D * (d["nsteps"] - min(5, d["nsteps"] - 1)) or d["n_solves"] > 0
    ls = out["log_signal"]
    assert ls.min() >= 0.0                                   # shift_signal clamps at 0 (fast_binless.cpp:67-83)
    for k in range(D):
        assert ls[k * nc:(k + 1) * nc].min() == 0.0
    assert np.allclose(out["signal"], np.exp(np.sign(ls) * np.maximum(np.abs(ls) - np.repeat([0.13, 0.10], nc), 0)))
    assert np.allclose(out["binless"], out["decaymat"] * out["signal"])
    # bias estimate is centred with the nobs weights before the +-3 sd cap (fast_fitter_mean.hpp:55-79)
    assert abs(np.mean(out["biases"]["estimate"])) < 0.2
    assert out["decay"]["lambda_"] > 0
    # decay is non-increasing beyond free_decay (decreasing_constraint, gam.cpp:6-13)
    assert np.all(np.diff(out["decay"]["beta"][25:]) <= 1e-9)
    assert out["patchno"].min() == 0


def test_oracle_is_deterministic(oracle):
    mat = synth(40, 2, 5000, 6)
    with oracle.create(mat, 40) as h:
        a = h.run_signal(0.5, 2.5, 0.0)
    with oracle.create(mat, 40) as h:
        b = h.run_signal(0.5, 2.5, 0.0)
    for k in ("log_signal", "phihat", "weight", "binless"):
        assert np.array_equal(a[k], b[k])


def test_oracle_port_equals_reference_linked_build(oracle_port, oracle_ref):
    mat = synth(60, 2, 5000, 2)
    outs = []
    for lib in (oracle_port, oracle_ref):
        with lib.create(mat, 60) as h:
            o = h.run_signal(0.49, [2.53, 4.51], 0.1)
            o["diff"] = h.run_difference(2, 0.49, 3.38)
            outs.append(o)
    a, b = outs
    assert list(a["diagnostics"]["solve_iters"]) == list(b["diagnostics"]["solve_iters"])
    for k in ("log_signal", "phihat", "weight", "binless", "background", "patchno"):
        assert np.array_equal(a[k], b[k]), k
    for k in ("difference", "deltahat", "phi_ref", "patchno"):
        assert np.array_equal(a["diff"][k], b["diff"][k]), k


def test_oracle_input_errors(oracle):
    """Rcpp::stop messages of FastDataCore (FastData.ipp:20-25) and FastData<Difference> (FastData.hpp:90)."""
    mat = synth(10, 2, 5000, 1)
    bad = dict(mat); bad["name"] = mat["name"] + 1
    with pytest.raises(FastBinlessError, match="Input size should be"):
        oracle.create(bad, 10)
    with pytest.raises(FastBinlessError, match="Input size should be"):
        oracle.create(mat, 11)
    bad = dict(mat); bad["bin1"] = mat["bin1"].copy(); bad["bin1"][bad["bin1"] == 10] = 9
    with pytest.raises(FastBinlessError, match="bins should range"):
        oracle.create(bad, 10)
    with oracle.create(mat, 10) as h:
        with pytest.raises(FastBinlessError):
            h.run_difference(1, 0.5, 2.0)       # no fitted / restored state yet
        h.run_signal(0.5, 2.5)
        with pytest.raises(FastBinlessError, match="Ref is not in the allowed range"):
            h.run_difference(3, 0.5, 2.0)
        with pytest.raises(FastBinlessError, match="Provide only 1 dataset"):
            h.eval_cv(0.5, [1.0, 2.0])


def test_oracle_restore_roundtrip(oracle):
    """fast_binless_difference rebuilds the estimators from the returned list (fast_binless.cpp:244-260)."""
    mat = synth(30, 3, 5000, 8)
    with oracle.create(mat, 30) as h:
        out = h.run_signal(0.5, 2.5, 0.0)
        d1 = h.run_difference(2, 0.5, 3.0)
    with oracle.create(mat, 30) as h2:
        h2.restore(out["log_signal"], out["biases"], out["decay"], out["exposures"])
        d2 = h2.run_difference(2, 0.5, 3.0)
    for k in ("difference", "deltahat", "weight", "phi_ref", "patchno"):
        assert np.array_equal(d1[k], d2[k]), k
