"""CPU: pins the oracle's GFL core (oracle/gfl_port.c) against the reference's own C sources compiled in place
(oracle/_ref/libgfl_ref.so = /root/reference/binless/src/gfl_graph_fl.c + gfl_tf.c) and against first principles."""
import ctypes as C
import os

import numpy as np
import pytest

from binless_b200._abi import gfl_solve

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
c_dp = C.POINTER(C.c_double)


def _dp(a):
    return a.ctypes.data_as(c_dp)


@pytest.fixture(scope="module")
def port_c(oracle_port):
    lib = C.CDLL(os.path.join(ROOT, "oracle", "_build", "libfbo_port.so"))
    return lib


@pytest.fixture(scope="module")
def ref_c(oracle_port):
    path = os.path.join(ROOT, "oracle", "_ref", "libgfl_ref.so")
    if not os.path.exists(path):
        pytest.skip("oracle/_ref not built")
    return C.CDLL(path)


def tv1d_objective(x, y, w, lam):
    return 0.5 * np.sum(w * (x - y) ** 2) + lam * np.sum(np.abs(np.diff(x)))


def port_dp(lib, y, w, lam):
    n = y.size
    out = np.empty(n)
    bufs = [np.empty(2 * n) for _ in range(3)] + [np.empty(n), np.empty(n)]
    lib.gflp_tf_dp_weight(C.c_int(n), _dp(y), _dp(w), C.c_double(lam), _dp(out), *[_dp(b) for b in bufs])
    return out


def ref_dp(lib, y, w, lam):
    n = y.size
    out = np.empty(n)
    bufs = [np.empty(2 * n) for _ in range(3)] + [np.empty(n), np.empty(n)]
    # void tf_dp_weight(int n, double *y, double *w, double lam, double *beta, double *x, double *a, double *b, double *tm, double *tp)
    lib.tf_dp_weight(C.c_int(n), _dp(y), _dp(w), C.c_double(lam), _dp(out), *[_dp(b) for b in bufs])
    return out


@pytest.mark.parametrize("n,lam", [(1, 1.0), (2, 0.5), (3, 0.0), (17, 0.3), (200, 1.0), (1000, 5.0), (64, 1e6)])
def test_dp_port_bitwise_equals_reference(port_c, ref_c, n, lam):
    rng = np.random.default_rng(n)
    y = rng.normal(size=n)
    w = rng.uniform(0.1, 3.0, size=n)
    a, b = port_dp(port_c, y, w, lam), ref_dp(ref_c, y, w, lam)
    assert np.array_equal(a, b)


@pytest.mark.parametrize("n,lam", [(5, 0.7), (40, 0.2), (300, 2.0)])
def test_dp_is_the_minimiser(port_c, n, lam):
    """KKT of the 1-D weighted fused lasso: w(x-y) + lam * D'u = 0 with |u| <= 1 and u = sign(diff) where diff != 0."""
    rng = np.random.default_rng(100 + n)
    y = np.cumsum(rng.normal(size=n)) * 0.3 + rng.normal(size=n)
    w = rng.uniform(0.2, 2.0, size=n)
    x = port_dp(port_c, y, w, lam)
    g = w * (x - y)                 # gradient of the smooth part
    u = np.cumsum(g)[:-1] / lam     # dual variable of each difference
    assert abs(np.sum(g)) < 1e-9
    assert np.all(np.abs(u) <= 1 + 1e-9)
    d = np.diff(x)
    nz = np.abs(d) > 1e-12
    assert np.allclose(u[nz], np.sign(d[nz]), atol=1e-8)
    for _ in range(20):             # and no random perturbation improves the objective
        xp = x + rng.normal(scale=1e-3, size=n)
        assert tv1d_objective(xp, y, w, lam) >= tv1d_objective(x, y, w, lam) - 1e-12


def _trails(lib, n, maxdiag=0):
    tr = np.zeros(n * (n + 1), dtype=np.uint32); bp = np.zeros(2 * n, dtype=np.uint32)
    nt, tsz, nc = C.c_int(), C.c_uint(), C.c_uint()
    rc = lib.fbo_test_trails(n, maxdiag, tr.ctypes.data_as(C.POINTER(C.c_uint)), bp.ctypes.data_as(C.POINTER(C.c_uint)),
                             C.byref(nt), C.byref(tsz), C.byref(nc))
    assert rc == 0
    return tr[:tsz.value].copy(), bp[:nt.value].copy(), nc.value


@pytest.mark.parametrize("n", [2, 3, 5, 17, 64])
def test_triangle_trails_closed_form(port_c, n):
    """row-trail slot of cell p is p; column c starts at (N-1) + (c-1)(c+2)/2."""
    tr, bp, nc = _trails(port_c, n)
    N = n * (n + 1) // 2
    assert nc == N and tr.size == 2 * N - 2 and bp.size == 2 * (n - 1)
    assert np.array_equal(tr[:N - 1], np.arange(N - 1))
    rm = lambda i, j: i * n - i * (i - 1) // 2 + (j - i)
    for c in range(1, n):
        s = (N - 1) + (c - 1) * (c + 2) // 2
        assert list(tr[s:s + c + 1]) == [rm(i, c) for i in range(c + 1)]


@pytest.mark.parametrize("n,md", [(9, 3), (20, 7), (33, 17), (12, 2)])
def test_trapeze_trails_cover_band(port_c, n, md):
    tr, bp, nc = _trails(port_c, n, md)
    assert nc == md * (n - md + 1) + (md - 1) * md // 2
    counts = np.bincount(tr, minlength=nc)
    assert counts.max() <= 2 and counts.min() >= 1
    lens = np.diff(np.concatenate([[0], bp]))
    assert lens.min() >= 2 and lens.max() <= md


@pytest.mark.parametrize("n,lam,conv,maxsteps", [(6, 1.3, 1e-3, 30000), (40, 2.5, 1e-2, 30000), (25, 0.0, 1e-3, 5),
                                                   (60, 2.5, 1e-12, 120)])
def test_admm_port_bitwise_equals_reference(oracle_port, oracle_ref, n, lam, conv, maxsteps):
    rng = np.random.default_rng(n)
    nc = n * (n + 1) // 2
    y = rng.normal(size=nc); w = rng.uniform(0, 3, size=nc); w[rng.random(nc) < 0.2] = 0
    a = gfl_solve(oracle_port, n, y, w, lam, conv, maxsteps)
    b = gfl_solve(oracle_ref, n, y, w, lam, conv, maxsteps)
    assert a["steps"] == b["steps"] and a["alpha"] == b["alpha"]
    for k in ("beta", "z", "u"):
        assert np.array_equal(a[k], b[k]), k


def test_admm_penalty_rescale_is_exercised(oracle_port, oracle_ref):
    """The step%50 varying-penalty branch (gfl_graph_fl.c:168-179) needs > 50 iterations: tight tolerance."""
    n = 30; nc = n * (n + 1) // 2
    rng = np.random.default_rng(5)
    y = rng.normal(size=nc) * 3; w = rng.uniform(0.01, 0.1, size=nc)
    a = gfl_solve(oracle_port, n, y, w, 5.0, 1e-9, 400)
    b = gfl_solve(oracle_ref, n, y, w, 5.0, 1e-9, 400)
    assert a["steps"] > 51 and a["alpha"] != 5.0
    assert a["steps"] == b["steps"] and a["alpha"] == b["alpha"] and np.array_equal(a["beta"], b["beta"])


def test_admm_converges_to_the_2d_minimiser(oracle_port):
    """At convergence beta minimises sum w (y-b)^2 + lam sum_edges |b_i - b_j| over the triangle grid."""
    n = 12; nc = n * (n + 1) // 2
    rng = np.random.default_rng(2)
    y = rng.normal(size=nc); w = rng.uniform(0.5, 2, size=nc)
    lam = 0.8
    r = gfl_solve(oracle_port, n, y, w, lam, 1e-10, 30000)
    rm = lambda i, j: i * n - i * (i - 1) // 2 + (j - i)
    edges = [(rm(i, j), rm(i, j + 1)) for i in range(n) for j in range(i, n - 1)] + \
            [(rm(i, j), rm(i + 1, j)) for i in range(n) for j in range(i + 1, n)]
    e = np.array(edges)
    obj = lambda b: np.sum(w * (y - b) ** 2) + lam * np.sum(np.abs(b[e[:, 0]] - b[e[:, 1]]))
    b0 = r["beta"]
    for _ in range(50):
        assert obj(b0 + rng.normal(scale=1e-4, size=nc)) >= obj(b0) - 1e-7


def test_warm_start_chain_port_equals_reference(oracle_port, oracle_ref):
    n = 24; nc = n * (n + 1) // 2
    rng = np.random.default_rng(8)
    y = rng.normal(size=nc); w = rng.uniform(0, 2, size=nc)
    outs = []
    for lib in (oracle_port, oracle_ref):
        s1 = gfl_solve(lib, n, y, w, 2.0, 1e-2)
        s2 = gfl_solve(lib, n, y * 1.1, w, 4.0, 5e-3, state=s1)
        outs.append(s2)
    assert outs[0]["steps"] == outs[1]["steps"]
    assert np.array_equal(outs[0]["beta"], outs[1]["beta"])
