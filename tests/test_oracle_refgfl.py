"""CPU: rows a9 / a12 of SURVEY.md section 8 pinned against the reference ITSELF. oracle/Makefile compiles the reference's
GFLLibrary.hpp + GFLLibraryTriangleImpl.cpp + GFLLibraryTrapezeImpl.cpp where they lie (against the stand-in
oracle/stub/rcpp/Rcpp.h) into oracle/_ref/librefgfl.so; the restatement in oracle/fast_binless_oracle.cpp must give
the same trails / breakpoints and, linked to the same reference C core, the same beta (with fill_triangle), z, u, alpha
and step counts BIT FOR BIT, cold and warm (get_state / set_state), triangle and trapeze."""
import ctypes as C
import os

import numpy as np
import pytest

from binless_b200._abi import gfl_solve

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
dp = C.POINTER(C.c_double); up = C.POINTER(C.c_uint)


@pytest.fixture(scope="module")
def ref():
    path = os.path.join(ROOT, "oracle", "_ref", "librefgfl.so")
    if not os.path.exists(path):
        pytest.skip("oracle/_ref/librefgfl.so not built (needs /root/reference at build time)")
    lib = C.CDLL(path)
    lib.refgfl_new.restype = C.c_void_p; lib.refgfl_new.argtypes = [C.c_uint, C.c_uint]
    lib.refgfl_free.argtypes = [C.c_void_p]; lib.refgfl_reset.argtypes = [C.c_void_p]
    lib.refgfl_geometry.argtypes = [C.c_void_p, up, up, C.POINTER(C.c_int)]
    lib.refgfl_trails.argtypes = [C.c_void_p, up, up]
    lib.refgfl_optimize.argtypes = [C.c_void_p, C.c_uint, dp, dp, C.c_double, C.c_double, dp]
    lib.refgfl_get_state.argtypes = [C.c_void_p, dp, dp, dp, dp, up, C.POINTER(C.c_int)]
    lib.refgfl_set_state.argtypes = [C.c_void_p, C.c_uint, C.c_uint, dp, dp, dp, C.c_double, C.c_uint, C.c_int]
    lib.refgfl_ninner.restype = C.c_uint; lib.refgfl_ninner.argtypes = [C.c_void_p]
    lib.refgfl_alpha.restype = C.c_double; lib.refgfl_alpha.argtypes = [C.c_void_p]
    return lib


def _p(a):
    return a.ctypes.data_as(dp)


def _geometry(ref, h):
    N = C.c_uint(); tsz = C.c_uint(); nt = C.c_int()
    kind = ref.refgfl_geometry(h, C.byref(N), C.byref(tsz), C.byref(nt))
    return kind, N.value, tsz.value, nt.value


CASES = [(2, 0), (5, 0), (33, 0), (60, 0), (40, 7), (41, 21), (41, 22), (12, 2), (30, 1), (64, 40)]


@pytest.mark.parametrize("n,md", CASES)
def test_trails_and_breakpoints_equal_the_reference(ref, oracle_ref, n, md):
    """store_trails / triangle_grid_chain / trapezoidal_grid_chain (GFLLibraryTriangleImpl.cpp:7-49,
    GFLLibraryTrapezeImpl.cpp:13-81) and the dispatch of GFLLibrary.hpp:21-27."""
    h = ref.refgfl_new(n, md)
    assert h
    kind, N, tsz, nt = _geometry(ref, h)
    tr = np.zeros(tsz, dtype=np.uint32); bp = np.zeros(nt, dtype=np.uint32)
    ref.refgfl_trails(h, tr.ctypes.data_as(up), bp.ctypes.data_as(up))
    ref.refgfl_free(h)
    f = oracle_ref.lib.fbo_test_trails
    f.argtypes = [C.c_int, C.c_uint, up, up, C.POINTER(C.c_int), up, up]
    tr2 = np.zeros(2 * n * (n + 1), dtype=np.uint32); bp2 = np.zeros(2 * n + 4, dtype=np.uint32)
    nt2 = C.c_int(); tsz2 = C.c_uint(); nc2 = C.c_uint()
    assert f(n, md, tr2.ctypes.data_as(up), bp2.ctypes.data_as(up), C.byref(nt2), C.byref(tsz2), C.byref(nc2)) == 0
    assert (nt2.value, tsz2.value, nc2.value) == (nt, tsz, N)
    assert np.array_equal(tr, tr2[:tsz]) and np.array_equal(bp, bp2[:nt])
    assert kind == (1 if (1 < md <= (n + 1) // 2) else 0)


@pytest.mark.parametrize("n,md", CASES)
def test_solve_state_and_fill_equal_the_reference(ref, oracle_ref, n, md):
    """optimize + get_state / set_state + extract_trapeze / fill_triangle (GFLLibraryTriangleImpl.cpp:60-87,
    GFLLibraryTrapezeImpl.cpp:93-162): a cold solve, then a warm one from the exported state, on the reference object
    and on the restatement (both over the reference's gfl_graph_fl.c / gfl_tf.c)."""
    assert oracle_ref.kind == "reference"
    rng = np.random.default_rng(n * 100 + md)
    nc = n * (n + 1) // 2
    y = rng.normal(size=nc); w = rng.uniform(0, 3, size=nc); w[rng.random(nc) < 0.2] = 0
    h = ref.refgfl_new(n, md)
    kind, N, tsz, nt = _geometry(ref, h)
    full = np.empty(nc); z = np.empty(tsz); u = np.empty(tsz); beta = np.empty(N)
    alpha = C.c_double(); counter = C.c_uint(); mdo = C.c_int()
    ref.refgfl_optimize(h, nc, _p(y), _p(w), 1.7, 1e-3, _p(full))
    ref.refgfl_get_state(h, _p(z), _p(u), _p(beta), C.byref(alpha), C.byref(counter), C.byref(mdo))
    a = gfl_solve(oracle_ref, n, y, w, 1.7, 1e-3, 30000, md)
    assert a["steps"] == counter.value and a["alpha"] == alpha.value and mdo.value == (md if kind else 0)
    for k, v in (("beta", beta), ("z", z), ("u", u), ("beta_full", full)):
        assert np.array_equal(a[k], v), k
    # warm start on a fresh reference object from the exported state
    h2 = ref.refgfl_new(n, md)
    ref.refgfl_set_state(h2, N, tsz, _p(z), _p(u), _p(beta), alpha.value, counter.value, mdo.value)
    y2 = y * 1.1
    ref.refgfl_optimize(h2, nc, _p(y2), _p(w), 2.9, 5e-4, _p(full))
    z2 = np.empty(tsz); u2 = np.empty(tsz); b2 = np.empty(N)
    ref.refgfl_get_state(h2, _p(z2), _p(u2), _p(b2), C.byref(alpha), C.byref(counter), C.byref(mdo))
    b = gfl_solve(oracle_ref, n, y2, w, 2.9, 5e-4, 30000, md, state=a)
    assert counter.value == a["steps"] + b["steps"] and b["alpha"] == alpha.value
    for k, v in (("beta", b2), ("z", z2), ("u", u2), ("beta_full", full)):
        assert np.array_equal(b[k], v), k
    ref.refgfl_free(h); ref.refgfl_free(h2)
