"""GPU (-m gpu): fb_wgfl_signal_bic / fb_wgfl_diff_bic (the optimized path's performance iteration,
perf_iteration_signal.cpp:22-157, perf_iteration_diff.cpp:20-117) against the CPU oracle, whose scatter front end is the
reference's own cts_core.c. Same IRLS iteration counts, same candidates, same chosen (lambda1, eCprime, dof), patch labels
bit-exact, columns to 1e-8."""
import numpy as np
import pytest

from binless_b200._abi import wgfl_diff_bic, wgfl_signal_bic
from binless_b200.synth import synth_cts

pytestmark = pytest.mark.gpu
ATOL = 1e-8


def close(a, b, what):
    a = np.asarray(a); b = np.asarray(b)
    fin = np.isfinite(b)
    assert np.array_equal(np.isfinite(a), fin), what
    assert np.array_equal(a[~fin], b[~fin]), what
    if fin.any():
        assert float(np.max(np.abs(a[fin] - b[fin]) / (1 + np.abs(b[fin])))) < ATOL, what


def compare_common(g, o):
    assert g["converged"] == o["converged"] and g["nouter_done"] == o["nouter_done"]
    assert g["admm_iters"] == o["admm_iters"]
    assert g["ncandidates"] == o["ncandidates"] and g["dof"] == o["dof"]
    assert abs(g["lambda1"] - o["lambda1"]) < ATOL and abs(g["eCprime"] - o["eCprime"]) < ATOL
    assert abs(g["BIC"] - o["BIC"]) < 1e-9 * (1 + abs(o["BIC"])) and g["BIC_sd"] == o["BIC_sd"] == -1
    assert np.array_equal(g["patchno"], o["patchno"])
    for k in ("beta", "nobs", "weight"):
        close(g[k], o[k], k)
    for k in ("beta", "z", "u"):
        close(g["GFLState"][k], o["GFLState"][k], "state " + k)
    assert abs(g["GFLState"]["alpha"] - o["GFLState"]["alpha"]) == 0


SIGNAL_CASES = [
    # nbins, seed, lam2, tol, constrained, fixed, positive, lambda1_fixed
    (12, 3, 1.0, 0.1, True, True, True, False),
    (33, 4, 2.0, 0.1, True, True, False, False),
    (33, 5, 2.0, 0.1, False, False, True, False),       # EstimatedOffset
    (64, 6, 3.0, 0.2, True, True, True, True),          # FixedSparsity
    (129, 7, 2.5, 0.1, False, True, False, False),
]


@pytest.mark.parametrize("nbins,seed,lam2,tol,constrained,fixed,positive,l1fixed", SIGNAL_CASES)
def test_wgfl_signal_bic_matches_oracle(gpu_lib, oracle, nbins, seed, lam2, tol, constrained, fixed, positive, l1fixed):
    cts = synth_cts(nbins, seed)
    nc = nbins * (nbins + 1) // 2
    args = (cts, cts["dispersion"], 20, nbins, None, lam2, tol, cts["metadata"], np.zeros(nc), constrained, fixed, positive,
            l1fixed, 0.25)
    o = wgfl_signal_bic(oracle, *args)
    g = wgfl_signal_bic(gpu_lib, *args)
    compare_common(g, o)
    for k in ("phi", "phihat"):
        close(g[k], o[k], k)
    # second round, warm: the GFLState and beta of the first call go back in (what the optimized path does every iteration)
    args2 = (cts, cts["dispersion"], 20, nbins, o["GFLState"], lam2 * 1.1, tol, cts["metadata"], o["beta"], constrained, fixed,
             positive, l1fixed, 0.25)
    o2 = wgfl_signal_bic(oracle, *args2)
    g2 = wgfl_signal_bic(gpu_lib, *args2)
    compare_common(g2, o2)


@pytest.mark.parametrize("nouter", [1, 2])
def test_wgfl_signal_bic_cold_restart_path(gpu_lib, oracle, nouter):
    """nouter too small to converge: nwarm = nouter / 2 warm iterations, then the cold restart (perf_iteration_signal.cpp:45-57)"""
    nbins = 40
    cts = synth_cts(nbins, 9)
    nc = nbins * (nbins + 1) // 2
    rng = np.random.default_rng(1)
    args = (cts, cts["dispersion"], nouter, nbins, None, 2.0, 0.05, cts["metadata"], rng.normal(0, 1, nc), True, True, True)
    o = wgfl_signal_bic(oracle, *args)
    g = wgfl_signal_bic(gpu_lib, *args)
    assert not o["converged"]
    compare_common(g, o)


@pytest.mark.parametrize("nbins,seed,constrained,l1fixed", [(12, 2, True, False), (48, 3, False, False), (96, 4, True, True)])
def test_wgfl_diff_bic_matches_oracle(gpu_lib, oracle, nbins, seed, constrained, l1fixed):
    ref = synth_cts(nbins, seed); oth = synth_cts(nbins, seed + 100, shift=0.5)
    nc = nbins * (nbins + 1) // 2
    args = (oth, ref, ref["dispersion"], 20, nbins, None, 1.5, 0.1, ref["metadata"], np.zeros(nc), np.zeros(nc), constrained,
            l1fixed, 0.2)
    o = wgfl_diff_bic(oracle, *args)
    g = wgfl_diff_bic(gpu_lib, *args)
    compare_common(g, o)
    for k in ("phi_ref", "delta", "phihat", "phihat_var", "phihat_ref", "phihat_var_ref", "deltahat"):
        close(g[k], o[k], k)


def test_wgfl_rejects_bad_cts(gpu_lib):
    from binless_b200._abi import FastBinlessError
    cts = synth_cts(10, 1)
    bad = dict(cts); bad["bin2"] = cts["bin2"].copy(); bad["bin2"][0] = 11
    with pytest.raises(FastBinlessError, match="bins must satisfy"):
        wgfl_signal_bic(gpu_lib, bad, 2.0, 5, 10, None, 1.0, 0.1, cts["metadata"], np.zeros(55), True, True, True)
