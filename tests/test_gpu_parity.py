"""GPU (-m gpu): the CUDA library, called through the C ABI, against the CPU oracle on the same seeded inputs.

Tolerances (BASELINE.json north_star): log-decay, log-biases, log-signal within 1e-6 absolute after the same
IRLS / ADMM iteration counts; fused-patch labels bit-exact. Measured agreement is ~1e-12; the bound asserted
here is 1e-8 so that a regression is caught long before the contractual 1e-6."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tools"))
sys.path.insert(0, os.path.join(ROOT, "tests"))
from make_golden import CASES  # noqa: E402
from test_golden import check_against_golden  # noqa: E402
from binless_b200._abi import FastBinlessError, gfl_solve  # noqa: E402
from binless_b200.synth import synth  # noqa: E402

pytestmark = pytest.mark.gpu
ATOL = 1e-8


def maxabs(a, b):
    return float(np.max(np.abs(np.asarray(a) - np.asarray(b)))) if np.size(a) else 0.0


def compare_signal(og, oo):
    dg, do = og["diagnostics"], oo["diagnostics"]
    assert dg["nsteps"] == do["nsteps"] and dg["converged"] == do["converged"]
    assert list(dg["solve_iters"]) == list(do["solve_iters"])          # same ADMM iteration counts
    assert dg["admm_cell_iters"] == do["admm_cell_iters"]
    assert maxabs(og["log_signal"], oo["log_signal"]) < ATOL          # log-signal
    assert maxabs(og["biases"]["estimate"], oo["biases"]["estimate"]) < ATOL   # log-biases
    assert maxabs(og["decay"]["beta"], oo["decay"]["beta"]) < ATOL
    assert maxabs(np.log(og["decaymat"]), np.log(oo["decaymat"])) < ATOL       # log-decay per cell
    assert maxabs(og["exposures"]["estimate"], oo["exposures"]["estimate"]) < ATOL
    assert abs(og["decay"]["lambda_"] - oo["decay"]["lambda_"]) < 1e-7 * (1 + abs(oo["decay"]["lambda_"]))
    assert maxabs(og["phihat"], oo["phihat"]) < 1e-7 * (1 + np.abs(oo["phihat"]).max())
    for k in ("weight", "signal", "binless", "background", "biasmat", "decaymat"):
        assert float(np.max(np.abs(og[k] - oo[k]) / (1 + np.abs(oo[k])))) < ATOL, k
    assert np.array_equal(og["patchno"], oo["patchno"])                # labels bit-exact
    assert maxabs(dg["prec_abs"], do["prec_abs"]) < ATOL


def compare_diff(a, b):
    assert list(a["diagnostics"]["solve_iters"]) == list(b["diagnostics"]["solve_iters"])
    for k in ("difference", "deltahat", "weight", "phi_ref"):
        assert float(np.max(np.abs(a[k] - b[k]) / (1 + np.abs(b[k])))) < ATOL, k
    assert np.array_equal(a["patchno"], b["patchno"])


SIGNAL_CASES = [
    # n, D, seed, alpha, lam2, lam1, kwargs
    (2, 1, 3, 0.49, 2.5, 0.0, {}),                                      # smallest grid the solver accepts
    (3, 2, 4, 0.49, 2.5, 0.0, {}),
    (5, 2, 3, 0.49, 2.5, 0.0, {}),
    (33, 2, 3, 0.49, [2.53, 4.51], [0.13, 0.10], {}),                   # ragged vs the 32-wide tiles
    (129, 1, 12, 0.7, 1.0, 0.0, {}),                                    # one past the 128-wide tile
    (200, 2, 1, 0.49, [2.53, 4.51], [0.13, 0.10], {}),                  # config C1 (SEMA3C-like)
    (200, 1, 1, 0.49, 2.53, 0.1, dict(tol_val=0.05, nouter=8)),         # no convergence: runs to nouter
    (131, 3, 5, 0.6, 1.5, 0.0, dict(maxdiag=40)),                       # trapeze band
    (64, 2, 6, 0.5, 2.5, 0.0, dict(maxdiag=33)),                        # maxdiag > (n+1)/2 falls back to triangle
    (257, 2, 6, 0.5, 2.5, 0.0, dict(bg_steps=2, nouter=6, tol_val=0.01)),   # > 50 ADMM iterations: penalty rescale
    (90, 2, 7, 0.5, 0.0, 0.0, dict(nouter=6)),                          # lam2 = 0: prox is the identity (one signal step:
                                                                        # unsmoothed per-cell IRLS amplifies rounding by |z| per step)
    (90, 2, 7, 0.5, 1e4, 0.0, dict(nouter=7)),                          # huge lam2: one constant per chain
    (70, 2, 8, 0.5, 2.5, 0.0, dict(compute_patchnos=False)),
    (70, 2, 8, 0.5, 2.5, 0.0, dict(bg_steps=0, nouter=3)),              # signal from the first step
    (70, 2, 8, 0.5, 2.5, 0.0, dict(nouter=1)),
    (150, 4, 9, 1.3, [1.0, 2.0, 3.0, 4.0], [0.0, 0.1, 0.2, 0.3], dict(free_decay=40000)),
]


@pytest.mark.parametrize("n,D,seed,alpha,lam2,lam1,kw", SIGNAL_CASES)
def test_fast_binless_matches_oracle(gpu_lib, oracle, n, D, seed, alpha, lam2, lam1, kw):
    mat = synth(n, D, 5000, seed)
    with gpu_lib.create(mat, n) as hg, oracle.create(mat, n) as ho:
        og = hg.run_signal(alpha, lam2, lam1, **kw)
        oo = ho.run_signal(alpha, lam2, lam1, **kw)
        assert hg.kernel_launches() > 0
        compare_signal(og, oo)
        if D > 1:
            md = kw.get("maxdiag", 0)
            for ref in (1, D):
                compare_diff(hg.run_difference(ref, alpha, 3.38, 0.14, maxdiag=md),
                             ho.run_difference(ref, alpha, 3.38, 0.14, maxdiag=md))
        else:
            cg = hg.eval_cv(alpha, [1.0, 5.0, 0.3], 0.05, group=1)
            co = ho.eval_cv(alpha, [1.0, 5.0, 0.3], 0.05, group=1)
            for a, b in zip(cg, co):
                for k in ("signal", "binless", "phihat", "weight", "log_signal"):
                    assert float(np.max(np.abs(a[k] - b[k]) / (1 + np.abs(b[k])))) < ATOL, k
                assert np.array_equal(a["patchno"], b["patchno"])


def test_sparse_and_dead_inputs(gpu_lib, oracle):
    """Rows without observables (nobs = 0 => z = w = 0, fast_residuals.ipp:55-62) and an all-zero dataset."""
    n = 60
    mat = synth(n, 2, 5000, 21, depth=-5.0)          # < 1 read per cell
    nc = n * (n + 1) // 2
    mat["observed"][nc:] = 0                          # second dataset is empty except the corner read
    mat["observed"][nc + n - 1] = 1
    with gpu_lib.create(mat, n) as hg, oracle.create(mat, n) as ho:
        compare_signal(hg.run_signal(0.5, 2.5, 0.0), ho.run_signal(0.5, 2.5, 0.0))


@pytest.mark.parametrize("name", sorted(CASES))
def test_cuda_reproduces_golden(gpu_lib, name):
    check_against_golden(gpu_lib, name, exact=False)


@pytest.mark.parametrize("n,lam,conv,maxsteps,md", [(2, 1.3, 1e-3, 30000, 0), (17, 1.3, 1e-3, 30000, 0),
                                                      (100, 1.3, 1e-3, 30000, 0), (64, 0.0, 1e-3, 5, 0),
                                                      (200, 2.5, 1e-12, 120, 0), (200, 2.5, 1e-12, 52, 0),
                                                      (120, 2.0, 1e-3, 30000, 30), (300, 50.0, 1e-4, 30000, 0)])
@pytest.mark.parametrize("prox", ["hybrid", "aset", "aset_fallback", "aset_2passes", "team", "fold"])
def test_gfl_solve_matches_oracle(gpu_lib, oracle, monkeypatch, n, lam, conv, maxsteps, md, prox):
    """The stand-alone fused-lasso solve: same `step`, alpha trajectory, beta / z / u in the reference's slot order.
    Every variant of the chain prox: the default (taut-string team kernel for the first steps of a cold start, parallel
    active set afterwards), the parallel active set from the first step, its sequential fallback forced for every
    batch (FB_ASET_MAXPASS=0) or after two passes, and the team kernel alone (FB_PROX_ASET=0)."""
    for k, v in {"hybrid": {}, "aset": {"FB_ASET_FROM": "1"}, "aset_fallback": {"FB_ASET_FROM": "1", "FB_ASET_MAXPASS": "0"},
                 "aset_2passes": {"FB_ASET_FROM": "1", "FB_ASET_MAXPASS": "2"}, "team": {"FB_PROX_ASET": "0"},
                 "fold": {"FB_CTL_FOLD": "1"}}[prox].items():   # the last CTA of the chain prox ends the iteration (fb_ctl.cuh)
        monkeypatch.setenv(k, v)
    rng = np.random.default_rng(n + maxsteps)
    nc = n * (n + 1) // 2
    y = rng.normal(size=nc); w = rng.uniform(0, 3, size=nc); w[rng.random(nc) < 0.2] = 0
    a = gfl_solve(oracle, n, y, w, lam, conv, maxsteps, md)
    b = gfl_solve(gpu_lib, n, y, w, lam, conv, maxsteps, md)
    assert a["steps"] == b["steps"] and a["alpha"] == b["alpha"]
    for k in ("beta", "z", "u", "beta_full"):
        assert maxabs(a[k], b[k]) < ATOL, k


@pytest.mark.parametrize("n,lam,rtab", [(200, 2.5, 4), (300, 50.0, 16), (120, 1.3, 2)])
def test_gfl_long_chain_mode_matches_oracle(gpu_lib, oracle, monkeypatch, n, lam, rtab):
    """The long-chain variant of the prox kernel (reciprocal table shorter than the chain, as used above ~14 000
    bins) forced at a small size: identical steps and solution."""
    rng = np.random.default_rng(n)
    nc = n * (n + 1) // 2
    y = rng.normal(size=nc); w = rng.uniform(0, 3, size=nc); w[rng.random(nc) < 0.2] = 0
    monkeypatch.setenv("FB_PROX_ASET", "0")
    a = gfl_solve(oracle, n, y, w, lam, 1e-4, 30000, 0)
    ref = gfl_solve(gpu_lib, n, y, w, lam, 1e-4, 30000, 0)
    monkeypatch.setenv("FB_PROX_RTAB", str(rtab))
    b = gfl_solve(gpu_lib, n, y, w, lam, 1e-4, 30000, 0)
    assert a["steps"] == b["steps"] and a["alpha"] == b["alpha"]
    for k in ("beta", "z", "u", "beta_full"):
        assert maxabs(a[k], b[k]) < ATOL, k
        assert np.array_equal(ref[k], b[k]), k       # 1/m is the same correctly rounded value either way


@pytest.mark.parametrize("n,lam,steps,check_oracle", [(1500, 2.5, 14, True), (1300, 400.0, 14, False), (700, 2.5, 30, False)])
def test_gfl_scratch_mode_matches_shared_memory_mode(gpu_lib, oracle, monkeypatch, n, lam, steps, check_oracle):
    """The team kernel with the chains' prefix sums in global scratch (what runs above ~4500 bins), forced at sizes the
    oracle still finishes, with short and with very long segments (lam = 400): bit-identical to the shared-memory mode."""
    rng = np.random.default_rng(n)
    nc = n * (n + 1) // 2
    y = np.repeat(rng.normal(0, 1, nc // 40 + 1), 40)[:nc] + rng.normal(0, 0.5, nc)
    w = rng.uniform(0, 3, size=nc); w[rng.random(nc) < 0.1] = 0
    monkeypatch.setenv("FB_PROX_ASET", "0")
    ref = gfl_solve(gpu_lib, n, y, w, lam, 1e-12, steps, 0)
    monkeypatch.setenv("FB_PROX_GS", "1")
    b = gfl_solve(gpu_lib, n, y, w, lam, 1e-12, steps, 0)
    assert ref["steps"] == b["steps"] and ref["alpha"] == b["alpha"]
    for k in ("beta", "z", "u", "beta_full"):
        assert np.array_equal(ref[k], b[k]), k
    if check_oracle:
        a = gfl_solve(oracle, n, y, w, lam, 1e-12, steps, 0)
        assert a["steps"] == b["steps"] and a["alpha"] == b["alpha"]
        for k in ("beta", "z", "u", "beta_full"):
            assert maxabs(a[k], b[k]) < ATOL, k


def test_gfl_warm_start_chain(gpu_lib, oracle):
    n = 120; nc = n * (n + 1) // 2
    rng = np.random.default_rng(3)
    y = rng.normal(size=nc); w = rng.uniform(0, 2, size=nc)
    res = []
    for lib in (gpu_lib, oracle):
        s1 = gfl_solve(lib, n, y, w, 2.53, 1e-2)
        s2 = gfl_solve(lib, n, y * 1.1, w, 4.0, 5e-3, state=s1)
        res.append(s2)
    assert res[0]["steps"] == res[1]["steps"] and res[0]["alpha"] == res[1]["alpha"]
    assert maxabs(res[0]["beta"], res[1]["beta"]) < ATOL


def test_restore_then_difference(gpu_lib, oracle):
    """fast_binless_difference from a saved result (set_state path, fast_binless.cpp:244-260)."""
    mat = synth(80, 3, 5000, 8)
    with oracle.create(mat, 80) as ho:
        out = ho.run_signal(0.5, 2.5, 0.0)
        want = ho.run_difference(2, 0.5, [3.0, 1.0])
    with gpu_lib.create(mat, 80) as hg:
        hg.restore(out["log_signal"], out["biases"], out["decay"], out["exposures"])
        compare_diff(hg.run_difference(2, 0.5, [3.0, 1.0]), want)


def test_input_errors(gpu_lib):
    """Same messages as the reference's Rcpp::stop calls (FastData.ipp:20-25, FastData.hpp:90, fast_binless.cpp:198)."""
    mat = synth(10, 2, 5000, 1)
    bad = dict(mat); bad["name"] = mat["name"] + 1
    with pytest.raises(FastBinlessError, match="Input size should be"):
        gpu_lib.create(bad, 10)
    with pytest.raises(FastBinlessError, match="Input size should be"):
        gpu_lib.create(mat, 11)
    bad = dict(mat); bad["bin1"] = mat["bin1"].copy(); bad["bin1"][bad["bin1"] == 10] = 11
    with pytest.raises(FastBinlessError, match="bins should range"):
        gpu_lib.create(bad, 10)
    bad = dict(mat); bad["bin2"] = mat["bin2"][::-1].copy()
    with pytest.raises(FastBinlessError, match="sorted"):
        gpu_lib.create(bad, 10)
    bad = dict(mat); bad["distance"] = mat["distance"].copy(); bad["distance"][3] += 1
    with gpu_lib.create(bad, 10) as h:      # not a bin_data() layout: accepted, binned per cell
        assert h.run_signal(0.5, 2.5)["diagnostics"]["nsteps"] >= 1
    bad = dict(mat); bad["distance"] = mat["distance"].copy(); bad["distance"][3] = 0.0
    with pytest.raises(FastBinlessError, match="distance must be positive"):
        gpu_lib.create(bad, 10)
    with gpu_lib.create(mat, 10) as h:
        with pytest.raises(FastBinlessError) as ei:
            h.run_difference(1, 0.5, 2.0)
        assert ei.value.code == 3
        h.run_signal(0.5, 2.5)
        with pytest.raises(FastBinlessError, match="Ref is not in the allowed range"):
            h.run_difference(3, 0.5, 2.0)
        with pytest.raises(FastBinlessError, match="Provide only 1 dataset"):
            h.eval_cv(0.5, [1.0, 2.0])
    # canonical order is implied when name / bin1 / bin2 are not passed
    slim = {k: (None if k in ("name", "bin1", "bin2") else v) for k, v in mat.items()}
    with gpu_lib.create(slim, 10) as h:
        assert h.run_signal(0.5, 2.5)["diagnostics"]["nsteps"] >= 1


def test_interrupt_callback(gpu_lib):
    import ctypes as C
    mat = synth(40, 1, 5000, 1)
    CB = C.CFUNCTYPE(C.c_int, C.c_void_p)
    calls = []
    cb = CB(lambda arg: (calls.append(1), int(len(calls) >= 2))[1])
    with gpu_lib.create(mat, 40) as h:
        gpu_lib.lib.fb_set_interrupt_cb.argtypes = [C.c_void_p, CB, C.c_void_p]
        gpu_lib.lib.fb_set_interrupt_cb(h.h, cb, None)
        with pytest.raises(FastBinlessError) as ei:
            h.run_signal(0.5, 2.5)
        assert ei.value.code == 5 and len(calls) == 2


def test_python_entry_points_mirror_r_signatures(oracle):
    from binless_b200 import fast_binless, fast_binless_difference
    mat = synth(64, 2, 5000, 7)
    out = fast_binless(mat, 64, 0.49, [2.53, 4.51], [0.13, 0.10])
    assert set(out) >= {"background", "biasmat", "decaymat", "signal", "binless", "phihat", "weight", "patchno",
                        "biases", "decay", "exposures", "log_signal", "nbins"}
    diff = fast_binless_difference(out, mat, 1, 0.49, 3.38, 0.14)
    assert set(diff) >= {"difference", "deltahat", "weight", "phi_ref", "patchno"}
    with oracle.create(mat, 64) as ho:
        compare_signal(out, ho.run_signal(0.49, [2.53, 4.51], [0.13, 0.10]))
        compare_diff(diff, ho.run_difference(1, 0.49, 3.38, 0.14))


def test_full_size_c2_properties_and_determinism(gpu_lib):
    """BASELINE config 2 (2400 bins x 4 datasets) without the oracle: size-independent properties."""
    n, D = 2400, 4
    mat = synth(n, D, 5000, 2)
    nc = n * (n + 1) // 2
    slim = {k: (None if k in ("name", "bin1", "bin2") else v) for k, v in mat.items()}
    with gpu_lib.create(slim, n) as h:
        a = h.run_signal(0.5, 2.5, 0.0)
        b = h.run_signal(0.5, 2.5, 0.0)
    for k in ("log_signal", "phihat", "weight", "binless", "patchno"):   # fixed-shape reductions: same bits every run
        assert np.array_equal(a[k], b[k]), k
    ls = a["log_signal"]
    nobs = mat["nobs"].astype(np.float64)
    for d in range(D):
        s = slice(d * nc, (d + 1) * nc)
        assert ls[s].min() == 0.0                                        # shift_signal
        lab = a["patchno"][s]
        assert lab.min() == 0 and lab[0] == 0
        # labels are numbered in order of each component's lowest cell
        first = np.full(lab.max() + 1, nc, dtype=np.int64)
        np.minimum.at(first, lab, np.arange(nc))
        assert np.all(np.diff(first) > 0)
    assert np.allclose(a["binless"], a["decaymat"] * a["signal"], rtol=1e-12)
    assert np.allclose(a["background"][nobs == 0], 0.0)
    assert np.all(np.diff(a["decay"]["beta"][30:]) <= 1e-9)
    assert a["diagnostics"]["admm_cell_iters"] == int(np.sum(a["diagnostics"]["solve_iters"])) * nc


@pytest.mark.timeout(900)
def test_full_size_c2_matches_oracle(gpu_lib, oracle):
    """BASELINE config 2 against the oracle end to end (~20-30 s of single-thread CPU)."""
    n, D = 2400, 4
    mat = synth(n, D, 5000, 2)
    with gpu_lib.create(mat, n) as hg:
        og = hg.run_signal(0.5, 2.5, 0.0)
    with oracle.create(mat, n) as ho:
        oo = ho.run_signal(0.5, 2.5, 0.0)
    compare_signal(og, oo)


def test_sweep_lambda_matches_oracle(gpu_lib, oracle):
    """fb_sweep_lambda (config 3): cold-start solve + BIC per (lambda2, dataset); dof and step counts exact."""
    n, D = 64, 2
    mat = synth(n, D, 5000, 11)
    lams = 10.0 ** np.linspace(-1, 2, 7)
    with oracle.create(mat, n) as ho:
        fit = ho.run_signal(0.5, 2.5, 0.0)
        ro = ho.sweep_lambda(0.5, lams, [0.05, 0.1], tol_val=0.2)
    with gpu_lib.create(mat, n) as hg:
        hg.restore(fit["log_signal"], fit["biases"], fit["decay"], fit["exposures"])
        rg = hg.sweep_lambda(0.5, lams, [0.05, 0.1], tol_val=0.2)
        assert hg.kernel_launches() > 0
    assert np.array_equal(rg["steps"], ro["steps"])
    assert np.array_equal(rg["dof"], ro["dof"])
    assert np.allclose(rg["chisq"], ro["chisq"], rtol=1e-9, atol=0)
    assert np.allclose(rg["bic"], ro["bic"], rtol=1e-9, atol=0)
    # band variant
    with oracle.create(mat, n) as ho:
        ho.restore(fit["log_signal"], fit["biases"], fit["decay"], fit["exposures"])
        ro = ho.sweep_lambda(0.5, lams[:3], 0.0, tol_val=0.2, maxdiag=20)
    with gpu_lib.create(mat, n) as hg:
        hg.restore(fit["log_signal"], fit["biases"], fit["decay"], fit["exposures"])
        rg = hg.sweep_lambda(0.5, lams[:3], 0.0, tol_val=0.2, maxdiag=20)
    assert np.array_equal(rg["steps"], ro["steps"]) and np.array_equal(rg["dof"], ro["dof"])
    assert np.allclose(rg["bic"], ro["bic"], rtol=1e-9, atol=0)


def test_csv_out_matches_oracle(gpu_lib, oracle, tmp_path):
    """csv_out (fast_dataframe.cpp:23-57): the reference as built streams gzip (csvfile.hpp:46-48, USE_BOOST = 1 in
    binless/src/Makevars:1-2): quoted header + one "%.4f" row per cell; gunzip of the product's file = the oracle's text."""
    import gzip
    n, D = 33, 2
    mat = synth(n, D, 5000, 3)
    fg, fo = str(tmp_path / "gpu.csv.gz"), str(tmp_path / "cpu.csv")
    with gpu_lib.create(mat, n) as hg:
        hg.run_signal(0.49, [2.53, 4.51], [0.13, 0.10], csv_out=fg)
    with oracle.create(mat, n) as ho:
        ho.run_signal(0.49, [2.53, 4.51], [0.13, 0.10], csv_out=fo)
    assert open(fg, "rb").read(3) == b"\x1f\x8b\x08"
    lg, lo = gzip.open(fg, "rt").read().splitlines(), open(fo).read().splitlines()
    assert lg[0] == lo[0] == '"phihat","weight","biasmat","decaymat","signal","binless","background"'
    assert len(lg) == len(lo) == 1 + D * n * (n + 1) // 2
    a = np.array([[float(x) for x in l.split(",")] for l in lg[1:]])
    b = np.array([[float(x) for x in l.split(",")] for l in lo[1:]])
    assert np.abs(a - b).max() <= 1.0001e-4            # 4 printed decimals: at most one unit in the last place
    assert np.mean(np.array(lg[1:]) == np.array(lo[1:])) > 0.999
    assert all(len(l.split(",")) == 7 and all(len(x.split(".")[1]) == 4 for x in l.split(",")) for l in lg[1:50])


def test_cv_score_matches_oracle(gpu_lib, oracle):
    """fb_cv_score (ScoreAssembler.hpp:15-39 on checkerboard hold-out solves) against the oracle:
    same ADMM step counts of every held-out solve, CV and CV_sd to 1e-9 relative."""
    n, D = 40, 2
    mat = synth(n, D, 5000, 8)
    lams = [0.8, 2.5, 7.0]
    with oracle.create(mat, n) as ho:
        fit = ho.run_signal(0.5, 2.5, 0.0)
        ro = ho.cv_score(0.5, lams, [0.05, 0.0], tol_val=0.2)
    with gpu_lib.create(mat, n) as hg:
        hg.restore(fit["log_signal"], fit["biases"], fit["decay"], fit["exposures"])
        rg = hg.cv_score(0.5, lams, [0.05, 0.0], tol_val=0.2)
        assert hg.kernel_launches() > 0
    assert np.array_equal(rg["steps"], ro["steps"])
    assert np.allclose(rg["cv"], ro["cv"], rtol=1e-9, atol=0)
    assert np.allclose(rg["cv_sd"], ro["cv_sd"], rtol=1e-7, atol=1e-12)
    assert rg["admm_cell_iters"] == ro["admm_cell_iters"]


@pytest.mark.parametrize("n,D,far,maxdiag", [(64, 2, 150000.0, 10), (129, 1, 1e9, 200), (96, 3, 100000.0, 0)])
def test_band_export_matches_oracle(gpu_lib, oracle, n, D, far, maxdiag):
    """fb_band_export (R/chromosome.R:262-330): lam1.sig by an exact device-side median, band triplets,
    sparse log10-signal triplets: integer columns bit-exact, doubles to 1e-12."""
    mat = synth(n, D, 5000, 31)
    with oracle.create(mat, n) as ho:
        fit = ho.run_signal(0.5, 2.5, 0.05)
        eo = ho.band_export(far, maxdiag, 0.05, lam1_final=0.07)
    with gpu_lib.create(mat, n) as hg:
        hg.restore(fit["log_signal"], fit["biases"], fit["decay"], fit["exposures"])
        eg = hg.band_export(far, maxdiag, 0.05, lam1_final=0.07)
        assert hg.kernel_launches() > 0
    assert np.allclose(eg["lam1_sig"], eo["lam1_sig"], rtol=0, atol=1e-12)
    if far >= 1e9:
        assert np.all(eg["lam1_sig"] == 0.07)                 # no far cell: lam1.final
    for k in ("band_bin1", "band_bin2", "band_observed", "nsig", "sig_bin1", "sig_bin2"):
        assert np.array_equal(eg[k], eo[k]), k
    assert np.allclose(eg["band_binless"], eo["band_binless"], rtol=1e-12, atol=0)
    assert np.allclose(eg["sig_log10"], eo["sig_log10"], rtol=1e-10, atol=1e-13)
    assert eg["band_bin1"].size == sum(min(maxdiag + 1, n - i) for i in range(n))


@pytest.mark.parametrize("n,D,seed,kw", [(200, 2, 1, {}), (129, 1, 12, {}), (96, 3, 7, {"nouter": 2}), (150, 2, 5, {"tol_val": 0.02}),
                                          (64, 2, 9, {"compute_patchnos": False})])
def test_early_download_into_pinned_buffers_is_identical(gpu_lib, monkeypatch, n, D, seed, kw):
    """With pinned output buffers fb_run_signal sends phihat / weight / biasmat / decaymat to the host on a second stream
    while the last fused-lasso solves run. Every column equals, bit for bit, the call with pageable buffers (no early
    copies) and the call with the early copies switched off; also when the loop stops at nouter or takes many steps."""
    import torch
    mat = synth(n, D, 5000, seed)
    N = D * n * (n + 1) // 2
    names = ("background", "biasmat", "decaymat", "signal", "binless", "phihat", "weight", "log_signal")

    def pinned():
        t = {k: torch.full((N,), float("nan"), dtype=torch.float64).pin_memory() for k in names}
        t["patchno"] = torch.zeros(N, dtype=torch.int32).pin_memory()
        return t, {k: v.numpy() for k, v in t.items()}

    with gpu_lib.create(mat, n) as h:
        plain = h.run_signal(0.49, 2.5, 0.1, **kw)
        keep1, b1 = pinned()
        early = h.run_signal(0.49, 2.5, 0.1, buffers=b1, **kw)
        again = h.run_signal(0.49, 2.5, 0.1, buffers=b1, **kw)          # on buffers that hold the previous results
        monkeypatch.setenv("FB_NO_EARLY_DOWNLOAD", "1")
        keep2, b2 = pinned()
        late = h.run_signal(0.49, 2.5, 0.1, buffers=b2, **kw)
    assert plain["diagnostics"]["nsteps"] == early["diagnostics"]["nsteps"] == late["diagnostics"]["nsteps"]
    for k in names + ("patchno",):
        assert np.array_equal(plain[k], early[k]), k
        assert np.array_equal(plain[k], again[k]), k
        assert np.array_equal(plain[k], late[k]), k
    del keep1, keep2


def _jittered(n, D, seed):
    """a matrix that is NOT a bin_data() layout: distance varies inside a diagonal, positions differ between datasets"""
    mat = synth(n, D, 5000, seed)
    rng = np.random.default_rng(seed)
    nc = n * (n + 1) // 2
    mat = {k: np.array(v, copy=True) for k, v in mat.items()}
    mat["distance"] = mat["distance"] * np.exp(rng.uniform(-0.02, 0.02, mat["distance"].size))
    for d in range(1, D):
        mat["pos1"][d * nc:(d + 1) * nc] += 7 * d
        mat["pos2"][d * nc:(d + 1) * nc] += 7 * d
    return mat


@pytest.mark.parametrize("n,D,seed", [(33, 2, 3), (129, 1, 5), (200, 2, 1)])
def test_per_cell_binner_fallback_matches_oracle(gpu_lib, oracle, n, D, seed):
    """inputs whose pos / distance are not functions of the bin / diagonal are binned cell by cell
    (util.cpp:56-111, fast_decay.hpp:48-68, fast_bias_mean.hpp:37-68) instead of being refused."""
    mat = _jittered(n, D, seed)
    with oracle.create(mat, n) as ho:
        oo = ho.run_signal(0.49, 2.5, 0.1)
        do = ho.run_difference(1, 0.49, 3.0, 0.1) if D > 1 else None
    with gpu_lib.create(mat, n) as hg:
        og = hg.run_signal(0.49, 2.5, 0.1)
        dg = hg.run_difference(1, 0.49, 3.0, 0.1) if D > 1 else None
    compare_signal(og, oo)
    if D > 1:
        compare_diff(dg, do)


def test_per_cell_binner_equals_closed_form_on_bin_data_layout(gpu_lib, monkeypatch):
    """FB_FORCE_CELL_BINS=1 sends a regular bin_data() matrix through the per-cell binner: same outer steps and ADMM
    iteration counts, outputs equal up to the summation order of the leg sums."""
    n, D = 96, 2
    mat = synth(n, D, 5000, 21)
    with gpu_lib.create(mat, n) as h:
        a = h.run_signal(0.5, 2.5, 0.0)
    monkeypatch.setenv("FB_FORCE_CELL_BINS", "1")
    with gpu_lib.create(mat, n) as h:
        b = h.run_signal(0.5, 2.5, 0.0)
    compare_signal(b, a)


def test_trail_block_sharding_on_two_gpus():
    """Two processes, one per GPU, share every fused-lasso solve over the library's NCCL communicator; the results
    must equal the single-GPU ones bit for bit (tools/shard_check.py). Needs two devices."""
    import subprocess
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs (run on a --gpus 2 box); the host-side logic is covered by tests/test_trail_shard.py")
    port = 29600 + os.getpid() % 300
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                        "--master-addr", "127.0.0.1", "--master-port", str(port),
                        os.path.join(ROOT, "tools", "shard_check.py")], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "SHARD CHECK PASS" in r.stdout, r.stdout[-3000:] + r.stderr[-3000:]


@pytest.mark.timeout(600)
def test_single_process_multi_gpu_is_bit_identical(gpu_lib):
    """fb_create_multi / fb_gfl_solve_multi (multi-GPU chosen by the caller of ONE process, as a stock R session
    would): one host thread and one NCCL communicator per device share every fused-lasso solve; the results must equal the
    single-GPU ones bit for bit. Needs two devices."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs (run on a --gpus 2 box)")
    from binless_b200.dist import _nccl_library_path
    if "FB_NCCL_LIB" not in os.environ and _nccl_library_path():
        os.environ["FB_NCCL_LIB"] = _nccl_library_path()
    n = 300
    rng = np.random.default_rng(7)
    nc = n * (n + 1) // 2
    y = np.repeat(rng.normal(0, 1, nc // 25 + 1), 25)[:nc] + rng.normal(0, 0.4, nc); w = rng.uniform(0, 3, nc)
    a = gfl_solve(gpu_lib, n, y, w, 2.5, 1e-12, 20, 0)
    b = gfl_solve(gpu_lib, n, y, w, 2.5, 1e-12, 20, 0, ngpus=2)
    assert a["steps"] == b["steps"] and a["alpha"] == b["alpha"]
    for k in ("beta", "z", "u", "beta_full"):
        assert np.array_equal(a[k], b[k]), k
    mat = synth(96, 2, 5000, 4)
    with gpu_lib.create(mat, 96) as h1:
        o1 = h1.run_signal(0.5, 2.5, 0.1)
        d1 = h1.run_difference(1, 0.5, 3.0, 0.1)
    with gpu_lib.create(mat, 96, ngpus=2) as h2:
        o2 = h2.run_signal(0.5, 2.5, 0.1)
        d2 = h2.run_difference(1, 0.5, 3.0, 0.1)
    assert list(o1["diagnostics"]["solve_iters"]) == list(o2["diagnostics"]["solve_iters"])
    for k in ("log_signal", "phihat", "weight", "signal", "binless", "background", "biasmat", "decaymat", "patchno"):
        assert np.array_equal(o1[k], o2[k]), k
    for k in ("difference", "deltahat", "weight", "phi_ref", "patchno"):
        assert np.array_equal(d1[k], d2[k]), k


@pytest.mark.parametrize("prox", ["aset", "team"])
def test_gfl_4300_bins_matches_oracle(gpu_lib, oracle, monkeypatch, prox):
    """4300 bins, eight ADMM iterations against the oracle (~10 s of CPU). aset: one chain per batch (two CTAs per
    SM). team: fewer than six chains fit in shared memory, so k_prox keeps the prefix sums in global scratch (15 teams
    per SM), its default above ~4150 bins."""
    monkeypatch.setenv("FB_PROX_ASET", "0" if prox == "team" else "1")     # above 3071 bins k_prox_aset is opt-in
    monkeypatch.setenv("FB_ASET_FROM", "1")
    n = 4300; nc = n * (n + 1) // 2
    rng = np.random.default_rng(43)
    y = np.repeat(rng.normal(0, 1, nc // 30 + 1), 30)[:nc] + rng.normal(0, 0.5, nc)
    w = rng.uniform(0, 3, nc)
    a = gfl_solve(oracle, n, y, w, 2.5, 1e-12, 9, 0)
    b = gfl_solve(gpu_lib, n, y, w, 2.5, 1e-12, 9, 0)
    assert a["steps"] == b["steps"] == 9 and a["alpha"] == b["alpha"]
    for k in ("beta", "z", "u"):
        assert maxabs(a[k], b[k]) < ATOL, k


def test_pipelined_pageable_copies_change_nothing(oracle):
    """Large copies to / from pageable host memory go through pinned staging in 32 MB chunks (fb_model.cu:
    upload_bytes / download_bytes). FB_STAGE_MIN_MB=1 (read once per process, hence the subprocess) forces that path
    at a small size; inputs, outputs and labels must equal the direct path bit for bit."""
    import subprocess
    code = (
        "import sys, zlib, numpy as np; sys.path.insert(0, %r)\n"
        "from binless_b200 import cuda_lib\n"
        "from binless_b200.synth import synth\n"
        "mat = synth(700, 3, 5000, 5)\n"
        "with cuda_lib().create(mat, 700) as h:\n"
        "    o = h.run_signal(0.5, 2.5, 0.05)\n"
        "print(' '.join(str(zlib.crc32(np.ascontiguousarray(o[k]).tobytes())) for k in\n"
        "      ('log_signal', 'phihat', 'weight', 'biasmat', 'decaymat', 'signal', 'binless', 'background', 'patchno')))\n" % ROOT)
    outs = []
    # staged + the early copies of phihat / weight / biasmat / decaymat by a second host thread (fb_run_signal), then without
    for env_extra in ({}, {"FB_STAGE_MIN_MB": "1"}, {"FB_STAGE_MIN_MB": "1", "FB_NO_EARLY_DOWNLOAD": "1"}):
        env = dict(os.environ); env.pop("FB_STAGE_MIN_MB", None); env.update(env_extra)
        r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, env=env, timeout=300)
        assert r.returncode == 0, r.stderr[-2000:]
        outs.append(r.stdout.strip().splitlines()[-1])
    assert outs[0] == outs[1] == outs[2]
