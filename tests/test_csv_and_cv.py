"""CPU: the gzip csvfile writer of the product (host code, no device) against the oracle's plain-text restatement of
csvfile.hpp:30-117, and the oracle's checkerboard CV score against an independent recomputation from binless_eval_cv."""
import gzip
import os

import numpy as np
import pytest

from binless_b200 import _abi
from binless_b200.synth import synth

HEADER = '"phihat","weight","biasmat","decaymat","signal","binless","background"'


@pytest.fixture(scope="module")
def product_lib():
    from binless_b200 import build as b
    return _abi.FastBinlessLib(b.build(), "fb_")


@pytest.mark.parametrize("nrows", [0, 1, 17, 70001])
def test_gzip_csv_equals_reference_text(product_lib, oracle_port, tmp_path, nrows):
    rng = np.random.default_rng(nrows)
    cols = [rng.normal(0, 10.0 ** rng.integers(-3, 6), nrows) for _ in range(7)]
    if nrows > 3:
        cols[0][0] = 0.0; cols[1][1] = -0.00004; cols[2][2] = 1e15; cols[3][3] = 0.00005
    fz, fp, fo = str(tmp_path / "a.csv.gz"), str(tmp_path / "a_plain.csv"), str(tmp_path / "o.csv")
    product_lib.csv_write(fz, HEADER, cols, gzip=True)
    product_lib.csv_write(fp, HEADER, cols, gzip=False)
    oracle_port.csv_write(fo, HEADER, cols, gzip=False)
    want = open(fo, "rb").read()
    assert gzip.open(fz, "rb").read() == want               # what read.csv() sees (R/chromosome.R:238)
    assert open(fp, "rb").read() == want                    # the USE_BOOST = 0 variant
    raw = open(fz, "rb").read()
    assert raw[:4] == b"\x1f\x8b\x08\x00"                    # one gzip member, deflate, no optional fields
    if nrows > 1000:
        assert len(raw) < len(want)


def test_csv_values_print_like_printf(product_lib, oracle_port, tmp_path):
    """The product prints "%.4f" with its own exact integer arithmetic (fb_csv.cpp: fmt_fixed4), the oracle with the C
    library like the reference's ostream. Ties at the fifth decimal (k / 32: 10^4 / 32 = 312.5), values next to them,
    subnormals, signed zeros, huge values and non-finite ones must give the same characters."""
    rng = np.random.default_rng(7)
    ties = rng.integers(0, 100000, 20000) + (2 * rng.integers(0, 16, 20000) + 1) / 32.0
    near = rng.integers(0, 10 ** 9, 20000) / 10000.0 + 0.00005
    mags = np.ldexp(1.0 + rng.random(20000), rng.integers(-60, 70, 20000))
    bits = rng.integers(0, 2 ** 63 - 1, 20000, dtype=np.int64).view(np.float64)
    bits = bits[np.isfinite(bits)]
    special = np.array([0.0, -0.0, 5e-324, -5e-324, 2.2250738585072014e-308, 0.03125, -0.03125, 0.09375, 0.99995, 9999.99995,
                        1125899906842623.9, 1125899906842624.0, 1e15, 1e16, 1e22, 1e300, -1e300, np.inf, -np.inf, np.nan,
                        0.5, 1.00005, 2.00005, 3.00015, 123456.78125])
    x = np.concatenate([special, ties, -ties, near, np.nextafter(near, 0), np.nextafter(near, 1e30), mags, -mags, bits])
    fp, fo = str(tmp_path / "p.csv"), str(tmp_path / "o.csv")
    product_lib.csv_write(fp, '"x"', [x], gzip=False)
    oracle_port.csv_write(fo, '"x"', [x], gzip=False)
    got, want = open(fp, "rb").read().split(b"\n"), open(fo, "rb").read().split(b"\n")
    assert len(got) == len(want) == x.size + 2
    bad = [(float(x[k - 1]), got[k], want[k]) for k in range(1, x.size + 1) if got[k] != want[k]]
    assert not bad, bad[:5]
    # and against Python's own correctly rounded formatting, for the finite ones
    for k in range(special.size, special.size + 2000):
        assert got[k + 1] == ("%.4f" % x[k]).encode()


def test_gzip_csv_many_slabs(product_lib, oracle_port, tmp_path):
    """more rows than one formatting slab (2^20): the pieces of all slabs form ONE deflate stream"""
    n = (1 << 20) + 12345
    cols = [np.arange(n) * 0.25, np.linspace(-1, 1, n)]
    fz, fo = str(tmp_path / "b.csv.gz"), str(tmp_path / "b.csv")
    product_lib.csv_write(fz, '"a","b"', cols, gzip=True)
    oracle_port.csv_write(fo, '"a","b"', cols, gzip=False)
    import zlib
    raw = open(fz, "rb").read()
    d = zlib.decompressobj(16 + zlib.MAX_WBITS)
    text = d.decompress(raw)
    assert d.eof and d.unused_data == b""                   # a single member, nothing after it
    assert text == open(fo, "rb").read()


def test_cv_score_oracle_matches_eval_cv_recomputation(oracle):
    """fbo_cv_score against the same quantity assembled in numpy from binless_eval_cv (fast_binless.cpp:186-233) run with
    group = 1 and group = 2 on a fresh handle each (cold start), ScoreAssembler.hpp:24-36."""
    n, lam2, lam1 = 24, 1.7, 0.05
    mat = synth(n, 1, 5000, 5)
    with oracle.create(mat, n) as h:
        fit = h.run_signal(0.5, lam2, 0.0)
        got = h.cv_score(0.5, [lam2, 4.0], lam1, tol_val=0.2)
        y = fit["phihat"]; w = fit["weight"]
    i, j = np.triu_indices(n)
    grp = ((i + 1) + (j + 1)) % 2
    # the IRLS (phihat, w) of the fitted state, not held out: cv_score computes them from the state like sweep_lambda
    with oracle.create(mat, n) as h:
        h.restore(fit["log_signal"], fit["biases"], fit["decay"], fit["exposures"])
        full = h.eval_cv(0.5, [lam2], lam1, group=0, tol_val=0.2)[0]
    yf, wf = full["phihat"], full["weight"]
    beta_cv = np.zeros(yf.size)
    for g in (0, 1):
        with oracle.create(mat, n) as h:
            h.restore(fit["log_signal"], fit["biases"], fit["decay"], fit["exposures"])
            r = h.eval_cv(0.5, [lam2], lam1, group=g + 1, tol_val=0.2)[0]
        beta_cv[grp == g] = r["log_signal"][grp == g]
    soft = np.where(beta_cv > 0, np.maximum(0, beta_cv - lam1), np.minimum(0, beta_cv + lam1))
    chisq = wf * (yf - soft) ** 2 + lam1 ** 2
    Ng = np.array([np.sum(grp == g) for g in (0, 1)], dtype=float)
    cvg = np.array([chisq[grp == g].sum() / Ng[g] for g in (0, 1)])
    cv = (Ng * cvg).sum() / Ng.sum()
    sd = np.sqrt((Ng * (cvg - cv) ** 2).sum() / (Ng.sum() - 1))
    assert abs(got["cv"][0, 0] - cv) < 1e-9 * (1 + abs(cv))
    assert abs(got["cv_sd"][0, 0] - sd) < 1e-9 * (1 + abs(sd))
    assert got["cv"].shape == (2, 1) and np.all(got["steps"] > 1)
