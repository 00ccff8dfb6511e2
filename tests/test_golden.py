"""CPU: the oracle against the committed golden vectors (tests/golden/*.npz, written by tools/make_golden.py from
the oracle linked with the reference's own C solver)."""
import glob
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tools"))
from make_golden import BIN_CASES, CASES  # noqa: E402
from binless_b200.synth import synth  # noqa: E402


def check_against_golden(lib, name, exact):
    c = CASES[name]
    g = np.load(os.path.join(ROOT, "tests", "golden", name + ".npz"))
    mat = synth(c["n"], c["D"], 5000, c["seed"])
    assert int(mat["observed"].astype(np.int64).sum()) == int(g["observed_sum"]), "synthetic generator drifted"
    close = (lambda a, b: np.array_equal(a, b)) if exact else \
        (lambda a, b: float(np.max(np.abs(a - b) / (1 + np.abs(b)))) < 1e-6)
    with lib.create(mat, c["n"]) as h:
        out = h.run_signal(c["alpha"], c["lam2"], c["lam1"], **c["kw"])
        d = out["diagnostics"]
        assert d["nsteps"] == int(g["nsteps"])
        assert list(d["solve_iters"]) == list(g["solve_iters"])
        assert close(out["log_signal"], g["log_signal"])
        assert close(out["binless"], g["binless"])
        assert close(out["biases"]["estimate"], g["bias"])
        assert close(out["decay"]["beta"], g["decay_beta"])
        assert close(np.array([out["decay"]["lambda_"], out["decay"]["mean"]]), np.array([g["decay_lambda"], g["decay_mean"]]))
        assert close(out["exposures"]["estimate"], g["expo"])
        assert np.array_equal(out["patchno"], g["patchno"])     # labels are integers: always bit-exact
        for k in ("phihat", "weight", "background"):
            if k in g.files:
                assert close(out[k], g[k]), k
        if c["D"] > 1:
            df = h.run_difference(1, c["alpha"], 3.38, 0.14, maxdiag=c["kw"].get("maxdiag", 0))
            assert list(df["diagnostics"]["solve_iters"]) == list(g["diff_iters"])
            assert close(df["difference"], g["diff_difference"])
            assert np.array_equal(df["patchno"], g["diff_patchno"])


@pytest.mark.parametrize("name", sorted(CASES))
def test_oracle_reproduces_golden(oracle_port, name):
    check_against_golden(oracle_port, name, exact=True)


def test_golden_files_present():
    files = {os.path.basename(f)[:-4] for f in glob.glob(os.path.join(ROOT, "tests", "golden", "*.npz"))}
    assert files == set(CASES) | set(BIN_CASES)
