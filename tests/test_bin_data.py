"""bin_data() on the input side (/root/reference/binless/R/preprocess.R:114-146, R/common.R:115-132).

CPU: the oracle's loop restatement (fbo_create_binned) against the statement-by-statement numpy restatement
(oracle/bin_data_numpy.py) -- integer work, bit-exact -- and hand-checked edge cases. The reference's R code
cannot run here (no R), so this row is "parity unpinned" by the reference itself, like the driver rows.
GPU (-m gpu): fb_create_binned builds the same matrix on the device, bit for bit, and fast_binless on that handle equals
fast_binless on the host-binned matrix."""
import os
import sys

import numpy as np
import pytest

sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle"))
from bin_data_numpy import bin_reads_numpy  # noqa: E402

from binless_b200.synth import synth_reads

COLS = ("name", "bin1", "pos1", "bin2", "pos2", "observed", "nobs", "distance")
CASES = [(7, 1, 5000, 1), (33, 2, 2000, 2), (100, 3, 5000, 3), (64, 2, 777, 4), (1, 1, 5000, 5), (2, 2, 101, 6)]


def same_mat(h, want):
    assert h.nbins == want["nbins"]
    for k in COLS:
        assert np.array_equal(h.cols[k], want[k]), k


@pytest.mark.parametrize("n,D,res,seed", CASES)
def test_oracle_binning_matches_numpy_restatement(oracle_port, n, D, res, seed):
    r = synth_reads(n, D, res, seed)
    want = bin_reads_numpy(r, res, D)
    assert want["nbins"] == n
    with oracle_port.create_binned(r, res, D) as h:
        same_mat(h, want)


def test_hand_checked_edges(oracle_port):
    """3 bins of 10 from cut sites 101..131: borders 100,110,120,130 (131+1=132 < 140: the 4th bin is incomplete)."""
    z = np.zeros(4, dtype=np.int32)
    r = dict(cut_name=[1, 1, 1, 1, 1], cut_pos=[101, 105, 112, 130, 131],
             rd_name=[1, 1, 1, 1], rd_pos1=[101, 105, 112, 101], rd_pos2=[105, 112, 130, 131],
             contact_close=[1, 0, 2, 5], contact_far=[1, 0, 0, 0], contact_up=z, contact_down=z)
    with oracle_port.create_binned(r, 10, 1) as h:
        assert h.nbins == 3
        m = h.cols
        # cells (1,1) (1,2) (1,3) (2,2) (2,3) (3,3); cuts per bin: 2, 1, 1 (130 closes the last bin, 131 is NA)
        assert m["observed"].tolist() == [2, 0, 1, 0, 2, 0]          # read 4 ends beyond the last border: dropped; corner +1
        assert m["nobs"].tolist() == [8, 8, 0, 0, 4, 0]              # 4*2*2/2 on the diagonal; (1,2) has a row with zero counts: nobs kept
        assert m["pos1"].tolist() == [105, 105, 105, 115, 115, 125] and m["pos2"].tolist() == [105, 115, 125, 115, 125, 125]
        assert m["distance"].tolist() == [5.0, 15.0, 25.0, 5.0, 15.0, 5.0]


def check_bin_golden(lib, name):
    import os, sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    sys.path.insert(0, os.path.join(root, "tools"))
    from make_golden import BIN_CASES
    c = BIN_CASES[name]
    g = np.load(os.path.join(root, "tests", "golden", name + ".npz"))
    r = synth_reads(c["n"], c["D"], c["res"], c["seed"])
    assert sum(int(r[k].astype(np.int64).sum()) for k in r) == int(g["reads_sum"]), "synthetic read generator drifted"
    with lib.create_binned(r, c["res"], c["D"]) as h:
        assert h.nbins == int(g["nbins"])
        for k in ("observed", "nobs", "pos1", "pos2", "distance"):
            assert np.array_equal(h.cols[k], g[k]), k


@pytest.mark.parametrize("name", ["bin_data_n33_d2", "bin_data_n64_d2_res777"])
def test_oracle_binning_reproduces_golden(oracle_port, name):
    check_bin_golden(oracle_port, name)


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["bin_data_n33_d2", "bin_data_n64_d2_res777"])
def test_device_binning_reproduces_golden(gpu_lib, name):
    check_bin_golden(gpu_lib, name)


def test_errors(oracle_port):
    from binless_b200._abi import FastBinlessError
    r = synth_reads(5, 1, 1000, 1)
    with pytest.raises(FastBinlessError):
        oracle_port.create_binned(r, 10 ** 9, 1)                      # no complete bin


@pytest.mark.gpu
@pytest.mark.parametrize("n,D,res,seed", CASES[:5] + [(600, 2, 5000, 7)])
def test_device_binning_is_bit_exact(gpu_lib, oracle, n, D, res, seed):
    r = synth_reads(n, D, res, seed, reads_per_dataset=20000 if n < 500 else 400000)
    want = bin_reads_numpy(r, res, D)
    with gpu_lib.create_binned(r, res, D) as hg:
        assert hg.kernel_launches() > 0
        same_mat(hg, want)
        if n >= 7:
            og = hg.run_signal(0.5, 2.5, 0.0)
            with gpu_lib.create(want, n) as hh:                       # host-binned matrix through fb_create
                oh = hh.run_signal(0.5, 2.5, 0.0)
            for k in ("log_signal", "phihat", "weight", "binless", "patchno"):
                assert np.array_equal(og[k], oh[k]), k
            if n <= 100:
                with oracle.create_binned(r, res, D) as ho:
                    oo = ho.run_signal(0.5, 2.5, 0.0)
                assert list(og["diagnostics"]["solve_iters"]) == list(oo["diagnostics"]["solve_iters"])
                assert float(np.max(np.abs(og["log_signal"] - oo["log_signal"]))) < 1e-8
