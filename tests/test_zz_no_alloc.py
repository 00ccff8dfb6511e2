"""fb_far_median_noalloc: the far-cell median of fb_band_export for callers that cannot allocate (scratch in a static
device array, nothing allocated, kernels only enqueued). The entry point has no counterpart in the reference; the check
is against the existing entry point, fb_band_export's lam1_sig, to the tolerance of the band-export test (1e-12).

This file is named so that pytest collects it LAST."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from binless_b200._abi import FastBinlessError  # noqa: E402
from binless_b200.synth import synth  # noqa: E402

ATOL = 1e-12          # test_band_export_matches_oracle's bound on lam1_sig
MAX_NBINS = 65535     # FB_NOALLOC_MAX_NBINS


def far_from_diag(mat, n, far_cutoff):
    """first diagonal whose distance exceeds far.cutoff (bin_data layout: the distance is a function of bin2 - bin1)"""
    dist = np.asarray(mat["distance"][:n])          # cells (1, 1..n) of the first dataset
    assert np.all(np.diff(dist) > 0)
    idx = np.nonzero(dist > far_cutoff)[0]
    return int(idx[0]) if idx.size else n


def test_oracle_side_of_the_entry_point_matches_band_export(oracle):
    """CPU: the oracle exports the same entry point on host pointers; it agrees with the oracle's band_export, which pins
    the argument convention (far_from_diag, one dataset's triangle) the GPU test below relies on."""
    n, D = 64, 2
    mat = synth(n, D, 5000, 31)
    nc = n * (n + 1) // 2
    with oracle.create(mat, n) as ho:
        fit = ho.run_signal(0.5, 2.5, 0.05, nouter=3)
        for far, lam1 in ((150000.0, 0.05), (100000.0, 0.0)):
            eo = ho.band_export(far, 10, lam1, lam1_final=0.07)
            ff = far_from_diag(mat, n, far)
            assert 0 < ff < n
            for d in range(D):
                x = np.ascontiguousarray(fit["log_signal"][d * nc:(d + 1) * nc])
                med = np.full(1, 123.0)
                oracle.far_median_noalloc(n, x.ctypes.data, lam1, ff, med.ctypes.data)
                assert abs(med[0] - eo["lam1_sig"][d]) <= ATOL
    med = np.full(1, 123.0)
    oracle.far_median_noalloc(n, x.ctypes.data, 0.0, n, med.ctypes.data)
    assert np.isnan(med[0])


@pytest.mark.gpu
def test_far_median_noalloc_matches_band_export(gpu_lib):
    """GPU, through the C ABI: two calls in a row (then more) on caller-owned device buffers equal fb_band_export's
    lam1_sig; the shared scratch keeps nothing from one call to the next; device memory in use does not change."""
    import torch
    n, D = 129, 2
    mat = synth(n, D, 5000, 31)
    nc = n * (n + 1) // 2
    cases = [(150000.0, 0.05), (100000.0, 0.0), (40000.0, 0.2)]        # (far.cutoff, lam1): odd and even far-cell counts
    with gpu_lib.create(mat, n) as hg:
        fit = hg.run_signal(0.5, 2.5, 0.05)
        want = [hg.band_export(far, 10, lam1, lam1_final=0.07)["lam1_sig"] for far, lam1 in cases]
    ffs = [far_from_diag(mat, n, far) for far, _ in cases]
    assert {((n - f) * (n - f + 1) // 2) & 1 for f in ffs} == {0, 1}
    dev = torch.device("cuda:0")
    x = torch.from_numpy(np.ascontiguousarray(fit["log_signal"]).reshape(D, nc)).to(dev)
    out = torch.full((len(cases), D), 123.0, dtype=torch.float64, device=dev)
    torch.cuda.synchronize()

    def enqueue_all(stream=None):
        for c, (far, lam1) in enumerate(cases):
            for d in range(D):                                       # back to back, no synchronisation in between
                gpu_lib.far_median_noalloc(n, x[d].data_ptr(), lam1, ffs[c], out[c, d:d + 1].data_ptr(), stream)

    enqueue_all()                                                    # first calls: kernels load, set-up runs
    torch.cuda.synchronize()
    got = out.cpu().numpy()
    assert np.all(np.isfinite(got))
    assert np.allclose(got, np.asarray(want), rtol=0, atol=ATOL), (got, want)

    free0 = torch.cuda.mem_get_info()[0]
    out.fill_(-7.0)
    enqueue_all()                                                    # again, on what the calls above left in the scratch
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info()[0] == free0                     # nothing was allocated (or freed)
    assert np.allclose(out.cpu().numpy(), np.asarray(want), rtol=0, atol=ATOL)

    side = torch.cuda.Stream()                                       # a caller's own stream
    out.fill_(-7.0)
    torch.cuda.synchronize()
    enqueue_all(side.cuda_stream)
    side.synchronize()
    assert np.allclose(out.cpu().numpy(), np.asarray(want), rtol=0, atol=ATOL)

    # no far cell: NaN (R's NA), which fb_band_export replaces by lam1_final
    gpu_lib.far_median_noalloc(n, x[0].data_ptr(), 0.0, n, out[0, 0:1].data_ptr())
    torch.cuda.synchronize()
    assert np.isnan(out[0, 0].item())
    # and a real call right after it
    gpu_lib.far_median_noalloc(n, x[1].data_ptr(), cases[0][1], ffs[0], out[0, 0:1].data_ptr())
    torch.cuda.synchronize()
    assert abs(out[0, 0].item() - want[0][1]) <= ATOL


@pytest.mark.gpu
def test_far_median_noalloc_rejects_sizes_beyond_the_scratch(gpu_lib):
    import torch
    buf = torch.zeros(8, dtype=torch.float64, device="cuda:0")
    with pytest.raises(FastBinlessError) as e:
        gpu_lib.far_median_noalloc(MAX_NBINS + 1, buf.data_ptr(), 0.0, 1, buf[0:1].data_ptr())
    assert e.value.code == 1 and "FB_NOALLOC_MAX_NBINS" in str(e.value)
    with pytest.raises(FastBinlessError):
        gpu_lib.far_median_noalloc(4, 0, 0.0, 1, buf[0:1].data_ptr())
    with pytest.raises(FastBinlessError):
        gpu_lib.far_median_noalloc(0, buf.data_ptr(), 0.0, 1, buf[0:1].data_ptr())
    torch.cuda.synchronize()
    assert float(buf.abs().sum().item()) == 0.0                      # refused calls touch nothing
