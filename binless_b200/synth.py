"""Deterministic synthetic Hi-C matrices in the layout ``binless:::bin_data`` produces.

The reference's bundled example inputs are git-LFS stubs, so every test and
benchmark runs on matrices from this generator. The
column layout follows ``bin_data`` (/root/reference/binless/R/preprocess.R:104-148)
and ``add_bin_bounds_and_distance`` (/root/reference/binless/R/common.R:115-132):
dense upper triangle per dataset, sorted by (name, bin1, bin2); ``begin_k = 1 +
(k-1)*res``, ``pos = begin + res/2``, ``distance = begin2 - begin1 + res/2``;
``nobs = 4*c_i*c_j`` (halved on the diagonal) with ``c`` the cut sites per bin;
``observed[(1, nbins)] += 1`` so the decay fit is not degenerate.
"""
from __future__ import annotations

import numpy as np

__all__ = ["synth", "synth_reads", "synth_cts", "tri_indices", "CONFIGS"]

# the named workloads of BASELINE.json / SURVEY.md section 8
CONFIGS = {
    "C1": dict(nbins=200, ndatasets=2, res=5000, seed=1),
    "C2": dict(nbins=2400, ndatasets=4, res=5000, seed=2),
    "C4": dict(nbins=25000, ndatasets=1, res=2000, seed=4),
    "C5": dict(nbins=50000, ndatasets=1, res=5000, seed=5),
}


def tri_indices(nbins: int):
    """0-based (row, col) of every cell of the packed upper triangle, row-major."""
    i, j = np.triu_indices(nbins)
    return i.astype(np.int32), j.astype(np.int32)


def _smooth3(x):
    y = x.copy()
    y[1:-1] = (x[:-2] + x[1:-1] + x[2:]) / 3.0
    return y


def synth(nbins: int, ndatasets: int, res: int = 5000, seed: int = 1, alpha_true: float = 0.5,
          depth: float | None = None):
    """Return a dict of numpy columns (the ``mat`` data frame of ``fast_binless``).

    Keys: name, bin1, pos1, bin2, pos2 (int32), distance (float64), observed, nobs
    (int32), plus ``nbins``. ``depth`` overrides the per-dataset exposure draw
    (log of expected reads per observable at distance ``res``).
    """
    rng = np.random.Generator(np.random.PCG64(seed))
    bi, bj = tri_indices(nbins)
    ncells = bi.size
    begin = 1 + np.arange(nbins, dtype=np.int64) * res
    pos = (begin + res // 2).astype(np.int32)
    dist = ((bj - bi).astype(np.float64)) * res + res / 2.0

    # cut sites per bin -> number of observables
    cuts = rng.poisson(res / 256.0, size=nbins).astype(np.int64)
    dead = rng.random(nbins) < 0.02
    cuts[dead] = 0
    nobs = 4 * cuts[bi] * cuts[bj]
    nobs[bi == bj] //= 2

    # genomic biases, decay, signal
    iota = _smooth3(rng.normal(0.0, 0.2, size=nbins))
    ld = np.log(dist / res * 2.0)  # = ln((d + 1/2) * 2) >= 0
    decay = -1.0 * ld - 0.3 * ld * ld / np.log(max(nbins, 3))
    # TAD squares tiled along the diagonal
    tad_of_bin = np.zeros(nbins, dtype=np.int64)
    k = 0
    t = 0
    while k < nbins:
        size = int(rng.integers(20, 101))
        tad_of_bin[k:k + size] = t
        k += size
        t += 1
    ntads = t
    same_tad = tad_of_bin[bi] == tad_of_bin[bj]
    # Gaussian loops: one per 200 bins
    nloops = max(1, nbins // 200)
    loops = []
    for _ in range(nloops):
        a = int(rng.integers(0, max(1, nbins - 30)))
        b = min(nbins - 1, a + int(rng.integers(10, 60)))
        loops.append((a, b))

    cols = {k_: [] for k_ in ("name", "bin1", "pos1", "bin2", "pos2", "distance", "observed", "nobs")}
    for d in range(ndatasets):
        e_d = rng.uniform(-2.0, -1.0) if depth is None else depth
        tad_on = np.ones(ntads, dtype=bool)
        if d > 0:
            tad_on = rng.random(ntads) >= 0.10  # flip 10 % of TADs so differences exist
        phi = np.where(same_tad & tad_on[tad_of_bin[bi]], 0.7, 0.0)
        for (a, b) in loops:
            r2 = (bi - a).astype(np.float64) ** 2 + (bj - b).astype(np.float64) ** 2
            phi = phi + 1.5 * np.exp(-r2 / (2.0 * 3.0 ** 2))
        eta = e_d + iota[bi] + iota[bj] + decay + phi
        mean = nobs * np.exp(eta)
        size = alpha_true * np.maximum(nobs, 1)
        # NegBin(mean, size) as a gamma-Poisson mixture
        lam = rng.gamma(shape=size, scale=np.where(nobs > 0, mean / size, 0.0))
        obs = rng.poisson(lam).astype(np.int64)
        obs[nobs == 0] = 0
        obs[nbins - 1] += 1  # cell (1, nbins): preprocess.R:140
        cols["name"].append(np.full(ncells, d + 1, dtype=np.int32))
        cols["bin1"].append(bi + 1)
        cols["bin2"].append(bj + 1)
        cols["pos1"].append(pos[bi])
        cols["pos2"].append(pos[bj])
        cols["distance"].append(dist)
        cols["observed"].append(np.minimum(obs, 2 ** 31 - 1).astype(np.int32))
        cols["nobs"].append(np.minimum(nobs, 2 ** 31 - 1).astype(np.int32))
    mat = {k_: np.ascontiguousarray(np.concatenate(v)) for k_, v in cols.items()}
    mat["nbins"] = nbins
    return mat


def synth_reads(nbins: int, ndatasets: int, res: int = 5000, seed: int = 1, reads_per_dataset: int = 20000):
    """A CSnorm-like object before binning: cut sites (obj@biases: name, pos) and cut-site pairs with their four
    contact counts (obj@counts: name, pos1 < pos2, contact.close/far/up/down), /root/reference/binless/R/preprocess.R:114-146.
    Positions start at an arbitrary offset; a few cut sites lie in the incomplete last bin, a few pairs cross it."""
    rng = np.random.Generator(np.random.PCG64(seed))
    start = int(rng.integers(1000, 100000))
    span = nbins * res + int(rng.integers(0, res - 1))           # the tail beyond the last complete bin is discarded
    cut_name, cut_pos = [], []
    for d in range(ndatasets):
        m = int(nbins * max(2.0, res / 256.0))
        pos = np.unique(rng.integers(start, start + span, size=m))
        if d == 0:
            pos = np.unique(np.concatenate([pos, [start, start + span - 1]]))
        cut_name.append(np.full(pos.size, d + 1)); cut_pos.append(pos)
    rd = {k: [] for k in ("rd_name", "rd_pos1", "rd_pos2", "contact_close", "contact_far", "contact_up", "contact_down")}
    for d in range(ndatasets):
        pos = cut_pos[d]
        i = rng.integers(0, pos.size, size=reads_per_dataset)
        off = np.minimum(rng.geometric(0.02, size=reads_per_dataset), pos.size - 1)
        j = np.minimum(i + off, pos.size - 1)
        rd["rd_name"].append(np.full(reads_per_dataset, d + 1)); rd["rd_pos1"].append(pos[i]); rd["rd_pos2"].append(pos[j])
        for k in ("contact_close", "contact_far", "contact_up", "contact_down"):
            rd[k].append(rng.poisson(0.4, size=reads_per_dataset))
    out = {"cut_name": np.concatenate(cut_name).astype(np.int32), "cut_pos": np.concatenate(cut_pos).astype(np.int32)}
    out.update({k: np.concatenate(v).astype(np.int32) for k, v in rd.items()})
    return out


def synth_cts(nbins: int, seed: int = 1, rows_per_cell: float = 2.5, dispersion: float = 2.0, shift: float = 0.0):
    """The `cts` data frame and `metadata` list of the optimized path (what binless:::csnorm_predict_binned hands to
    wgfl_signal_BIC, /root/reference/binless/src/perf_iteration_signal.cpp:22-30; columns read by cts_to_mat.cpp:43-50):
    several rows per bin pair (one per cut-site category), 1-based bins, sorted by (bin1, bin2); a few cells have no row.
    `shift` moves the planted signal so that a second call gives the `ref` / `cts` pair of wgfl_diff_BIC."""
    rng = np.random.Generator(np.random.PCG64(seed))
    bi, bj = tri_indices(nbins)
    k = rng.poisson(rows_per_cell, size=bi.size)
    k[rng.random(bi.size) < 0.03] = 0
    b1 = np.repeat(bi + 1, k).astype(np.int32); b2 = np.repeat(bj + 1, k).astype(np.int32)
    n = b1.size
    d = (b2 - b1).astype(np.float64)
    log_decay = -0.9 * np.log1p(d) + rng.normal(0, 0.02, n)
    tad = (np.arange(nbins) // max(4, nbins // 5))
    phi_cell = np.where(tad[bi] == tad[bj], 0.8 + shift, 0.0)
    a, b = nbins // 3, (2 * nbins) // 3
    phi_cell = phi_cell + (1.2 - shift) * np.exp(-((bi - a) ** 2 + (bj - b) ** 2) / 6.0)
    phi = np.repeat(phi_cell, k)
    lmu_nosig = 1.0 + log_decay + rng.normal(0, 0.1, n)
    nobs = rng.integers(1, 40, size=n).astype(np.float64)
    mu = np.exp(lmu_nosig + phi)
    lam = rng.gamma(shape=dispersion, scale=mu / dispersion)
    count = rng.poisson(lam).astype(np.float64)
    diag_grp = np.floor(np.log2(np.arange(nbins) + 1)).astype(np.int32)
    return {"bin1": b1, "bin2": b2, "count": count, "lmu_nosig": lmu_nosig, "nobs": nobs, "log_decay": log_decay,
            "metadata": {"bad_rows": np.array([max(1, nbins // 2)], dtype=np.int32),
                         "bad_diags": np.array([nbins - 1], dtype=np.int32), "diag_grp": diag_grp},
            "nbins": nbins, "dispersion": dispersion}
