"""TEST INFRASTRUCTURE (oracle): bin_data() for a CSnorm object restated with numpy, statement by statement.

Follows /root/reference/binless/R/preprocess.R:114-146 and add_bin_bounds_and_distance, R/common.R:115-132. Only tests/,
tools/make_golden.py and the timing aid tools/bin_data_time.py import this; the product path is fb_create_binned
(binless_b200/csrc/fb_bin.cu). Parity unpinned by the reference itself: R is not available in this image, so the pin is
this literal restatement plus the hand-checked case in tests/test_bin_data.py."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from binless_b200.synth import tri_indices  # noqa: E402


def bin_reads_numpy(reads, res: int, ndatasets: int):
    """Returns the ``mat`` columns bin_data() builds (plus ``nbins``)."""
    pos = reads["cut_pos"].astype(np.int64)
    borders = np.arange(pos.min() - 1, pos.max() + 1 + 1, res)         # seq(min(pos)-1, max(pos)+1, resolution)
    nb = borders.size - 1                                               # the last incomplete bin is discarded
    def cut(p):                                                         # cut(.., right=F, include.lowest=T); -1 = NA
        p = np.asarray(p, dtype=np.int64)
        b = np.searchsorted(borders, p, side="right") - 1
        b[p == borders[-1]] = nb - 1
        b[(p < borders[0]) | (p > borders[-1])] = -1
        return b
    bi, bj = tri_indices(nb)
    nc = bi.size
    cols = {k: [] for k in ("name", "bin1", "pos1", "bin2", "pos2", "distance", "observed", "nobs")}
    for d in range(ndatasets):
        sel = reads["cut_name"] == d + 1
        b = cut(pos[sel]); b = b[b >= 0]
        ncut = np.bincount(b, minlength=nb).astype(np.int64)
        nobs = 4 * ncut[bi] * ncut[bj]
        nobs[bi == bj] //= 2                                            # as.integer(nobs / 2)
        rs = reads["rd_name"] == d + 1
        b1 = cut(reads["rd_pos1"][rs]); b2 = cut(reads["rd_pos2"][rs])
        tot = (reads["contact_close"][rs].astype(np.int64) + reads["contact_far"][rs] + reads["contact_up"][rs] + reads["contact_down"][rs])
        ok = (b1 >= 0) & (b2 >= 0) & (b1 <= b2)
        cell = b1[ok] * nb - b1[ok] * (b1[ok] - 1) // 2 + (b2[ok] - b1[ok])
        observed = np.bincount(cell, weights=tot[ok], minlength=nc).astype(np.int64)
        present = np.bincount(cell, minlength=nc) > 0                   # rows of poscounts; absent -> is.na(observed)
        nobs[~present] = 0; observed[~present] = 0
        observed[nb - 1] += 1
        begin = borders[:-1]; end = borders[1:]
        p = ((end + begin) // 2).astype(np.int32)                       # as.integer((end + begin) / 2)
        cols["name"].append(np.full(nc, d + 1, dtype=np.int32)); cols["bin1"].append(bi + 1); cols["bin2"].append(bj + 1)
        cols["pos1"].append(p[bi]); cols["pos2"].append(p[bj])
        cols["distance"].append((begin[bj] - begin[bi]) + res / 2.0)
        cols["observed"].append(observed.astype(np.int32)); cols["nobs"].append(nobs.astype(np.int32))
    mat = {k: np.ascontiguousarray(np.concatenate(v)) for k, v in cols.items()}
    mat["nbins"] = int(nb)
    return mat
