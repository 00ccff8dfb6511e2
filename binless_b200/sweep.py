"""lambda2 / BIC sweep of the signal fused lasso, sharded over ranks (BASELINE.json config 3).

Every (lambda2, dataset) point is an independent cold-start solve (fb_sweep_lambda, include/fastbinless.h), so the
lambda2 grid is dealt round-robin to the ranks (one process per GPU); the only communication is the gather of the
small per-point rows. ``lib`` is the CUDA library on a GPU box; tests pass the CPU oracle, which exports the same ABI.
"""
from __future__ import annotations

import numpy as np

from .shard import partition_units


def lambda_grid(n: int = 64, lo: float = -1.0, hi: float = 2.0) -> np.ndarray:
    """10^linspace(-1, 2, 64): the grid of config 3."""
    return 10.0 ** np.linspace(lo, hi, n)


def bic_sweep(lib, mat, nbins, fit, alpha, lam2_grid, lam1=0.0, tol_val=0.2, maxdiag=0, rank=0, world_size=1):
    """BIC of every lambda2 in ``lam2_grid`` for every dataset, from the fitted state ``fit`` (the value returned by
    fast_binless). Returns ``{"lam2", "bic", "chisq", "dof", "steps", "best_lam2"}`` with (nlam, ndatasets) arrays,
    identical on every rank."""
    lam2_grid = np.asarray(lam2_grid, dtype=np.float64)
    mine = partition_units(lam2_grid.size, world_size, rank)
    with lib.create(mat, nbins) as h:
        h.restore(fit["log_signal"], fit["biases"], fit["decay"], fit["exposures"])
        part = h.sweep_lambda(alpha, lam2_grid[mine], lam1, tol_val=tol_val, maxdiag=maxdiag) if mine else None
        D = h.ndatasets
    rows = [(u, {k: part[k][i] for k in ("bic", "chisq", "dof", "steps")}) for i, u in enumerate(mine)]
    if world_size > 1:
        import torch.distributed as dist
        bucket = [None] * world_size
        dist.all_gather_object(bucket, rows)
        rows = [r for b in bucket for r in b]
    rows.sort(key=lambda r: r[0])
    out = {"lam2": lam2_grid}
    for k, dt in (("bic", np.float64), ("chisq", np.float64), ("dof", np.int32), ("steps", np.int32)):
        out[k] = np.array([r[1][k] for r in rows], dtype=dt).reshape(lam2_grid.size, D)
    out["best_lam2"] = lam2_grid[np.argmin(out["bic"], axis=0)] if lam2_grid.size else np.full(D, np.nan)
    return out
