"""Host-side sharding of independent fast-binless work units over ranks (one process per GPU).

The units of this path that are independent of each other are whole ``fast_binless`` calls on different loci
(what ``chromosome_binless`` does with chunks of a chromosome, /root/reference/binless/R/chromosome.R:106-148)
and cold-start fused-lasso solves at different lambda2 (the lambda/BIC grid). Datasets
*within* one call are not independent in the reference: all ``flos`` copies share one GFL state
(GFLLibrary.hpp:59), so dataset d warm-starts from dataset d-1; sharding them would change the iterates.
No collective is on the data path; ``torch.distributed`` only gathers the small per-unit summaries.
"""
from __future__ import annotations

from typing import Callable, List, Sequence


def partition_units(n_units: int, world_size: int, rank: int) -> List[int]:
    """Round-robin assignment: unit u goes to rank u % world_size (units are sorted by decreasing cost)."""
    if world_size < 1 or not (0 <= rank < world_size):
        raise ValueError("bad rank/world_size")
    return list(range(rank, n_units, world_size))


def run_sharded(units: Sequence, fn: Callable, rank: int = 0, world_size: int = 1, gather: bool = True):
    """Run ``fn(unit)`` on this rank's share; with ``gather`` return all results in unit order on every rank."""
    mine = partition_units(len(units), world_size, rank)
    local = [(u, fn(units[u])) for u in mine]
    if not gather or world_size == 1:
        return [r for _, r in sorted(local, key=lambda p: p[0])] if world_size == 1 else local
    import torch.distributed as dist
    bucket = [None] * world_size
    dist.all_gather_object(bucket, local)
    flat = [p for part in bucket for p in part]
    return [r for _, r in sorted(flat, key=lambda p: p[0])]


# ---- trail-block sharding of ONE fused-lasso solve (include/fastbinless.h: fb_comm_init, fb_shard_plan) -------------
# Host-side mirror of the geometry the CUDA library uses (binless_b200/csrc/fb_gfl.cu: GflSolver::setup_shard); the
# tests check the library's plan against it and run the exchange pattern over gloo.
def band_width(nbins: int, maxdiag: int = 0) -> int:
    """Width of the solved band: the trapeze variant iff 1 < maxdiag <= (nbins+1)/2 (GFLLibrary.hpp:21-27)."""
    md = min(int(maxdiag), nbins)
    return nbins if (md > (nbins + 1) // 2 or md <= 1) else md


def crossing_cells(nbins: int, md: int, row_begin, col_begin, sender: int, receiver: int, direction: int):
    """Cells (i, j) whose consensus term z - u the sender owns and the receiver's chains need, in packing order.

    direction 0: the sender's ROW chains (rows of its row block) crossing the receiver's column block;
    direction 1: the sender's COLUMN chains crossing the receiver's row block."""
    import numpy as np
    out = []
    if direction == 0:
        for i in range(row_begin[sender], row_begin[sender + 1]):
            lo = max(col_begin[receiver], i); hi = min(col_begin[receiver + 1], i + md, nbins)
            out += [(i, j) for j in range(lo, hi)]
    else:
        for j in range(col_begin[sender], col_begin[sender + 1]):
            lo = max(row_begin[receiver], j - md + 1, 0); hi = min(row_begin[receiver + 1], j + 1)
            out += [(i, j) for i in range(lo, hi)]
    return np.array(out, dtype=np.int64).reshape(-1, 2)
