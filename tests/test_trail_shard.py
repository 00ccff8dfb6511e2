"""CPU: host-side logic of the trail-block sharded ADMM (one fused-lasso solve shared by the ranks).

* fb_shard_plan (product library, host only) against the oracle's cell-counting restatement and its properties;
* world_size-2 gloo run of the exchange pattern: every rank forms z - u for the chains it owns, the crossing cells
  travel with send/recv in the library's packing order, and the consensus sum r = (z-u)_row + (z-u)_col that the beta
  update (update_beta_weight, gfl_graph_fl.c:241-264) needs comes out identical to the unsharded one on every cell
  of every rank's chains. The bit-exact GPU run of the same scheme is tools/shard_check.py (2 B200)."""
import os
import sys

import numpy as np
import pytest
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

from binless_b200 import _abi  # noqa: E402
from binless_b200.shard import band_width, crossing_cells  # noqa: E402


@pytest.fixture(scope="module")
def product_lib():
    from binless_b200 import build as b
    return _abi.FastBinlessLib(b.build(), "fb_")


def cells_in_rows(n, md, a, b):
    return sum(min(md, n - i) for i in range(a, b))


def cells_in_cols(n, md, a, b):
    return sum(min(md, j + 1) for j in range(a, b))


@pytest.mark.parametrize("n,maxdiag", [(2, 0), (3, 0), (17, 0), (200, 0), (2400, 0), (131, 40), (96, 30), (64, 33), (2400, 300)])
@pytest.mark.parametrize("world", [1, 2, 3, 4, 8])
def test_shard_plan_matches_oracle_and_balances(product_lib, oracle_port, n, maxdiag, world):
    rb, cb = product_lib.shard_plan(n, world, maxdiag)
    ro, co = oracle_port.shard_plan(n, world, maxdiag)
    assert np.array_equal(rb, ro) and np.array_equal(cb, co)
    md = band_width(n, maxdiag)
    total = cells_in_rows(n, md, 0, n)
    assert rb[0] == cb[0] == 0 and rb[-1] == cb[-1] == n
    assert np.all(np.diff(rb) >= 0) and np.all(np.diff(cb) >= 0)
    for g in range(world):      # each block within one line of the ideal share
        assert abs(cells_in_rows(n, md, rb[g], rb[g + 1]) - total / world) <= md + 1
        assert abs(cells_in_cols(n, md, cb[g], cb[g + 1]) - total / world) <= md + 1


def test_shard_plan_needs_no_device(product_lib):
    rb, cb = product_lib.shard_plan(25000, 8)          # BASELINE config 4
    assert rb.tolist()[0] == 0 and rb.tolist()[-1] == 25000 and len(rb) == 9
    assert rb[1] < 25000 // 8 < cb[1]                   # long rows first, short columns first


def test_crossing_cells_partition_the_needed_terms():
    """Every cell of a rank's chains gets its row term and its column term exactly once: locally or from one peer."""
    for n, maxdiag, world in [(23, 0, 2), (23, 0, 3), (40, 12, 4), (9, 0, 8)]:
        md = band_width(n, maxdiag)
        rb = np.zeros(world + 1, int); cb = np.zeros(world + 1, int)
        lib = _abi.FastBinlessLib(os.path.join(ROOT, "oracle", "_build", "libfbo_port.so"), "fbo_")
        rb, cb = lib.shard_plan(n, world, maxdiag)
        row_owner = np.searchsorted(rb, np.arange(n), side="right") - 1
        col_owner = np.searchsorted(cb, np.arange(n), side="right") - 1
        for me in range(world):
            got_row, got_col = {}, {}
            for p in range(world):
                if p == me:
                    continue
                for (i, j) in crossing_cells(n, md, rb, cb, p, me, 0):
                    assert (i, j) not in got_row; got_row[(i, j)] = p
                for (i, j) in crossing_cells(n, md, rb, cb, p, me, 1):
                    assert (i, j) not in got_col; got_col[(i, j)] = p
            for i in range(n):
                for j in range(i, min(n, i + md)):
                    mine_row, mine_col = row_owner[i] == me, col_owner[j] == me
                    if not (mine_row or mine_col):
                        continue                          # not on any chain of this rank
                    # the row term is local or received from the row owner, never both; same for the column term
                    assert ((i, j) in got_row) == (not mine_row) and ((i, j) in got_col) == (not mine_col)
                    if not mine_row:
                        assert got_row[(i, j)] == row_owner[i]
                    if not mine_col:
                        assert got_col[(i, j)] == col_owner[j]


def _worker(rank, world, port, n, maxdiag, q):
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist
    from binless_b200 import _abi as abi
    from binless_b200.shard import band_width as bw, crossing_cells as cc
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    lib = abi.FastBinlessLib(os.path.join(ROOT, "binless_b200", "libfastbinless_cuda.so"), "fb_")
    md = bw(n, maxdiag)
    rb, cb = lib.shard_plan(n, world, maxdiag)
    rng = np.random.default_rng(7)                        # same global state on every rank ...
    zr, ur, zc, uc = (rng.normal(size=(n, n)) for _ in range(4))
    mine_r = np.zeros((n, n), bool); mine_r[rb[rank]:rb[rank + 1], :] = True
    mine_c = np.zeros((n, n), bool); mine_c[:, cb[rank]:cb[rank + 1]] = True
    t_row = np.where(mine_r, zr - ur, np.nan)             # ... of which a rank only knows the chains it owns
    t_col = np.where(mine_c, zc - uc, np.nan)
    reqs, recv = [], []
    for p in range(world):
        if p == rank:
            continue
        s0 = cc(n, md, rb, cb, rank, p, 0); s1 = cc(n, md, rb, cb, rank, p, 1)
        sbuf = torch.from_numpy(np.concatenate([t_row[s0[:, 0], s0[:, 1]], t_col[s1[:, 0], s1[:, 1]]]))
        r0 = cc(n, md, rb, cb, p, rank, 0); r1 = cc(n, md, rb, cb, p, rank, 1)
        rbuf = torch.empty(len(r0) + len(r1), dtype=torch.float64)
        if sbuf.numel():
            reqs.append(dist.isend(sbuf, p))
        if rbuf.numel():
            reqs.append(dist.irecv(rbuf, p))
        recv.append((r0, r1, rbuf))
    for r in reqs:
        r.wait()
    for r0, r1, rbuf in recv:
        t_row[r0[:, 0], r0[:, 1]] = rbuf[:len(r0)].numpy()
        t_col[r1[:, 0], r1[:, 1]] = rbuf[len(r0):].numpy()
    band = np.triu(np.ones((n, n), bool)) & ~np.triu(np.ones((n, n), bool), md)
    need = band & (mine_r | mine_c)
    want = (zr - ur) + (zc - uc)
    got = t_row + t_col
    ok = bool(np.all(np.isfinite(got[need])) and np.array_equal(got[need], want[need]))
    q.put((rank, ok, int(need.sum())))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("n,maxdiag", [(37, 0), (48, 14)])
def test_consensus_exchange_over_two_gloo_ranks(product_lib, n, maxdiag):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 31500 + (os.getpid() + n) % 2000
    procs = [ctx.Process(target=_worker, args=(r, 2, port, n, maxdiag, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in range(2))
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    assert [r[1] for r in res] == [True, True]
    md = band_width(n, maxdiag)
    assert sum(r[2] for r in res) >= cells_in_rows(n, md, 0, n)    # every band cell is on some rank's chains
