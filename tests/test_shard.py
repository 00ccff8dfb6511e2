"""CPU: host-side sharding of independent units over ranks (gloo, world_size 2). The compute stand-in is the CPU
oracle (allowed in tests); on the GPU box the same code path runs the CUDA library, one process per GPU."""
import os
import sys

import numpy as np
import pytest
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

from binless_b200.shard import partition_units, run_sharded  # noqa: E402


def test_partition_covers_all_units_once():
    for n, ws in [(0, 1), (5, 1), (5, 2), (64, 8), (7, 8)]:
        seen = sorted(u for r in range(ws) for u in partition_units(n, ws, r))
        assert seen == list(range(n))
        sizes = [len(partition_units(n, ws, r)) for r in range(ws)]
        assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        partition_units(4, 2, 2)


def _worker(rank, world_size, port, q):
    sys.path.insert(0, ROOT)
    import torch.distributed as dist
    from oracle.loader import oracle_lib
    from binless_b200._abi import gfl_solve
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world_size)
    ora = oracle_lib("port")
    n = 16; nc = n * (n + 1) // 2
    rng = np.random.default_rng(0)
    y = rng.normal(size=nc); w = rng.uniform(0.1, 2, size=nc)
    lams = list(10 ** np.linspace(-1, 1, 7))          # a small lambda2 grid: cold-start solves are independent
    fn = lambda lam: (float(lam), int(gfl_solve(ora, n, y, w, lam, 1e-2)["steps"]))
    res = run_sharded(lams, fn, rank, world_size)
    if rank == 0:
        q.put(res)
    dist.barrier()
    dist.destroy_process_group()


def test_lambda_grid_sharded_over_two_ranks_matches_serial(oracle_port):
    from binless_b200._abi import gfl_solve
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + os.getpid() % 2000
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = q.get(timeout=120)
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    n = 16; nc = n * (n + 1) // 2
    rng = np.random.default_rng(0)
    y = rng.normal(size=nc); w = rng.uniform(0.1, 2, size=nc)
    lams = list(10 ** np.linspace(-1, 1, 7))
    serial = [(float(l), int(gfl_solve(oracle_port, n, y, w, l, 1e-2)["steps"])) for l in lams]
    assert res == serial
