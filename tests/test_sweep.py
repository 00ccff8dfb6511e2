"""CPU: the lambda2 / BIC sweep (config 3) on the oracle -- BIC identities, and the sharded sweep over two gloo ranks
equals the serial one. On the GPU box the same module (binless_b200/sweep.py) drives the CUDA library, one process
per GPU; tests/test_gpu_parity.py::test_sweep_lambda_matches_oracle checks the CUDA entry point itself."""
import os
import sys

import numpy as np
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

from binless_b200.sweep import bic_sweep, lambda_grid  # noqa: E402
from binless_b200.synth import synth  # noqa: E402

N_BINS, N_DS, SEED = 24, 2, 5


def _fit(ora, mat):
    with ora.create(mat, N_BINS) as h:
        return h.run_signal(0.5, 2.5, 0.0)


def test_bic_identities(oracle_port):
    mat = synth(N_BINS, N_DS, 5000, SEED)
    fit = _fit(oracle_port, mat)
    grid = lambda_grid(5)
    r = bic_sweep(oracle_port, mat, N_BINS, fit, 0.5, grid, lam1=0.05)
    assert r["bic"].shape == (5, N_DS)
    ncells = N_BINS * (N_BINS + 1) // 2
    nobs = np.asarray(mat["nobs"]).reshape(N_DS, ncells).sum(axis=1)
    assert np.allclose(r["bic"], r["chisq"] + np.log(nobs)[None, :] * r["dof"], rtol=1e-12)
    assert (r["dof"] >= 0).all() and (r["dof"] <= ncells).all()
    assert (r["dof"][-1] <= r["dof"][0]).all()            # a heavier penalty never leaves more selected patches here
    assert (r["chisq"] >= ncells * 0.05 ** 2 - 1e-9).all()  # the lambda1^2 prior term alone
    assert r["best_lam2"].shape == (N_DS,)
    # an empty grid is legal
    e = bic_sweep(oracle_port, mat, N_BINS, fit, 0.5, [], lam1=0.05)
    assert e["bic"].shape == (0, N_DS)


def _worker(rank, world_size, port, q):
    sys.path.insert(0, ROOT)
    import torch.distributed as dist
    from oracle.loader import oracle_lib
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world_size)
    ora = oracle_lib("port")
    mat = synth(N_BINS, N_DS, 5000, SEED)
    fit = _fit(ora, mat)
    r = bic_sweep(ora, mat, N_BINS, fit, 0.5, lambda_grid(5), lam1=0.05, rank=rank, world_size=world_size)
    q.put((rank, r["bic"], r["dof"], r["steps"]))
    dist.barrier()
    dist.destroy_process_group()


def test_sharded_sweep_matches_serial(oracle_port):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 31500 + os.getpid() % 2000
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = [q.get(timeout=180) for _ in range(2)]
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    mat = synth(N_BINS, N_DS, 5000, SEED)
    serial = bic_sweep(oracle_port, mat, N_BINS, _fit(oracle_port, mat), 0.5, lambda_grid(5), lam1=0.05)
    for _, bic, dof, steps in got:                       # every rank holds the complete, identical table
        assert np.array_equal(bic, serial["bic"]) and np.array_equal(dof, serial["dof"])
        assert np.array_equal(steps, serial["steps"])
