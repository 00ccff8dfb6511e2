"""BASELINE.json config 3: 64-point lambda2 / BIC sweep of the signal fused lasso on the fitted C2 model (2400 bins x
4 datasets), the lambda grid dealt round-robin to the ranks (independent cold-start solves, no data-path collective).

    python tools/c3_sweep.py [--nlam 64] [--nbins 2400]                                            # 1 GPU
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29519 tools/c3_sweep.py
Prints one JSON line on rank 0: wall time of the sweep (max over ranks), M cell*iter/s over all ranks, best lambda2."""
import argparse, json, os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
from binless_b200 import cuda_lib
from binless_b200.shard import partition_units
from binless_b200.sweep import lambda_grid
from binless_b200.synth import CONFIGS, synth

ap = argparse.ArgumentParser()
ap.add_argument("--nlam", type=int, default=64)
ap.add_argument("--nbins", type=int, default=CONFIGS["C2"]["nbins"])
args = ap.parse_args()
rank = int(os.environ.get("RANK", 0)); world = int(os.environ.get("WORLD_SIZE", 1)); local = int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
lib = cuda_lib()
cfg = CONFIGS["C2"]; n, D = args.nbins, cfg["ndatasets"]
mat = synth(n, D, cfg["res"], cfg["seed"])
slim = {k: (None if k in ("name", "bin1", "bin2") else v) for k, v in mat.items()}
grid = lambda_grid(args.nlam)
mine = partition_units(grid.size, world, rank)
with lib.create(slim, n) as h:
    fit = h.run_signal(0.5, 2.5, 0.0, columns="none")           # every rank fits the same model (deterministic)
    h.sweep_lambda(0.5, grid[mine][:1], 0.0)                     # warm-up
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    part = h.sweep_lambda(0.5, grid[mine], 0.0)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
tt = torch.tensor([dt, float(part["admm_cell_iters"]), part["ms_solver"]], dtype=torch.float64, device="cuda")
rows = [(int(u), part["bic"][i].tolist(), part["steps"][i].tolist()) for i, u in enumerate(mine)]
if world > 1:
    tmax = tt.clone(); dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
    tsum = tt.clone(); dist.all_reduce(tsum, op=dist.ReduceOp.SUM)
    bucket = [None] * world
    dist.all_gather_object(bucket, rows)
    rows = sorted(r for b in bucket for r in b)
else:
    tmax = tsum = tt
if rank == 0:
    bic = np.array([r[1] for r in rows]); steps = np.array([r[2] for r in rows])
    print(json.dumps({"workload": f"C3: {args.nlam}-point lambda2/BIC sweep, {n} bins x {D} datasets, cold-start solves",
                      "n_gpus": world, "seconds": float(tmax[0]), "Mcell_iter_per_s": float(tsum[1]) / float(tmax[0]) / 1e6,
                      "admm_iterations_total": int(steps.sum() - steps.size), "solves": int(steps.size),
                      "best_lam2": grid[np.argmin(bic, axis=0)].tolist(), "solver_ms_max_rank": float(tmax[2])}), flush=True)
if world > 1:
    dist.destroy_process_group()
