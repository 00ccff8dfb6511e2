"""Chromosome-scale end-to-end run of fast_binless on ONE GPU (BASELINE.json configs 4 / 5), with the synthetic input
generated on the device by torch (plumbing only; the numpy generator of binless_b200/synth.py takes minutes at this size)
in the same bin_data layout, then handed to the C ABI as host arrays like any caller's.

    python tools/chrom_full.py --nbins 25000 --res 2000                  # C4: full triangle, 3.1e8 cells
    python tools/chrom_full.py --nbins 50000 --res 5000 --maxdiag 2000   # C5: 1.25e9 cells, lasso on the 2000-wide band
Checks size-independent properties of the result (shift_signal leaves min(log_signal) == 0, finite legs, decay
non-increasing beyond free_decay) and prints one JSON line with the timings."""
import argparse, json, os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from binless_b200 import cuda_lib

ap = argparse.ArgumentParser()
ap.add_argument("--nbins", type=int, default=25000)
ap.add_argument("--res", type=int, default=2000)
ap.add_argument("--maxdiag", type=int, default=0)
ap.add_argument("--seed", type=int, default=4)
ap.add_argument("--rows-per-block", type=int, default=128)
ap.add_argument("--nouter", type=int, default=25)
ap.add_argument("--sharded", action="store_true", help="under torchrun: the ranks share every fused-lasso solve (trail blocks)")
args = ap.parse_args()
n, res = args.nbins, args.res
nc = n * (n + 1) // 2
rank = int(os.environ.get("RANK", 0)); world = int(os.environ.get("WORLD_SIZE", 1)); local = int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
if args.sharded:
    import torch.distributed as dist
    from binless_b200.dist import init_comm
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
dev = torch.device("cuda", local)
torch.manual_seed(args.seed); torch.cuda.manual_seed_all(args.seed)      # _standard_gamma draws from the default generator
g = torch.Generator(device=dev); g.manual_seed(args.seed)
t0 = time.time()
cuts = torch.poisson(torch.full((n,), res / 256.0, device=dev), generator=g)
cuts[torch.rand(n, device=dev, generator=g) < 0.02] = 0
iota = torch.randn(n, device=dev, generator=g, dtype=torch.float64) * 0.2
tad = torch.cumsum((torch.rand(n, device=dev, generator=g) < 1.0 / 60.0).long(), 0)      # TADs of ~60 bins
pos = (1 + torch.arange(n, device=dev, dtype=torch.int64) * res + res // 2).to(torch.int32)
cols = {k: np.empty(nc, dtype=np.int32) for k in ("pos1", "pos2", "observed", "nobs")}
cols["distance"] = np.empty(nc, dtype=np.float64)
off = 0
for i0 in range(0, n, args.rows_per_block):
    i1 = min(n, i0 + args.rows_per_block)
    ii = torch.arange(i0, i1, device=dev).view(-1, 1); jj = torch.arange(i0, n, device=dev).view(1, -1)
    m = jj >= ii
    bi = ii.expand(-1, jj.numel())[m]; bj = jj.expand(ii.numel(), -1)[m]
    d = (bj - bi).to(torch.float64)
    nobs = 4.0 * cuts[bi] * cuts[bj]
    nobs = torch.where(bi == bj, torch.floor(nobs / 2), nobs)
    ld = torch.log(2.0 * d + 1.0)
    eta = -1.5 + iota[bi] + iota[bj] - ld - 0.3 * ld * ld / np.log(n) + 0.7 * (tad[bi] == tad[bj])
    mean = nobs.to(torch.float64) * torch.exp(eta)
    size = 0.5 * torch.clamp(nobs, min=1.0).to(torch.float64)
    lam = torch._standard_gamma(size) * (mean / size)
    obs = torch.poisson(lam.to(torch.float32), generator=g)
    if i0 == 0:
        obs[n - 1] += 1                                   # cell (1, nbins): preprocess.R:140
    k = bi.numel()
    cols["observed"][off:off + k] = obs.to(torch.int32).cpu().numpy()
    cols["nobs"][off:off + k] = nobs.to(torch.int32).cpu().numpy()
    cols["pos1"][off:off + k] = pos[bi].cpu().numpy(); cols["pos2"][off:off + k] = pos[bj].cpu().numpy()
    cols["distance"][off:off + k] = (d * res + res / 2.0).cpu().numpy()
    off += k
assert off == nc
del bi, bj, d, nobs, ld, eta, mean, size, lam, obs, m
torch.cuda.empty_cache()
t_gen = time.time() - t0
if rank == 0:
    print(f"generated {nc} cells in {t_gen:.1f} s, {cols['observed'].sum()} reads", flush=True)
mat = dict(cols, name=None, bin1=None, bin2=None)
lib = cuda_lib()
import zlib
in_crc = zlib.crc32(cols["observed"].tobytes()) ^ zlib.crc32(cols["nobs"].tobytes())
if args.sharded:
    t_in = torch.tensor([in_crc], dtype=torch.int64, device=dev)
    lo = t_in.clone(); dist.all_reduce(lo, op=dist.ReduceOp.MIN); hi = t_in.clone(); dist.all_reduce(hi, op=dist.ReduceOp.MAX)
    if lo.item() != hi.item():
        raise SystemExit("the ranks generated different inputs: the sharded solve needs identical calls on every rank")
    init_comm(lib)      # every rank generated the same matrix (same seed, same GPU model) and makes the same calls
    dist.barrier()
t0 = time.time()
h = lib.create(mat, n)
t_create = time.time() - t0
t0 = time.time()
out = h.run_signal(0.5, 2.5, 0.0, columns=False, maxdiag=args.maxdiag, nouter=args.nouter,
                   compute_patchnos=(args.maxdiag == 0))
t_run = time.time() - t0
dg = out["diagnostics"]
ls = out["log_signal"]
ok = bool(np.isfinite(ls).all() and ls.min() == 0.0 and np.isfinite(out["biases"]["estimate"]).all()
          and np.all(np.diff(out["decay"]["beta"][30:]) <= 1e-9))
h.close()
if args.sharded:
    import zlib
    sig = torch.tensor([zlib.crc32(ls.tobytes())], dtype=torch.int64, device=dev)
    lo = sig.clone(); dist.all_reduce(lo, op=dist.ReduceOp.MIN); hi = sig.clone(); dist.all_reduce(hi, op=dist.ReduceOp.MAX)
    ok = ok and bool(lo.item() == hi.item())      # every rank holds the same log_signal, bit for bit
if rank == 0:
  print(json.dumps({"workload": f"fast_binless, {n} bins x 1 dataset ({nc} cells), res {res}, maxdiag {args.maxdiag}, lam2 2.5",
                  "properties_ok": ok, "outer_steps": int(dg["nsteps"]), "converged": bool(dg["converged"]),
                  "admm_iters_per_solve": [int(x) for x in dg["solve_iters"]], "seconds_create": t_create,
                  "seconds_run": t_run, "ms_device": dg["ms_total"], "ms_solver": dg["ms_solver"],
                  "Mcell_iter_per_s": dg["admm_cell_iters"] / t_run / 1e6,
                  "log_signal_max": float(ls.max()), "frac_signal_nonzero": float((ls > 0).mean()),
                  "n_gpus": world if args.sharded else 1, "input_crc32": in_crc, "log_signal_crc32": __import__("zlib").crc32(ls.tobytes())}), flush=True)
if args.sharded:
    dist.destroy_process_group()
sys.exit(0 if ok else 1)
