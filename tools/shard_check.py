"""Multi-GPU check of the trail-block sharded ADMM (run under torchrun, one process per GPU):

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 \
        tools/shard_check.py [--big N]

Every rank first solves on its own GPU alone (communicator off), then the ranks share the same solves; results must be
identical BIT FOR BIT (same iterates: the exchange carries exactly the terms the single-GPU kernel reads locally).
"""
import argparse, os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
from binless_b200 import cuda_lib
from binless_b200._abi import gfl_solve
from binless_b200.dist import init_comm, shutdown_comm
from binless_b200.synth import synth

ap = argparse.ArgumentParser()
ap.add_argument("--big", type=int, default=0, help="also time a fixed number of iterations at this many bins")
ap.add_argument("--iters", type=int, default=17)
args = ap.parse_args()

rank = int(os.environ.get("RANK", 0)); world = int(os.environ.get("WORLD_SIZE", 1)); local = int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
lib = cuda_lib()
ok = True

def say(*a):
    if rank == 0:
        print(*a, flush=True)

def same(a, b, keys, label):
    global ok
    for k in keys:
        if not np.array_equal(np.asarray(a[k]), np.asarray(b[k])):
            ok = False
            d = np.abs(np.asarray(a[k], dtype=np.float64) - np.asarray(b[k], dtype=np.float64)).max()
            print(f"[rank {rank}] MISMATCH {label}.{k}: max |d| = {d:.3e}", flush=True)

SOLVES = [(2, 1.3, 1e-3, 30000, 0), (5, 1.3, 1e-3, 30000, 0), (33, 1.3, 1e-3, 30000, 0), (100, 1.3, 1e-3, 30000, 0),
          (200, 2.5, 1e-12, 120, 0), (257, 2.5, 1e-4, 30000, 0), (120, 2.0, 1e-3, 30000, 30), (300, 50.0, 1e-4, 30000, 0)]
single = []
for (n, lam, conv, ms, md) in SOLVES:
    rng = np.random.default_rng(n + ms)
    nc = n * (n + 1) // 2
    y = rng.normal(size=nc); w = rng.uniform(0, 3, size=nc); w[rng.random(nc) < 0.2] = 0
    s1 = gfl_solve(lib, n, y, w, lam, conv, ms, md)
    s2 = gfl_solve(lib, n, y * 1.1, w, lam * 1.5, conv, ms, md, state=s1)      # warm start from exported state
    single.append((y, w, s1, s2))
mat = synth(96, 3, 5000, 5)
with lib.create(mat, 96) as h:
    sig1 = h.run_signal(0.5, [2.5, 1.5, 3.0], [0.1, 0.0, 0.2])
    dif1 = h.run_difference(2, 0.5, 3.38, 0.14)
with lib.create(mat, 96) as h:
    band1 = h.run_signal(0.5, 2.5, 0.0, maxdiag=30)

init_comm(lib)
say(f"communicator up: world {world}; shard plan n=300: rows {lib.shard_plan(300, world)[0].tolist()} cols {lib.shard_plan(300, world)[1].tolist()}")
for (n, lam, conv, ms, md), (y, w, s1, s2) in zip(SOLVES, single):
    t1 = gfl_solve(lib, n, y, w, lam, conv, ms, md)
    t2 = gfl_solve(lib, n, y * 1.1, w, lam * 1.5, conv, ms, md, state=t1)
    same(s1, t1, ("steps", "alpha", "beta", "z", "u", "beta_full"), f"solve n={n} md={md}")
    same(s2, t2, ("steps", "alpha", "beta", "z", "u", "beta_full"), f"warm solve n={n} md={md}")
    say(f"solve n={n} lam={lam} md={md}: steps {t1['steps']} / warm {t2['steps']}  sharded == single")
with lib.create(mat, 96) as h:
    sig2 = h.run_signal(0.5, [2.5, 1.5, 3.0], [0.1, 0.0, 0.2])
    dif2 = h.run_difference(2, 0.5, 3.38, 0.14)
with lib.create(mat, 96) as h:
    band2 = h.run_signal(0.5, 2.5, 0.0, maxdiag=30)
cols = ("log_signal", "phihat", "weight", "signal", "binless", "background", "biasmat", "decaymat", "patchno")
same(sig1, sig2, cols, "fast_binless"); same(band1, band2, cols, "fast_binless band")
same(dif1, dif2, ("difference", "deltahat", "weight", "phi_ref", "patchno"), "difference")
if list(sig1["diagnostics"]["solve_iters"]) != list(sig2["diagnostics"]["solve_iters"]):
    ok = False; print(f"[rank {rank}] iteration counts differ", flush=True)
say(f"fast_binless 96 bins x 3: outer steps {sig2['diagnostics']['nsteps']}, ADMM iterations {list(sig2['diagnostics']['solve_iters'])}")

if args.big:
    n = args.big; nc = n * (n + 1) // 2
    rng = np.random.default_rng(4)
    y = np.repeat(rng.normal(0, 1, nc // 40 + 1), 40)[:nc] + rng.normal(0, 0.5, nc)
    w = rng.uniform(0, 3, nc)
    import ctypes as C
    from binless_b200._abi import fb_gfl_state, _dp
    def run(ms):
        beta = np.zeros(nc)
        st = fb_gfl_state(beta=_dp(beta), z=None, u=None, alpha=5.0, steps=0)
        dist.barrier(); torch.cuda.synchronize()
        t = time.time()
        lib._check(lib._gfl_solve(n, 0, _dp(y), _dp(w), 2.5, 1e-12, ms, 0, C.byref(st), None), None)
        torch.cuda.synchronize(); dist.barrier()
        return time.time() - t, st.steps, beta
    run(2)
    ta, sa, _ = run(2)
    tb, sb, beta_sh = run(args.iters)
    per = (tb - ta) / max(1, sb - sa)
    say(f"BIG n={n} cells={nc} world={world}: {per*1e3:.3f} ms per ADMM iteration ({nc/per/1e6:.0f} Mcell*iter/s, "
        f"{104*nc/per/1e9:.0f} GB/s algorithmic over {world} GPUs)")
    shutdown_comm(lib)
    run(2)
    ta, sa, _ = run(2)
    tb, sb, beta_1 = run(args.iters)
    per1 = (tb - ta) / max(1, sb - sa)
    say(f"BIG n={n} single GPU: {per1*1e3:.3f} ms per ADMM iteration  => speed-up {per1/per:.2f}x on {world} GPUs")
    if not np.array_equal(beta_sh, beta_1):
        ok = False; print(f"[rank {rank}] BIG beta mismatch {np.abs(beta_sh-beta_1).max():.3e}", flush=True)
else:
    shutdown_comm(lib)

flag = torch.tensor([0 if ok else 1], device="cuda")
dist.all_reduce(flag)
say("SHARD CHECK", "PASS" if flag.item() == 0 else "FAIL")
dist.destroy_process_group()
sys.exit(0 if flag.item() == 0 else 1)
