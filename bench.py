#!/usr/bin/env python
"""bench.py -- fast-binless hot path on B200: throughput, end-to-end, roofline, CPU baseline.

    python bench.py --gpus N --steps K --warmup W            # the CUDA library (this repo)
    python bench.py --impl reference --gpus N --steps K ...  # the reference's CPU path (oracle/_ref), rank 0 only

Workload (config.workload): BASELINE.json configs[1] -- synthetic chr7 12 Mb locus at 5 kb: 2400 bins
(2 881 200 cells) x 4 datasets, alpha 0.5, lam2 2.5, reference defaults otherwise. One "step" is one complete
fast_binless call on that batch. With N > 1 every rank runs its own locus of the same shape (independent units,
no data-path collective; weak scaling) and `value` is the sum over ranks divided by the slowest rank's time.

Timing: whole calls between barrier + cudaDeviceSynchronize pairs (max over ranks); `device_ms_per_step` is the same
call measured with CUDA events on the library's own stream (rank 0), which torch.cuda.Event cannot see.

metric = M cell*iter/s: ADMM cell-iterations of the whole call (sum over fused-lasso solves of cells x iterations,
identical to the oracle's counts) divided by the time of the whole call.
  value : inputs resident in HBM, results left in HBM (fb_run_signal on an existing handle)
  e2e   : the public API with host buffers: fb_create (validation + H2D) + fb_run_signal + every output column D2H
e2e_pageable : the same call with pageable numpy buffers and every input column passed (what the Rcpp shim supplies)
c3_sweep, c4_trailblock, c5_chromosome : the lambda-grid, trail-block and whole-chromosome configurations of BASELINE.json,
         sharded over the N ranks (strong scaling), with the bit-identity of a sharded solve against the single-GPU one
roofline: the ADMM iteration of the fused-lasso solve (the dominant device work), HBM bound, 104 algorithmic
bytes per cell per iteration, timed with CUDA events on the solver's stream inside the timed steps.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

from binless_b200.synth import CONFIGS, synth  # noqa: E402

WORKLOAD = "C2: synthetic chr7 12 Mb @ 5 kb, 2400 bins x 4 datasets, alpha=0.5 lam2=2.5 lam1=0 nouter=25 tol_val=0.2 bg_steps=5"
METRIC = "fast_binless_Mcell_iter_per_s"
UNIT = "Mcell*iter/s"
ALPHA, LAM2, LAM1 = 0.5, 2.5, 0.0
ADMM_BYTES_PER_CELL_ITER = 104  # DESIGN.md section 4 / SURVEY.md section 8(d)


def ncu_traffic():
    """DRAM bytes (read + write) of one ADMM iteration from the committed `ncu --set full` captures (profiles/): the mean over
    an early, a middle and a late iteration of the first solve (k_beta + the chain-prox kernel that ran + k_ctl)."""
    tot = []
    for it in (2, 9, 17):
        path = os.path.join(ROOT, "profiles", f"r2_admm_it{it}_ncu_full_summary.json")
        try:
            rows = json.load(open(path))
            tot.append(sum((r["dram_read_MB"] + r["dram_write_MB"]) * 1e6 for r in rows))
        except Exception:
            pass
    return int(sum(tot) / len(tot)) if tot else None


def measured_peak():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        try:
            return float(json.load(open(path))["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.rows, self.proc = [], None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={index}", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], None, set()
        for r in self.rows:
            try:
                sm.append(float(r[0])); mx = float(r[1])
            except Exception:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(sm)}


def pinned_like(a):
    import torch
    t = torch.from_numpy(np.ascontiguousarray(a)).pin_memory()
    return t.numpy(), t


def pinned_empty(n, dtype):
    import torch
    t = torch.empty(n, dtype=dtype).pin_memory()
    return t.numpy(), t


def run_cuda(args):
    import torch
    import torch.distributed as dist
    from binless_b200 import cuda_lib
    from oracle.loader import oracle_lib

    rank = int(os.environ.get("RANK", 0)); world = int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (the product has no CPU fallback)")
    torch.cuda.set_device(local)
    numa = None
    if world > 1:
        from binless_b200.dist import bind_to_gpu_numa_node
        numa = bind_to_gpu_numa_node(local)          # before any pinned allocation
        os.environ.setdefault("FB_HOST_THREADS", str(max(1, (os.cpu_count() or 8) // world)))
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    lib = cuda_lib()
    cfg = CONFIGS["C2"]
    n, D = cfg["nbins"], cfg["ndatasets"]
    mat = synth(n, D, cfg["res"], cfg["seed"] + rank)           # each rank: its own locus of the same shape
    keep = []
    pin = {}
    for k in ("pos1", "pos2", "observed", "nobs", "distance"):   # host buffers of the call, pinned (contract)
        pin[k], t = pinned_like(mat[k]); keep.append(t)
    pin.update(name=None, bin1=None, bin2=None)                  # canonical (name, bin1, bin2) order is implied
    N = pin["distance"].size
    ncells = n * (n + 1) // 2
    outbuf = {}
    for k in ("background", "biasmat", "decaymat", "signal", "binless", "phihat", "weight", "log_signal"):
        outbuf[k], t = pinned_empty(N, torch.float64); keep.append(t)
    outbuf["patchno"], t = pinned_empty(N, torch.int32); keep.append(t)
    h2d_bytes = int(pin["observed"].nbytes + pin["nobs"].nbytes)
    d2h_bytes = int(sum(v.nbytes for v in outbuf.values()))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def sum_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())

    # ---- value: inputs resident in HBM ------------------------------------------------------------------
    h = lib.create(pin, n)
    for _ in range(args.warmup):
        h.run_signal(ALPHA, LAM2, LAM1, columns="none")
    clocks = ClockSampler(local) if rank == 0 else None
    cell_iters = 0; iters = 0; launches = 0; ms_solver = 0.0; ms_dev = 0.0
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        o = h.run_signal(ALPHA, LAM2, LAM1, columns="none")
        d = o["diagnostics"]
        cell_iters += d["admm_cell_iters"]; iters += d["admm_iters"]; ms_solver += d["ms_solver"]; ms_dev += d["ms_total"]
        launches += h.kernel_launches()
    barrier()
    t_value = max_over_ranks(time.perf_counter() - t0)
    clk = clocks.stop() if clocks else None
    nsteps_outer = d["nsteps"]; solve_iters = [int(x) for x in d["solve_iters"]]
    h.close()
    total_cell_iters = sum_over_ranks(float(cell_iters))
    value = total_cell_iters / t_value / 1e6

    # ---- e2e: public API, host buffers, copies inside the timed region -----------------------------------
    def e2e_step():
        with lib.create(pin, n) as hh:
            oo = hh.run_signal(ALPHA, LAM2, LAM1, columns=True, buffers=outbuf)
            return oo["diagnostics"]["admm_cell_iters"], float(oo["log_signal"][ncells - 1])
    for _ in range(max(1, min(args.warmup, 2))):
        e2e_step()
    barrier()
    t0 = time.perf_counter()
    e2e_ci = 0
    for _ in range(args.steps):
        ci, _probe = e2e_step()
        e2e_ci += ci
    barrier()
    t_e2e = max_over_ranks(time.perf_counter() - t0)
    e2e_value = sum_over_ranks(float(e2e_ci)) / t_e2e / 1e6

    peak, peak_src = measured_peak()
    # ---- e2e again with what an R caller has: pageable buffers, every input column passed (rank-local, 2 steps) ----
    pageable = {k: np.array(v, copy=True) for k, v in mat.items()}
    outp = {k: np.empty_like(v) for k, v in outbuf.items()}
    def e2e_pageable_step():
        with lib.create(pageable, n) as hh:
            return hh.run_signal(ALPHA, LAM2, LAM1, columns=True, buffers=outp)["diagnostics"]["admm_cell_iters"]
    e2e_pageable_step()
    barrier()
    t0 = time.perf_counter()
    pg_ci = sum(e2e_pageable_step() for _ in range(2))
    barrier()
    t_pg = max_over_ranks(time.perf_counter() - t0)
    pg_value = sum_over_ranks(float(pg_ci)) / t_pg / 1e6
    del pageable, outp
    extra = {} if args.no_extra else extra_legs(lib, rank, world, local, peak)
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    sec_per_iter = (ms_solver / 1e3) / max(1, iters)
    achieved = ADMM_BYTES_PER_CELL_ITER * ncells / sec_per_iter / 1e9
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": t_value / args.steps * 1e3, "device_ms_per_step": ms_dev / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic",
        "config": {"workload": WORKLOAD, "nbins": n, "ndatasets": D, "cells_per_dataset": ncells,
                   "outer_steps": int(nsteps_outer), "admm_iters_per_solve": solve_iters,
                   "l2": "inputs larger than L2 (working set > 1 GB per call)", "units_per_rank": 1,
                   "host": {"cpus": os.cpu_count(), "numa_node_of_rank0_gpu": numa, "host_threads_per_rank": os.environ.get("FB_HOST_THREADS")}},
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d_bytes, "d2h_bytes_per_step": d2h_bytes,
                "ms_per_step": t_e2e / args.steps * 1e3,
                "what": "fb_create (validation + H2D) + fb_run_signal + fb_destroy per step, pinned host buffers; every output column "
                        "is copied to the host inside the call: phihat / weight / biasmat / decaymat on a second stream while the "
                        "last solves run, the rest after them (FB_NO_EARLY_DOWNLOAD=1: all after)"},
        "e2e_pageable": {"value": pg_value, "unit": UNIT, "ms_per_step": t_pg / 2 * 1e3, "steps": 2,
                         "what": "numpy (pageable) host buffers, name / bin1 / bin2 passed and validated: what the Rcpp shim supplies"},
        "gpu_launches": int(launches),
        "clocks": clk,
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                     "traffic": ncu_traffic(), "peak_source": peak_src,
                     "traffic_source": "ncu --set full, dram__bytes_read.sum + dram__bytes_write.sum of k_beta + chain prox + k_ctl, mean of "
                                       "iterations 2 / 9 / 17 of the first solve (profiles/r2_admm_it*_ncu_full_summary.json)",
                     "kernel": "ADMM iteration of the fused-lasso solve (GflSolver::iteration = k_beta + k_prox | k_prox_aset + k_ctl)",
                     "bytes_per_launch": ADMM_BYTES_PER_CELL_ITER * ncells, "us_per_launch": sec_per_iter * 1e6,
                     "solver_share_of_device_time": ms_solver / max(ms_dev, 1e-9)},
    }
    # ---- CPU baseline: the oracle on the same workload, one full call, rank 0 only, N = 1 only ------------
    line.update(extra)
    if world == 1 and not args.no_cpu:
        line["cpu_baseline"] = cpu_baseline(mat, n, full=True)
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def extra_legs(lib, rank, world, local, peak):
    """The other configurations BASELINE.json names, in the same run (rank 0 reports):
       c3_sweep      : 64-point lambda2 / BIC sweep of the fitted C2 model, cold-start solves dealt round-robin to the
                       ranks (strong scaling, no data-path collective; binless_b200/sweep.py)
       c4_trailblock : ONE fused-lasso solve at 25 000 bins (3.1e8 cells) shared by the ranks through the library's
                       NCCL communicator (trail blocks, fb_comm_init): ms per ADMM iteration; and whether a sharded
                       solve equals the single-GPU one bit for bit (3000 bins, 15 steps, both chain-prox kernels)."""
    import ctypes as C
    import torch
    import torch.distributed as dist
    from binless_b200._abi import _dp, fb_gfl_state, gfl_solve
    from binless_b200.dist import init_comm, shutdown_comm
    from binless_b200.shard import partition_units
    from binless_b200.sweep import lambda_grid
    out = {}

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def reduce(x, op):
        if world == 1:
            return float(x)
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=op)
        return float(t.item())

    # ---- C3 ------------------------------------------------------------------------------------------------
    cfg = CONFIGS["C2"]; n, D = cfg["nbins"], cfg["ndatasets"]
    mat = synth(n, D, cfg["res"], cfg["seed"])                      # the same locus on every rank
    slim = {k: (None if k in ("name", "bin1", "bin2") else v) for k, v in mat.items()}
    grid = lambda_grid(64)
    mine = partition_units(grid.size, world, rank)
    with lib.create(slim, n) as h:
        h.run_signal(ALPHA, LAM2, LAM1, columns="none")             # every rank fits the same model (deterministic)
        h.sweep_lambda(ALPHA, grid[mine][:1], 0.0)                  # warm-up
        barrier()
        t0 = time.perf_counter()
        part = h.sweep_lambda(ALPHA, grid[mine], 0.0)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
    rows = [(int(u), part["bic"][i].tolist()) for i, u in enumerate(mine)]
    if world > 1:
        bucket = [None] * world
        dist.all_gather_object(bucket, rows)
        rows = sorted(r for b in bucket for r in b)
    t_c3 = reduce(dt, dist.ReduceOp.MAX) if world > 1 else dt
    ci_c3 = reduce(float(part["admm_cell_iters"]), dist.ReduceOp.SUM) if world > 1 else float(part["admm_cell_iters"])
    bic = np.array([r[1] for r in rows])
    out["c3_sweep"] = {"workload": f"C3: 64-point lambda2/BIC sweep, {n} bins x {D} datasets, cold-start solves, lambda grid dealt round-robin",
                       "n_gpus": world, "scaling": "strong", "seconds": t_c3, "Mcell_iter_per_s": ci_c3 / t_c3 / 1e6,
                       "solves": int(grid.size * D), "best_lam2": grid[np.argmin(bic, axis=0)].tolist()}
    del mat, slim

    # ---- trail blocks -----------------------------------------------------------------------------------------
    def fixed_steps(nb, y, w, steps):
        beta = np.zeros(nb * (nb + 1) // 2)
        st = fb_gfl_state(beta=_dp(beta), z=None, u=None, alpha=5.0, steps=0)
        barrier()
        t = time.perf_counter()
        lib._check(lib._gfl_solve(nb, 0, _dp(y), _dp(w), 2.5, 1e-12, steps, 0, C.byref(st), None), None)
        barrier()
        # device time of the ADMM loop alone (events on the solver's stream), slowest rank: the call also uploads 5 GB and
        # sets the exchange up, whose run-to-run variation is larger than ten iterations on 8 GPUs
        return reduce(lib._gfl_last_solve_ms() / 1e3, dist.ReduceOp.MAX), st.steps, beta

    nchk = 3000; ncc = nchk * (nchk + 1) // 2
    rng = np.random.default_rng(44)
    yc = np.repeat(rng.normal(0, 1, ncc // 30 + 1), 30)[:ncc] + rng.normal(0, 0.5, ncc); wc = rng.uniform(0, 3, ncc)
    single = gfl_solve(lib, nchk, yc, wc, 2.5, 1e-12, 16, 0) if world > 1 else None     # communicator off: this GPU alone
    nb = CONFIGS["C4"]["nbins"]; nc4 = nb * (nb + 1) // 2
    rng = np.random.default_rng(4)
    tile = rng.normal(0, 0.5, 1 << 20)
    y4 = np.repeat(rng.normal(0, 1, nc4 // 40 + 1), 40)[:nc4]
    y4 += np.resize(tile, nc4)
    w4 = np.resize(rng.uniform(0, 3, 1 << 20), nc4)
    identical = None
    if world > 1:
        init_comm(lib)
        shared = gfl_solve(lib, nchk, yc, wc, 2.5, 1e-12, 16, 0)
        same = all(np.array_equal(single[k], shared[k]) for k in ("beta", "z", "u", "beta_full")) and \
            single["steps"] == shared["steps"] and single["alpha"] == shared["alpha"]
        identical = reduce(0.0 if same else 1.0, dist.ReduceOp.SUM) == 0.0
    fixed_steps(nb, y4, w4, 2)
    ta, sa, _ = fixed_steps(nb, y4, w4, 2)
    tb, sb, _ = fixed_steps(nb, y4, w4, 22)
    # ---- C5: chromosome scale end to end (human chr1 at 5 kb: 50 000 bins, 1.25e9 cells), the matrix built in HBM from
    # cut sites and read pairs (fb_create_binned: bin_data on the device), fast_binless with the lasso on the 2000-wide
    # band (R/chromosome.R:74,244-246: maxdiag = far.cutoff / base.res, compute_patchnos = F); the ranks share every solve
    from binless_b200.synth import synth_reads
    c5 = CONFIGS["C5"]
    reads = synth_reads(c5["nbins"], 1, c5["res"], c5["seed"], reads_per_dataset=30_000_000)
    barrier()
    t0 = time.perf_counter()
    h5 = lib.create_binned(reads, c5["res"], 1, fetch_mat=False)
    torch.cuda.synchronize()
    t_create5 = time.perf_counter() - t0
    barrier()
    t0 = time.perf_counter()
    o5 = h5.run_signal(ALPHA, LAM2, LAM1, columns="none", maxdiag=2000, compute_patchnos=False)
    barrier()
    t_call5 = time.perf_counter() - t0
    t0 = time.perf_counter()      # the same call again on the handle: without the one-off set-up (50 GB of first allocations,
    o5 = h5.run_signal(ALPHA, LAM2, LAM1, columns="none", maxdiag=2000, compute_patchnos=False)   # chain plans, peer mappings)
    barrier()
    t_rep5 = time.perf_counter() - t0
    d5 = o5["diagnostics"]
    out["c5_chromosome"] = {"workload": f"C5: fast_binless at {h5.nbins} bins ({h5.ncells} cells, {reads['rd_pos1'].size} read pairs binned on the "
                                        "device), maxdiag = 2000, compute_patchnos = F, trail blocks over the ranks",
                            "n_gpus": world, "scaling": "strong", "seconds_create_binned": t_create5, "seconds_fast_binless": t_call5,
                            "seconds_fast_binless_repeat": t_rep5,
                            "outer_steps": int(d5["nsteps"]), "admm_iters_per_solve": [int(x) for x in d5["solve_iters"]],
                            "solver_ms": float(d5["ms_solver"]), "Mcell_iter_per_s": d5["admm_cell_iters"] / max(t_rep5, 1e-9) / 1e6}
    h5.close(); del reads
    lib.lib.fb_release_cached_memory()
    if world > 1:
        shutdown_comm(lib)
    # ---- one PROCESS driving several GPUs (fb_gfl_solve_multi: host thread + communicator per device): rank 0, 2 devices
    if world > 1 and rank == 0:
        multi = gfl_solve(lib, nchk, yc, wc, 2.5, 1e-12, 16, 0, ngpus=2)
        out["single_process_2gpu_bit_identical"] = bool(all(np.array_equal(single[k], multi[k]) for k in ("beta", "z", "u", "beta_full"))
                                                         and single["steps"] == multi["steps"] and single["alpha"] == multi["alpha"])
    barrier()
    per = (tb - ta) / max(1, sb - sa)
    out["c4_trailblock"] = {"workload": f"C4: one fused-lasso solve, {nb} bins ({nc4} cells), trail blocks over the ranks (fb_comm_init)",
                            "n_gpus": world, "scaling": "strong", "ms_per_admm_iteration": per * 1e3,
                            "Mcell_iter_per_s": nc4 / per / 1e6,
                            "roofline_frac_all_gpus": ADMM_BYTES_PER_CELL_ITER * nc4 / per / 1e9 / (peak * world),
                            "bit_identical_to_single_gpu": identical,
                            "bit_identity_check": f"{nchk} bins, 15 ADMM iterations (team kernel, then active set), beta / z / u / steps / alpha"
                                                  if world > 1 else "n/a at 1 GPU"}
    return out


def cpu_baseline(mat, n, full=True, budget_s=None):
    """Times the reference's CPU path (oracle/_ref when built from /root/reference, else the port). 1 thread: the
    reference has no parallel region on this path."""
    from oracle.loader import oracle_lib
    ora = oracle_lib("auto")
    t0 = time.perf_counter()
    with ora.create(mat, n) as h:
        o = h.run_signal(ALPHA, LAM2, LAM1)
    dt = time.perf_counter() - t0
    ncells = n * (n + 1) // 2
    D = mat["distance"].size // ncells
    return {"value": o["diagnostics"]["admm_cell_iters"] / dt / 1e6, "unit": UNIT, "cores": 1, "kind": ora.kind,
            "sample": f"one full fast_binless call, {n} bins x {D} datasets ({dt:.1f} s, 1 thread of {os.cpu_count()})",
            "seconds": dt, "admm_cell_iters": int(o["diagnostics"]["admm_cell_iters"]),
            "solver_seconds": o["diagnostics"]["ms_solver"] / 1e3}


def run_reference(args):
    rank = int(os.environ.get("RANK", 0)); world = int(os.environ.get("WORLD_SIZE", 1))
    if rank != 0:
        return
    cfg = CONFIGS["C2"]
    nb, D = cfg["nbins"], cfg["ndatasets"]
    # every step is one full call on the full C2 workload (the same config as the CUDA arm: ~16 s per call on one
    # core; the reference has no parallel region on this path)
    mat = synth(nb, D, cfg["res"], cfg["seed"])
    for _ in range(args.warmup):
        cpu_baseline(mat, nb)
    t0 = time.perf_counter()
    ci = 0
    for _ in range(args.steps):
        r = cpu_baseline(mat, nb)
        ci += r["admm_cell_iters"]
    dt = time.perf_counter() - t0
    value = ci / dt / 1e6
    sample = f"the full workload every step: {nb} bins x {D} datasets, one complete fast_binless call, 1 thread of {os.cpu_count()}"
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": dt / max(1, args.steps) * 1e3, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": WORKLOAD, "nbins": nb, "ndatasets": D, "sample": sample},
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": 1, "kind": r["kind"], "sample": sample},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="cuda", choices=["cuda", "reference"])
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg (development)")
    ap.add_argument("--no-extra", action="store_true", help="skip the C3 / C4 legs (development)")
    args = ap.parse_args()
    if args.warmup < 3 and args.impl == "cuda":
        args.warmup = 3
    if args.impl == "reference":
        run_reference(args)
    else:
        run_cuda(args)


if __name__ == "__main__":
    main()
