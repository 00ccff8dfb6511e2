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
    """DRAM bytes (read + write) of one ADMM iteration from the committed `ncu --set full` capture (profiles/)."""
    path = os.path.join(ROOT, "profiles", "r1s_admm_iteration_ncu_full_summary.json")
    try:
        rows = json.load(open(path))
        return int(sum((r["dram_read_MB"] + r["dram_write_MB"]) * 1e6 for r in rows))
    except Exception:
        return None


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
    from binless_b200 import cuda_lib, oracle_lib

    rank = int(os.environ.get("RANK", 0)); world = int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (the product has no CPU fallback)")
    torch.cuda.set_device(local)
    if world > 1:
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

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    peak, peak_src = measured_peak()
    sec_per_iter = (ms_solver / 1e3) / max(1, iters)
    achieved = ADMM_BYTES_PER_CELL_ITER * ncells / sec_per_iter / 1e9
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": t_value / args.steps * 1e3, "device_ms_per_step": ms_dev / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic",
        "config": {"workload": WORKLOAD, "nbins": n, "ndatasets": D, "cells_per_dataset": ncells,
                   "outer_steps": int(nsteps_outer), "admm_iters_per_solve": solve_iters,
                   "l2": "inputs larger than L2 (working set > 1 GB per call)", "units_per_rank": 1},
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d_bytes, "d2h_bytes_per_step": d2h_bytes,
                "ms_per_step": t_e2e / args.steps * 1e3},
        "gpu_launches": int(launches),
        "clocks": clk,
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                     "traffic": ncu_traffic(), "peak_source": peak_src,
                     "traffic_source": "ncu --set full, dram__bytes_read.sum + dram__bytes_write.sum of k_beta + k_prox + k_ctl, "
                                       "one iteration (profiles/r1s_admm_iteration_ncu_full_summary.json)",
                     "kernel": "ADMM iteration of the fused-lasso solve (GflSolver::iteration = k_beta + k_prox + k_ctl)",
                     "bytes_per_launch": ADMM_BYTES_PER_CELL_ITER * ncells, "us_per_launch": sec_per_iter * 1e6,
                     "solver_share_of_device_time": ms_solver / max(ms_dev, 1e-9)},
    }
    # ---- CPU baseline: the oracle on the same workload, one full call, rank 0 only, N = 1 only ------------
    if world == 1 and not args.no_cpu:
        line["cpu_baseline"] = cpu_baseline(mat, n, full=True)
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def cpu_baseline(mat, n, full=True, budget_s=None):
    """Times the reference's CPU path (oracle/_ref when built from /root/reference, else the port). 1 thread: the
    reference has no parallel region on this path."""
    from binless_b200 import oracle_lib
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
    n = cfg["nbins"]
    # bounded sample per step so that (steps + warmup) finishes within a few minutes: the full call is ~16-30 s
    total = args.steps + args.warmup
    D = cfg["ndatasets"] if total <= 5 else (2 if total <= 10 else 1)
    nb = n if total <= 20 else 1200
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
    sample = f"{nb} bins x {D} datasets per step (of the 2400 x 4 workload), full fast_binless call, 1 thread"
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": dt / max(1, args.steps) * 1e3, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": WORKLOAD, "sample": sample},
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
    args = ap.parse_args()
    if args.warmup < 3 and args.impl == "cuda":
        args.warmup = 3
    if args.impl == "reference":
        run_reference(args)
    else:
        run_cuda(args)


if __name__ == "__main__":
    main()
