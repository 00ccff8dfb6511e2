"""Writes tests/golden/*.npz: outputs of the oracle built against the REFERENCE's own C solver (oracle/_ref,
compiled in place from /root/reference) on seeded synthetic inputs. The reference ships no golden vectors
(SURVEY.md F1/F2), so these pin the repo's own oracle + CUDA path against regressions; run in the container
that has /root/reference:   python tools/make_golden.py
"""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle.loader import oracle_lib
from binless_b200.synth import synth

CASES = {
    "n8_d1": dict(n=8, D=1, seed=11, alpha=0.49, lam2=[2.53], lam1=[0.13], kw={}),
    "n33_d2": dict(n=33, D=2, seed=3, alpha=0.49, lam2=[2.53, 4.51], lam1=[0.13, 0.10], kw={}),
    "n200_d2_c1": dict(n=200, D=2, seed=1, alpha=0.49, lam2=[2.53, 4.51], lam1=[0.13, 0.10], kw={}),  # config C1
    "n96_d3_band": dict(n=96, D=3, seed=5, alpha=0.6, lam2=[1.5], lam1=[0.0], kw=dict(maxdiag=30)),
    "n120_d1_tight": dict(n=120, D=1, seed=9, alpha=0.5, lam2=[2.5], lam1=[0.05], kw=dict(tol_val=0.01, nouter=6, bg_steps=2)),
}

# input side (bin_data): the numpy statement-by-statement restatement of R/preprocess.R:114-146 on seeded reads
BIN_CASES = {"bin_data_n33_d2": dict(n=33, D=2, res=2000, seed=2), "bin_data_n64_d2_res777": dict(n=64, D=2, res=777, seed=4)}


def make_bin_golden(outdir):
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    from bin_data_numpy import bin_reads_numpy
    from binless_b200.synth import synth_reads
    for name, c in BIN_CASES.items():
        r = synth_reads(c["n"], c["D"], c["res"], c["seed"])
        m = bin_reads_numpy(r, c["res"], c["D"])
        np.savez_compressed(os.path.join(outdir, name + ".npz"), observed=m["observed"], nobs=m["nobs"], pos1=m["pos1"],
                            pos2=m["pos2"], distance=m["distance"], nbins=np.int64(m["nbins"]),
                            reads_sum=np.int64(sum(int(r[k].astype(np.int64).sum()) for k in r)))
        print(name, "bins", m["nbins"], "reads", int(m["observed"].sum()))


def main():
    make_bin_golden(os.path.join(ROOT, "tests", "golden"))
    ora = oracle_lib("reference")
    outdir = os.path.join(ROOT, "tests", "golden")
    os.makedirs(outdir, exist_ok=True)
    for name, c in CASES.items():
        mat = synth(c["n"], c["D"], 5000, c["seed"])
        with ora.create(mat, c["n"]) as h:
            out = h.run_signal(c["alpha"], c["lam2"], c["lam1"], **c["kw"])
            d = out["diagnostics"]
            save = dict(log_signal=out["log_signal"], patchno=out["patchno"], phihat=out["phihat"], weight=out["weight"],
                        binless=out["binless"], background=out["background"],
                        bias=out["biases"]["estimate"], decay_beta=out["decay"]["beta"],
                        decay_lambda=out["decay"]["lambda_"], decay_mean=out["decay"]["mean"],
                        expo=out["exposures"]["estimate"], nsteps=d["nsteps"], solve_iters=d["solve_iters"],
                        prec_abs=d["prec_abs"], observed_sum=np.int64(mat["observed"].astype(np.int64).sum()))
            if c["D"] > 1:
                df = h.run_difference(1, c["alpha"], 3.38, 0.14, maxdiag=c["kw"].get("maxdiag", 0))
                save.update(diff_difference=df["difference"], diff_deltahat=df["deltahat"], diff_phi_ref=df["phi_ref"],
                            diff_patchno=df["patchno"], diff_iters=df["diagnostics"]["solve_iters"])
        if out["log_signal"].size > 20000:  # keep the fixtures small: drop the bulky secondary columns
            for k in ("phihat", "weight", "background", "diff_deltahat", "diff_phi_ref"):
                save.pop(k, None)
        np.savez_compressed(os.path.join(outdir, name + ".npz"), **save)
        print(name, "steps", d["nsteps"], "iters", list(d["solve_iters"]))

if __name__ == "__main__":
    main()
