"""Summaries of ncu captures for profiles/:
   python tools/ncu_summary.py launches gpurun_out/r1q_launches.csv            -> per-kernel totals / shares (text)
   python tools/ncu_summary.py full gpurun_out/r1q_admm_iter.ncu-rep           -> per-launch JSON rows (needs ncu here)"""
import collections, csv, io, json, re, subprocess, sys

def launches(path):
    rows = [r for r in csv.reader(l for l in open(path) if not l.startswith("==")) if r]
    hdr = rows[0]; ix = {h: i for i, h in enumerate(hdr)}
    per = collections.OrderedDict()
    for r in rows[1:]:
        if len(r) != len(hdr) or r[ix["Metric Name"]] != "gpu__time_duration.sum":
            continue
        name = re.sub(r"\(.*", "", r[ix["Kernel Name"]]).replace("fb::<unnamed>::", "")
        v = float(r[ix["Metric Value"]].replace(",", "")); u = r[ix["Metric Unit"]]
        v = v / 1e3 if u in ("ns", "nsecond") else v * (1e3 if u in ("ms", "msecond") else 1)
        per.setdefault(name, []).append(v)
    total = sum(sum(v) for v in per.values())
    # launches that return at once (the chain-prox kernel a step does not use; launches queued past convergence in the
    # polling loop) would dilute the means: a launch under 10 % of the kernel's longest one counts as "skipped"
    print(f"{'kernel':34s} {'real':>5s} {'skipped':>7s} {'total us':>10s} {'avg real us':>12s} {'share':>6s}")
    for k, v in sorted(per.items(), key=lambda kv: -sum(kv[1])):
        cut = 0.1 * max(v)
        real = [x for x in v if x >= cut]
        print(f"{k:34s} {len(real):5d} {len(v) - len(real):7d} {sum(v):10.1f} {sum(real) / len(real):12.1f} {100 * sum(v) / total:5.1f}%")
    print(f"total kernel time {total/1e3:.2f} ms (cold-cache, serialised by ncu: shares, not absolutes)")

def full(path):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units = rows[0], rows[1]; ix = {h: i for i, h in enumerate(hdr)}
    def g(r, k):
        try: return float(r[ix[k]].replace(",", ""))
        except Exception: return None
    def scaled(r, k, want):
        v = g(r, k); u = units[ix[k]]
        if v is None: return None
        f = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1e-3, "us": 1, "ms": 1e3, "nsecond": 1e-3, "usecond": 1, "msecond": 1e3}.get(u, 1)
        return v * f / (1e6 if want == "MB" else 1)
    res = []
    for r in rows[2:]:
        name = re.sub(r"\(.*", "", r[ix["Kernel Name"]]).replace("fb::<unnamed>::", "")
        res.append({"kernel": name, "us": scaled(r, "gpu__time_duration.sum", "us"),
                    "dram_read_MB": scaled(r, "dram__bytes_read.sum", "MB"), "dram_write_MB": scaled(r, "dram__bytes_write.sum", "MB"),
                    "dram_throughput_pct": g(r, "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed"),
                    "issue_active_pct": g(r, "smsp__issue_active.avg.pct_of_peak_sustained_active"),
                    "pipe_fp64_pct": g(r, "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active"),
                    "pipe_alu_pct": g(r, "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active"),
                    "pipe_lsu_pct": g(r, "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active"),
                    "warps_active_per_smsp": g(r, "smsp__warps_active.avg.per_cycle_active"),
                    "regs": g(r, "launch__registers_per_thread"), "inst": g(r, "smsp__inst_executed.sum")})
    print(json.dumps(res, indent=1))

if __name__ == "__main__":
    {"launches": launches, "full": full}[sys.argv[1]](sys.argv[2])
