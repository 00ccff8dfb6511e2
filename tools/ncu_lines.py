"""Per-source-line instruction counts and stall samples of one kernel from an ncu report (needs -lineinfo + --import-source on):
   python tools/ncu_lines.py gpurun_out/x.ncu-rep [top]"""
import csv, io, subprocess, sys
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
rows = list(csv.reader(io.StringIO(out)))
hdr = None; res = []; cur_file = ""
for r in rows:
    if len(r) >= 2 and r[0] == "File Path": cur_file = r[1].split("/")[-1]; continue
    if len(r) > 8 and r[0] == "Line No": hdr = {h: i for i, h in enumerate(r)}; continue
    if hdr and len(r) > 8 and r[0].isdigit():
        num = lambda v: float(v) if v not in ("", "-") else 0.0
        res.append((cur_file, int(r[0]), r[1].strip()[:110], int(num(r[7])), int(num(r[6])), num(r[10])))
tot = sum(x[3] for x in res); tots = sum(x[4] for x in res)
print(f"total warp instructions {tot}, samples {tots}")
for f, ln, src, inst, smp, thr in sorted(res, key=lambda x: -x[3])[:top]:
    print(f"{f}:{ln:4d} inst {100*inst/tot:5.1f}% smp {100*smp/max(1,tots):5.1f}% thr {thr:4.0f} | {src}")

if len(sys.argv) > 3:   # ranges: "name:lo-hi,name:lo-hi" aggregated for the file given as argv[4] (default fb_prox_lane.cu)
    fname = sys.argv[4] if len(sys.argv) > 4 else "fb_prox_lane.cu"
    print("--- by range")
    for spec in sys.argv[3].split(","):
        name, r = spec.split(":"); lo, hi = map(int, r.split("-"))
        sel = [x for x in res if x[0] == fname and lo <= x[1] <= hi]
        print(f"{name:14s} inst {100*sum(x[3] for x in sel)/tot:5.1f}%  samples {100*sum(x[4] for x in sel)/max(1,tots):5.1f}%")
    other = {}
    for x in res:
        if x[0] != fname: other[x[0]] = other.get(x[0], 0) + x[3]
    for k, v in sorted(other.items(), key=lambda kv: -kv[1])[:8]:
        print(f"{k:30s} inst {100*v/tot:5.1f}%")
