"""Development aid: a variant build of the library with extra compiler defines, next to the product build:
    python tools/build_variant.py <name> -DAS_UN=4 ...   ->  binless_b200/_variants/lib_<name>.so
Time it with tools/variant_time.py (FB_LIB is not needed: the path is an argument)."""
import os, subprocess, sys
from concurrent.futures import ThreadPoolExecutor
sys.path.insert(0, '.')
from binless_b200 import build as b
name, extra = sys.argv[1], sys.argv[2:]
out = os.path.join(b.HERE, "_variants"); obj = os.path.join(out, "obj_" + name)
os.makedirs(obj, exist_ok=True)
def cc(src):
    o = os.path.join(obj, os.path.basename(src) + ".o")
    r = subprocess.run([b.NVCC, *b.NVCC_FLAGS, *extra, "-Xptxas=-v", "-x", "cu", "-c", src, "-o", o], capture_output=True, text=True)
    if r.returncode: raise SystemExit(r.stderr)
    return o, r.stderr
with ThreadPoolExecutor(8) as ex: res = list(ex.map(cc, b.sources()))
for o, log in res:
    if "prox_aset" in o or "fb_gfl" in o:
        lines = log.splitlines()
        for i, l in enumerate(lines):
            if "Compiling entry function" in l and ("k_prox" in l or "k_beta" in l):
                print(l.split("'")[1][:60], "|", lines[i + 2].strip()[:110] if i + 2 < len(lines) else "")
lib = os.path.join(out, f"lib_{name}.so")
r = subprocess.run([b.NVCC, "-shared", "-o", lib, *[o for o, _ in res], "-gencode", "arch=compute_100a,code=sm_100a",
                    "-Xcompiler", "-fPIC", "-lcudart_static", "-lpthread", "-ldl", "-lrt", "-lz"], capture_output=True, text=True)
if r.returncode: raise SystemExit(r.stderr)
print(lib)
