"""Development aid: wall-clock split of one end-to-end fast_binless call (create / run / outputs / destroy)."""
import sys, time, numpy as np, torch
sys.path.insert(0, '.')
from binless_b200 import cuda_lib
from binless_b200.synth import synth
n = int(sys.argv[1]) if len(sys.argv) > 1 else 2400
D = int(sys.argv[2]) if len(sys.argv) > 2 else 4
mat = synth(n, D, 5000, 2)
pin = {}
keep = []
for k in ("pos1", "pos2", "observed", "nobs", "distance"):
    t = torch.from_numpy(np.ascontiguousarray(mat[k])).pin_memory(); keep.append(t); pin[k] = t.numpy()
pin.update(name=None, bin1=None, bin2=None)
N = pin["distance"].size
out = {}
for k in ("background", "biasmat", "decaymat", "signal", "binless", "phihat", "weight", "log_signal"):
    t = torch.empty(N, dtype=torch.float64).pin_memory(); keep.append(t); out[k] = t.numpy()
t = torch.empty(N, dtype=torch.int32).pin_memory(); keep.append(t); out["patchno"] = t.numpy()
lib = cuda_lib()
for rep in range(4):
    t0 = time.perf_counter(); h = lib.create(pin, n)
    t1 = time.perf_counter(); o = h.run_signal(0.5, 2.5, 0.0, columns="none")
    t2 = time.perf_counter(); o = h.run_signal(0.5, 2.5, 0.0, columns=True, buffers=out)
    t3 = time.perf_counter(); h.close()
    t4 = time.perf_counter()
    print(f"rep {rep}: create {1e3*(t1-t0):.2f} ms, run(no columns) {1e3*(t2-t1):.2f} ms, run(all columns) {1e3*(t3-t2):.2f} ms, destroy {1e3*(t4-t3):.2f} ms")
