"""Input-side timing: fb_create_binned (reads -> dense matrix on the device) against the host
route (numpy restatement of bin_data + fb_create with its 16 B/cell validation and 8 B/cell upload) at C2's shape."""
import json, sys, time
import numpy as np
sys.path.insert(0, '.')
from binless_b200 import cuda_lib
sys.path.insert(0, 'oracle')
from bin_data_numpy import bin_reads_numpy
from binless_b200.synth import synth_reads
n, D, res = 2400, 4, 5000
nreads = int(sys.argv[1]) if len(sys.argv) > 1 else 5_000_000
r = synth_reads(n, D, res, 2, reads_per_dataset=nreads)
lib = cuda_lib()
lib.create_binned(r, res, D).close()                       # warm-up (allocations)
t0 = time.perf_counter(); h = lib.create_binned(r, res, D); t_dev = time.perf_counter() - t0
t0 = time.perf_counter(); mat = bin_reads_numpy(r, res, D); t_np = time.perf_counter() - t0
t0 = time.perf_counter(); h2 = lib.create(mat, n); t_create = time.perf_counter() - t0
ok = all(np.array_equal(h.cols[k], mat[k]) for k in ("observed", "nobs", "pos1", "pos2", "distance"))
in_bytes = sum(v.nbytes for v in r.values())
print(json.dumps({"workload": f"bin_data: {D} x {nreads} read rows -> {n} bins x {D} datasets", "bit_exact": bool(ok),
                  "seconds_fb_create_binned_incl_get_mat": t_dev, "seconds_numpy_bin_data": t_np, "seconds_fb_create_dense": t_create,
                  "h2d_bytes_reads": in_bytes, "h2d_bytes_dense_route": int(mat["observed"].nbytes + mat["nobs"].nbytes)}))
h.close(); h2.close()
