import sys, numpy as np
sys.path.insert(0, '.')
from binless_b200 import cuda_lib
from oracle.loader import oracle_lib
from binless_b200._abi import gfl_solve
lib = cuda_lib()
n = int(sys.argv[1]) if len(sys.argv) > 1 else 100
rng = np.random.default_rng(n + 30000)
nc = n * (n + 1) // 2
y = rng.normal(size=nc); w = rng.uniform(0, 3, size=nc); w[rng.random(nc) < 0.2] = 0
b = gfl_solve(lib, n, y, w, 1.3, 1e-3, 30000, 0)
a = gfl_solve(oracle_lib("auto"), n, y, w, 1.3, 1e-3, 30000, 0)
print("steps", a["steps"], b["steps"], "max diff", np.abs(a["beta"] - b["beta"]).max())
