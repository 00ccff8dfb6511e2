"""CPU: the per-word / per-point passes of k_prox_aset -- the very source the kernel compiles
(binless_b200/csrc/fb_prox_aset_passes.cuh), built for the host with tests/aset_host/shim.h and run under
AddressSanitizer -- return the minimiser of the reference's tf_dp_weight (gfl_tf.c:184-321, compiled from
/root/reference into oracle/_ref) for batches of several chains, any knot hypothesis, and with the sequential fallback
forced (maxpass 0 and 2). An index leaving the shared-memory arrays aborts the run."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HOST = os.path.join(ROOT, "tests", "aset_host")

DRIVER = r'''
import sys, ctypes as C, numpy as np
sys.path.insert(0, %(root)r); sys.path.insert(0, %(root)r + "/tools")
from certificate_prox_prototype import tf_dp_reference
ref = tf_dp_reference()
lib = C.CDLL(%(so)r)
dp = C.POINTER(C.c_double); ip = C.POINTER(C.c_int)
lib.aset_passes_host.argtypes = [C.c_int, ip, dp, dp, C.c_double, dp, C.c_int, C.c_int]
lib.aset_passes_host.restype = C.c_int
rng = np.random.default_rng(1)
worst = 0.0; fell = 0
for trial in range(360):
    nch = int(rng.integers(1, 6))
    Ls = [int(rng.integers(1, 400)) for _ in range(nch)]
    if trial %% 40 == 0: Ls[0] = int(rng.integers(1500, 2500))
    off = np.concatenate([[0], np.cumsum(Ls)]).astype(np.int32)
    lam = float(rng.choice([0.0, 0.2, 1.0, 4.0]))
    v = rng.normal(size=off[-1]) + np.repeat(rng.normal(0, 1.5, off[-1] // 6 + 1), 6)[:off[-1]]
    if trial %% 7 == 3: v = np.round(v)
    sol = lambda x, l: ref(x, 1.0, l) if (len(x) > 1 and l > 0) else x.copy()
    want = np.concatenate([sol(v[off[c]:off[c+1]], lam) for c in range(nch)])
    kind = trial %% 4
    hyp = [np.zeros_like(v), want.copy(), np.concatenate([sol(v[off[c]:off[c+1]] + rng.normal(0, .05, Ls[c]), max(lam, 1e-3)) for c in range(nch)]),
           rng.normal(size=v.size).round()][kind]
    maxpass = [64, 64, 64, 0, 2][trial %% 5]
    out = np.empty_like(v)
    r = lib.aset_passes_host(nch, off.ctypes.data_as(ip), v.ctypes.data_as(dp), np.ascontiguousarray(hyp).ctypes.data_as(dp), lam,
                             out.ctypes.data_as(dp), maxpass, 0)
    fell += r < 0
    e = float(np.abs(out - want).max())
    worst = max(worst, e)
    assert e < 1e-9, (trial, kind, Ls, lam, r, e)
    if maxpass == 64: assert r >= 0, (trial, "did not settle")
    if kind == 1 and maxpass == 64 and lam > 0 and trial %% 7 != 3: assert r <= 1, (trial, r)
print("OK worst", worst, "fallbacks", fell)
'''


def test_device_passes_on_host_under_asan(oracle_ref, tmp_path):
    so = str(tmp_path / "libaset_host.so")
    src = os.path.join(HOST, "aset_passes_host.cpp")
    env = dict(os.environ)
    built = False
    for cxx in ("/usr/bin/g++", os.environ.get("CXX", "g++"), "g++"):      # first a compiler that ships libasan
        r = subprocess.run([cxx, "-O1", "-g", "-fsanitize=address", "-fno-omit-frame-pointer", "-std=c++17", "-shared", "-fPIC",
                            src, "-o", so], capture_output=True, text=True)
        asan = subprocess.run([cxx, "-print-file-name=libasan.so"], capture_output=True, text=True).stdout.strip()
        if r.returncode == 0 and os.path.isabs(asan):
            env.update(LD_PRELOAD=asan, ASAN_OPTIONS="detect_leaks=0")
            built = True
            break
    if not built:       # no sanitizer runtime on this box: the same run without the guard
        r = subprocess.run([os.environ.get("CXX", "g++"), "-O1", "-g", "-std=c++17", "-shared", "-fPIC", src, "-o", so],
                           capture_output=True, text=True)
        assert r.returncode == 0, r.stderr
    r = subprocess.run([sys.executable, "-c", DRIVER % {"root": ROOT, "so": so}], capture_output=True, text=True, env=env, timeout=600)
    assert r.returncode == 0 and "OK worst" in r.stdout, r.stdout[-2000:] + r.stderr[-4000:]
