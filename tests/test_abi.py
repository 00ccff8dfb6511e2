"""CPU: the product library loads, exports every symbol include/fastbinless.h declares, the ctypes mirror matches
the header's struct layout, and the compute entry points fail loudly without a GPU (no CPU fallback)."""
import ctypes as C
import os
import re
import subprocess
import sys

import numpy as np
import pytest

from binless_b200 import _abi
from binless_b200.synth import synth

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "fastbinless.h")


@pytest.fixture(scope="module")
def product_lib():
    from binless_b200 import build as b
    path = b.build()
    return _abi.FastBinlessLib(path, "fb_")


def declared_symbols():
    src = open(HEADER).read()
    return sorted(set(re.findall(r"FB_NAME\((\w+)\)\s*\(", src)))


def test_header_symbols_are_exported(product_lib):
    names = declared_symbols()
    assert set(names) == set(_abi.ENTRY_POINTS)
    for n in names:
        assert hasattr(product_lib.lib, "fb_" + n), n


def test_oracle_exports_same_abi(oracle_port):
    for n in declared_symbols():
        assert hasattr(oracle_port.lib, "fbo_" + n), n


def test_ctypes_mirror_matches_header_layout(tmp_path):
    """sizeof/offsetof of every struct, computed by the C compiler from the header, equals the ctypes mirror."""
    structs = {"fb_mat": _abi.fb_mat, "fb_params": _abi.fb_params, "fb_mean_state": _abi.fb_mean_state,
               "fb_gam_state": _abi.fb_gam_state, "fb_signal_out": _abi.fb_signal_out, "fb_diff_out": _abi.fb_diff_out,
               "fb_cv_out": _abi.fb_cv_out, "fb_gfl_state": _abi.fb_gfl_state, "fb_cv_score_out": _abi.fb_cv_score_out,
               "fb_cts": _abi.fb_cts, "fb_perf_meta": _abi.fb_perf_meta, "fb_perf_params": _abi.fb_perf_params,
               "fb_perf_signal_out": _abi.fb_perf_signal_out, "fb_perf_diff_out": _abi.fb_perf_diff_out,
               "fb_band_params": _abi.fb_band_params, "fb_band_out": _abi.fb_band_out, "fb_sweep_out": _abi.fb_sweep_out}
    lines = ['#include <stdio.h>', '#include <stddef.h>', f'#include "{HEADER}"', 'int main(void){']
    for name, st in structs.items():
        lines.append(f'printf("{name} %zu\\n", sizeof({name}));')
        for fname, _ in st._fields_:
            cname = "lambda" if fname == "lambda_" else fname
            lines.append(f'printf("{name}.{fname} %zu\\n", offsetof({name}, {cname}));')
    lines.append('return 0;}')
    src = tmp_path / "layout.c"; src.write_text("\n".join(lines))
    exe = tmp_path / "layout"
    subprocess.run(["gcc", str(src), "-o", str(exe)], check=True)
    got = dict(l.split() for l in subprocess.run([str(exe)], capture_output=True, text=True, check=True).stdout.splitlines())
    for name, st in structs.items():
        assert int(got[name]) == C.sizeof(st), name
        for fname, _ in st._fields_:
            assert int(got[f"{name}.{fname}"]) == getattr(st, fname).offset, (name, fname)


def _has_gpu():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


@pytest.mark.skipif(_has_gpu(), reason="checks the behaviour on a box without a CUDA device")
def test_no_cpu_fallback_without_gpu(product_lib):
    mat = synth(8, 1, 5000, 1)
    with pytest.raises(_abi.FastBinlessError, match="no CUDA device") as ei:
        product_lib.create(mat, 8)
    assert ei.value.code == 2  # FB_ERR_CUDA
    y = np.zeros(36); w = np.ones(36)
    with pytest.raises(_abi.FastBinlessError, match="no CUDA device"):
        _abi.gfl_solve(product_lib, 8, y, w, 1.0, 1e-3)


def test_product_never_links_the_oracle(product_lib):
    """The shipped library has no dependency on, and no symbol from, oracle/."""
    out = subprocess.run(["ldd", product_lib.path], capture_output=True, text=True).stdout
    assert "fbo" not in out and "oracle" not in out
    syms = subprocess.run(["nm", "-D", product_lib.path], capture_output=True, text=True).stdout
    assert "fbo_" not in syms and "gflp_" not in syms
    for f in os.listdir(os.path.join(ROOT, "binless_b200", "csrc")):
        assert "oracle" not in open(os.path.join(ROOT, "binless_b200", "csrc", f)).read().replace("the CPU oracle", ""), f
    # nor does the Python host side: the loader of the oracle lives in oracle/loader.py, outside the package
    for f in os.listdir(os.path.join(ROOT, "binless_b200")):
        if f.endswith(".py"):
            src = open(os.path.join(ROOT, "binless_b200", f)).read()
            assert "oracle.loader" not in src and "import oracle" not in src and "libfbo" not in src, f
