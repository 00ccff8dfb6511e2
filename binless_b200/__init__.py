"""binless_b200 -- B200-native fast-binless hot path.

The product is ``libfastbinless_cuda.so`` (C ABI in ``include/fastbinless.h``); this package is the Python
stand-in for the reference's R entry points (R is not available in this image):

    fast_binless(mat, nbins, alpha, lam2, lam1=0, nouter=25, tol_val=0.2, bg_steps=5, free_decay=10000,
                 compute_patchnos=True, csv_out="", maxdiag=0)          -> dict  (binless:::fast_binless)
    fast_binless_difference(out, mat, ref, alpha, lam2, lam1=0, tol_val=0.2, compute_patchnos=True,
                            maxdiag=0)                                   -> dict  (binless:::fast_binless_difference)

    wgfl_signal_BIC(cts, dispersion, nouter, nbins, GFLState, lam2, tol_val, metadata, beta, constrained, fixed, positive,
                    lambda1_fixed, lambda1_fix_value)                    -> dict  (binless:::wgfl_signal_BIC)
    wgfl_diff_BIC(cts, ref, dispersion, nouter, nbins, GFLState, lam2, tol_val, metadata, phi_ref, beta, constrained,
                  lambda1_fixed, lambda1_fix_value)                      -> dict  (binless:::wgfl_diff_BIC)

Same names, argument meaning, defaults and error behaviour as /root/reference/binless/src/csnorm_cpp.cpp:18-26 and
perf_iteration_signal.hpp / perf_iteration_diff.hpp.
There is no CPU fallback: the CUDA library must load and a GPU must be present.
"""
from __future__ import annotations

import os

from ._abi import FastBinlessError, FastBinlessLib

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("FB_LIB") or os.path.join(_HERE, "libfastbinless_cuda.so")   # FB_LIB: development builds
_lib = None


def cuda_lib() -> FastBinlessLib:
    """The product library. Fails loudly when it has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise FastBinlessError(-1, f"{LIB_PATH} is missing: run `python -m binless_b200.build` "
                                       "(there is no CPU fallback)")
        _lib = FastBinlessLib(LIB_PATH, "fb_")
    return _lib


def fast_binless(mat, nbins, alpha, lam2, lam1=0.0, nouter=25, tol_val=2e-1, bg_steps=5, free_decay=10000,
                 compute_patchnos=True, csv_out="", maxdiag=0, verbose=False, columns=True, ngpus=1):
    """binless:::fast_binless (csnorm_cpp.cpp:18-22 -> fast_binless.cpp:85-184) on the GPU; ngpus > 1 shares every
    fused-lasso solve between that many devices of this process (fb_create_multi)."""
    with cuda_lib().create(mat, nbins, ngpus) as h:
        out = h.run_signal(alpha, lam2, lam1, columns=columns, nouter=nouter, tol_val=tol_val, bg_steps=bg_steps,
                           free_decay=free_decay, compute_patchnos=compute_patchnos, csv_out=csv_out,
                           maxdiag=maxdiag, verbose=verbose)
        out["diagnostics"]["kernel_launches"] = h.kernel_launches()
    return out


def fast_binless_difference(out, mat, ref, alpha, lam2, lam1=0.0, tol_val=2e-1, compute_patchnos=True, maxdiag=0,
                            verbose=False):
    """binless:::fast_binless_difference (csnorm_cpp.cpp:23-26 -> fast_binless.cpp:235-279) on the GPU.

    ``out`` is the value returned by fast_binless; ``mat`` the input columns (``out$mat`` in R carries them)."""
    with cuda_lib().create(mat, out["nbins"]) as h:
        h.restore(out["log_signal"], out["biases"], out["decay"], out["exposures"])
        res = h.run_difference(ref, alpha, lam2, lam1, tol_val=tol_val, compute_patchnos=compute_patchnos,
                               maxdiag=maxdiag, verbose=verbose)
        res["diagnostics"]["kernel_launches"] = h.kernel_launches()
    return res


def wgfl_signal_BIC(cts, dispersion, nouter, nbins, GFLState, lam2, tol_val, metadata, beta, constrained, fixed,
                    positive=True, lambda1_fixed=False, lambda1_fix_value=0.0):
    """binless:::wgfl_signal_BIC (perf_iteration_signal.cpp:22-157) on the GPU; `cts` / `metadata` as dicts of columns."""
    from ._abi import wgfl_signal_bic
    return wgfl_signal_bic(cuda_lib(), cts, dispersion, nouter, nbins, GFLState, lam2, tol_val, metadata, beta, constrained,
                           fixed, positive, lambda1_fixed, lambda1_fix_value)


def wgfl_diff_BIC(cts, ref, dispersion, nouter, nbins, GFLState, lam2, tol_val, metadata, phi_ref, beta, constrained,
                  lambda1_fixed=False, lambda1_fix_value=0.0):
    """binless:::wgfl_diff_BIC (perf_iteration_diff.cpp:20-117) on the GPU."""
    from ._abi import wgfl_diff_bic
    return wgfl_diff_bic(cuda_lib(), cts, ref, dispersion, nouter, nbins, GFLState, lam2, tol_val, metadata, phi_ref, beta,
                         constrained, lambda1_fixed, lambda1_fix_value)
