"""ctypes mirror of include/fastbinless.h (structs + a thin caller).

``FastBinlessLib(path, prefix)`` binds one shared library that exports the C ABI
under ``prefix`` (``fb_`` for the CUDA product library). It performs the same
marshalling the Rcpp shim does: numpy columns in, freshly
allocated numpy columns out. No computation happens in Python.
"""
from __future__ import annotations

import ctypes as C
import numpy as np

FB_DECAY_K = 50
FB_MAX_OUTER = 256

c_dp = C.POINTER(C.c_double)
c_ip = C.POINTER(C.c_int32)


class fb_mat(C.Structure):
    _fields_ = [("N", C.c_int64), ("name", c_ip), ("bin1", c_ip), ("pos1", c_ip), ("bin2", c_ip),
                ("pos2", c_ip), ("observed", c_ip), ("nobs", c_ip), ("distance", c_dp)]


class fb_params(C.Structure):
    _fields_ = [("alpha", C.c_double), ("lam2", c_dp), ("n_lam2", C.c_int), ("lam1", c_dp),
                ("n_lam1", C.c_int), ("nouter", C.c_uint), ("tol_val", C.c_double),
                ("bg_steps", C.c_uint), ("free_decay", C.c_uint), ("compute_patchnos", C.c_int),
                ("csv_out", C.c_char_p), ("maxdiag", C.c_uint), ("verbose", C.c_int)]


class fb_mean_state(C.Structure):
    _fields_ = [("estimate", c_dp), ("n", C.c_int), ("mean", C.c_double)]


class fb_gam_state(C.Structure):
    _fields_ = [("beta", C.c_double * FB_DECAY_K), ("lambda_", C.c_double), ("mean", C.c_double)]


class fb_signal_out(C.Structure):
    _fields_ = [("background", c_dp), ("biasmat", c_dp), ("decaymat", c_dp), ("signal", c_dp),
                ("binless", c_dp), ("phihat", c_dp), ("weight", c_dp), ("patchno", c_ip),
                ("log_signal", c_dp), ("biases", fb_mean_state), ("decay", fb_gam_state),
                ("exposures", fb_mean_state), ("nsteps", C.c_int), ("converged", C.c_int),
                ("prec_abs", C.c_double * FB_MAX_OUTER), ("prec_rel", C.c_double * FB_MAX_OUTER),
                ("admm_iters", C.c_int64), ("admm_cell_iters", C.c_int64), ("n_solves", C.c_int64),
                ("solve_iters", c_ip), ("solve_iters_cap", C.c_int64),
                ("ms_total", C.c_double), ("ms_solver", C.c_double)]


class fb_diff_out(C.Structure):
    _fields_ = [("difference", c_dp), ("deltahat", c_dp), ("weight", c_dp), ("phi_ref", c_dp),
                ("patchno", c_ip), ("admm_iters", C.c_int64), ("admm_cell_iters", C.c_int64),
                ("n_solves", C.c_int64), ("solve_iters", c_ip), ("solve_iters_cap", C.c_int64),
                ("ms_total", C.c_double), ("ms_solver", C.c_double)]


class fb_cv_out(C.Structure):
    _fields_ = [("signal", c_dp), ("binless", c_dp), ("phihat", c_dp), ("weight", c_dp),
                ("log_signal", c_dp), ("patchno", c_ip)]


class fb_sweep_out(C.Structure):
    _fields_ = [("bic", c_dp), ("chisq", c_dp), ("dof", c_ip), ("steps", c_ip), ("ms_solver", C.c_double),
                ("admm_cell_iters", C.c_int64)]


class fb_cv_score_out(C.Structure):
    _fields_ = [("cv", c_dp), ("cv_sd", c_dp), ("steps", c_ip), ("ms_solver", C.c_double), ("admm_cell_iters", C.c_int64)]


class fb_cts(C.Structure):
    _fields_ = [("N", C.c_int64), ("bin1", c_ip), ("bin2", c_ip), ("count", c_dp), ("lmu_nosig", c_dp), ("nobs", c_dp),
                ("log_decay", c_dp)]


class fb_perf_meta(C.Structure):
    _fields_ = [("n_bad_rows", C.c_int), ("bad_rows", c_ip), ("n_bad_diags", C.c_int), ("bad_diags", c_ip), ("diag_grp", c_ip)]


class fb_perf_params(C.Structure):
    _fields_ = [("dispersion", C.c_double), ("nouter", C.c_int), ("lam2", C.c_double), ("tol_val", C.c_double),
                ("constrained", C.c_int), ("fixed", C.c_int), ("positive", C.c_int), ("lambda1_fixed", C.c_int),
                ("lambda1_fix_value", C.c_double)]


class fb_perf_signal_out(C.Structure):
    _fields_ = [("phi", c_dp), ("beta", c_dp), ("phihat", c_dp), ("nobs", c_dp), ("weight", c_dp), ("patchno", c_ip),
                ("dof", C.c_int), ("converged", C.c_int), ("nouter_done", C.c_int), ("ncandidates", C.c_int),
                ("BIC", C.c_double), ("BIC_sd", C.c_double), ("eCprime", C.c_double), ("lambda1", C.c_double),
                ("admm_iters", C.c_int64)]


class fb_perf_diff_out(C.Structure):
    _fields_ = [("phi_ref", c_dp), ("delta", c_dp), ("beta", c_dp), ("phihat", c_dp), ("phihat_var", c_dp),
                ("phihat_ref", c_dp), ("phihat_var_ref", c_dp), ("nobs", c_dp), ("deltahat", c_dp), ("weight", c_dp),
                ("patchno", c_ip), ("dof", C.c_int), ("converged", C.c_int), ("nouter_done", C.c_int),
                ("ncandidates", C.c_int), ("BIC", C.c_double), ("BIC_sd", C.c_double), ("lambda1", C.c_double),
                ("admm_iters", C.c_int64)]


class fb_band_params(C.Structure):
    _fields_ = [("far_cutoff", C.c_double), ("maxdiag", C.c_uint), ("lam1_final", C.c_double), ("lam1", c_dp), ("n_lam1", C.c_int)]


class fb_band_out(C.Structure):
    _fields_ = [("lam1_sig", c_dp), ("nband", C.c_int64), ("band_bin1", c_ip), ("band_bin2", c_ip), ("band_observed", c_dp),
                ("band_binless", c_dp), ("sig_cap", C.c_int64), ("nsig", C.POINTER(C.c_int64)), ("sig_bin1", c_ip),
                ("sig_bin2", c_ip), ("sig_log10", c_dp)]


class fb_gfl_state(C.Structure):
    _fields_ = [("beta", c_dp), ("z", c_dp), ("u", c_dp), ("alpha", C.c_double), ("steps", C.c_int)]


ENTRY_POINTS = ["create", "destroy", "last_error", "run_signal", "restore", "run_difference",
                "eval_cv", "sweep_lambda", "gfl_solve", "set_interrupt_cb", "kernel_launches", "release_cached_memory",
                "comm_unique_id", "comm_init", "comm_destroy", "comm_rank", "comm_world", "shard_plan",
                "create_binned", "get_mat", "mat_rows", "cv_score", "csv_write",
                "wgfl_signal_bic", "wgfl_diff_bic", "create_multi", "handle_ngpus", "gfl_solve_multi",
                "band_export", "gfl_last_solve_ms", "far_median_noalloc"]


def _dp(a):
    return a.ctypes.data_as(c_dp) if a is not None else None


def _ip(a):
    return a.ctypes.data_as(c_ip) if a is not None else None


def _i32(a):
    return np.ascontiguousarray(a, dtype=np.int32)


def _f64(a):
    return np.ascontiguousarray(a, dtype=np.float64)


class FastBinlessError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(msg)
        self.code = code


class fb_reads(C.Structure):
    _fields_ = [("ndatasets", C.c_int), ("resolution", C.c_int),
                ("ncuts", C.c_int64), ("cut_name", c_ip), ("cut_pos", c_ip),
                ("nreads", C.c_int64), ("rd_name", c_ip), ("rd_pos1", c_ip), ("rd_pos2", c_ip),
                ("contact_close", c_ip), ("contact_far", c_ip), ("contact_up", c_ip), ("contact_down", c_ip)]


class FastBinlessLib:
    """One loaded library exporting the fastbinless C ABI under ``prefix``."""

    def __init__(self, path: str, prefix: str = "fb_"):
        self.path = path
        self.prefix = prefix
        self.lib = C.CDLL(path)
        def f(n):
            try:
                return getattr(self.lib, prefix + n)
            except AttributeError:  # a partial library (development builds): fail at call time
                def missing(*a, **k):
                    raise FastBinlessError(-1, f"{path} does not export {prefix}{n}")
                return missing
        self._create = f("create")
        self._create.argtypes = [C.c_int, C.POINTER(fb_mat), C.POINTER(C.c_void_p)]
        self._create.restype = C.c_int
        self._destroy = f("destroy"); self._destroy.argtypes = [C.c_void_p]; self._destroy.restype = None
        self._last_error = f("last_error"); self._last_error.argtypes = [C.c_void_p]; self._last_error.restype = C.c_char_p
        self._run_signal = f("run_signal")
        self._run_signal.argtypes = [C.c_void_p, C.POINTER(fb_params), C.POINTER(fb_signal_out)]
        self._run_signal.restype = C.c_int
        self._restore = f("restore")
        self._restore.argtypes = [C.c_void_p, c_dp, C.POINTER(fb_mean_state), C.POINTER(fb_gam_state), C.POINTER(fb_mean_state)]
        self._restore.restype = C.c_int
        self._run_difference = f("run_difference")
        self._run_difference.argtypes = [C.c_void_p, C.c_int, C.POINTER(fb_params), C.POINTER(fb_diff_out)]
        self._run_difference.restype = C.c_int
        self._eval_cv = f("eval_cv")
        self._eval_cv.argtypes = [C.c_void_p, C.POINTER(fb_params), C.c_int, c_dp, C.c_uint, C.POINTER(fb_cv_out)]
        self._eval_cv.restype = C.c_int
        self._sweep_lambda = f("sweep_lambda")
        self._sweep_lambda.argtypes = [C.c_void_p, C.POINTER(fb_params), C.c_int, c_dp, C.POINTER(fb_sweep_out)]
        self._sweep_lambda.restype = C.c_int
        self._cv_score = f("cv_score")
        self._cv_score.argtypes = [C.c_void_p, C.POINTER(fb_params), C.c_int, c_dp, C.POINTER(fb_cv_score_out)]
        self._cv_score.restype = C.c_int
        self._csv_write = f("csv_write")
        self._csv_write.argtypes = [C.c_char_p, C.c_char_p, C.c_int64, C.c_int, C.POINTER(c_dp), C.c_int]
        self._csv_write.restype = C.c_int
        self._wgfl_signal_bic = f("wgfl_signal_bic")
        self._wgfl_signal_bic.argtypes = [C.c_int, C.POINTER(fb_cts), C.POINTER(fb_perf_meta), C.POINTER(fb_perf_params), c_dp,
                                          C.c_int, C.POINTER(fb_gfl_state), C.POINTER(fb_perf_signal_out)]
        self._wgfl_signal_bic.restype = C.c_int
        self._wgfl_diff_bic = f("wgfl_diff_bic")
        self._wgfl_diff_bic.argtypes = [C.c_int, C.POINTER(fb_cts), C.POINTER(fb_cts), C.POINTER(fb_perf_meta),
                                        C.POINTER(fb_perf_params), c_dp, c_dp, C.c_int, C.POINTER(fb_gfl_state), C.POINTER(fb_perf_diff_out)]
        self._wgfl_diff_bic.restype = C.c_int
        self._create_multi = f("create_multi")
        self._create_multi.argtypes = [C.c_int, C.POINTER(fb_mat), C.c_int, C.POINTER(C.c_void_p)]; self._create_multi.restype = C.c_int
        self._gfl_solve_multi = f("gfl_solve_multi")
        self._gfl_solve_multi.argtypes = [C.c_int, C.c_int, C.c_uint, c_dp, c_dp, C.c_double, C.c_double, C.c_int, C.c_int,
                                          C.POINTER(fb_gfl_state), c_dp]
        self._gfl_solve_multi.restype = C.c_int
        self._band_export = f("band_export")
        self._band_export.argtypes = [C.c_void_p, C.POINTER(fb_band_params), C.POINTER(fb_band_out)]; self._band_export.restype = C.c_int
        self._gfl_last_solve_ms = f("gfl_last_solve_ms"); self._gfl_last_solve_ms.argtypes = []; self._gfl_last_solve_ms.restype = C.c_double
        self._far_median_noalloc = f("far_median_noalloc")
        self._far_median_noalloc.argtypes = [C.c_int, C.c_void_p, C.c_double, C.c_int, C.c_void_p, C.c_void_p]
        self._far_median_noalloc.restype = C.c_int
        self._kernel_launches = f("kernel_launches")
        self._kernel_launches.argtypes = [C.c_void_p]; self._kernel_launches.restype = C.c_int64
        self._gfl_solve = f("gfl_solve")
        self._gfl_solve.argtypes = [C.c_int, C.c_uint, c_dp, c_dp, C.c_double, C.c_double, C.c_int, C.c_int,
                                    C.POINTER(fb_gfl_state), c_dp]
        self._gfl_solve.restype = C.c_int
        self._create_binned = f("create_binned")
        self._create_binned.argtypes = [C.POINTER(fb_reads), C.POINTER(C.c_int), C.POINTER(C.c_void_p)]; self._create_binned.restype = C.c_int
        self._get_mat = f("get_mat")
        self._get_mat.argtypes = [C.c_void_p, c_ip, c_ip, c_ip, c_ip, c_ip, c_dp, c_ip, c_ip]; self._get_mat.restype = C.c_int
        self._mat_rows = f("mat_rows"); self._mat_rows.argtypes = [C.c_void_p]; self._mat_rows.restype = C.c_int64
        self._comm_unique_id = f("comm_unique_id"); self._comm_unique_id.argtypes = [C.c_void_p]; self._comm_unique_id.restype = C.c_int
        self._comm_init = f("comm_init"); self._comm_init.argtypes = [C.c_int, C.c_int, C.c_void_p]; self._comm_init.restype = C.c_int
        self._comm_destroy = f("comm_destroy"); self._comm_destroy.argtypes = []; self._comm_destroy.restype = C.c_int
        self._shard_plan = f("shard_plan")
        self._shard_plan.argtypes = [C.c_int, C.c_uint, C.c_int, c_ip, c_ip]; self._shard_plan.restype = C.c_int

    def far_median_noalloc(self, nbins, log_signal_ptr, lam1, far_from_diag, median_ptr, stream=None):
        """fb_far_median_noalloc: the far-cell median of band_export on caller-owned DEVICE buffers (raw addresses, e.g.
        ``tensor.data_ptr()``); enqueues on ``stream`` (a raw cudaStream_t, None = default stream) and allocates nothing.
        With the oracle library the two addresses are host pointers."""
        self._check(self._far_median_noalloc(int(nbins), C.c_void_p(int(log_signal_ptr)), float(lam1), int(far_from_diag),
                                             C.c_void_p(int(median_ptr)), C.c_void_p(int(stream or 0))), None)

    def csv_write(self, path, header, cols, gzip=True):
        """The reference's csvfile for caller-owned columns (fb_csv_write): gzip as built, or plain text."""
        cols = [_f64(c) for c in cols]
        arr = (c_dp * len(cols))(*[_dp(c) for c in cols])
        self._check(self._csv_write(str(path).encode(), header.encode(), int(cols[0].size), len(cols), arr, int(bool(gzip))), None)

    # ---- multi-GPU (one process per GPU) --------------------------------
    def shard_plan(self, nbins, world, maxdiag=0):
        """(row_begin, col_begin): the equal-cell row / column blocks of the trail-block sharding (host only)."""
        rb = np.zeros(world + 1, dtype=np.int32); cb = np.zeros(world + 1, dtype=np.int32)
        self._check(self._shard_plan(int(nbins), int(maxdiag), int(world), _ip(rb), _ip(cb)), None)
        return rb, cb

    def comm_unique_id(self) -> bytes:
        buf = C.create_string_buffer(128)
        self._check(self._comm_unique_id(buf), None)
        return buf.raw

    def comm_init(self, rank, world, uid: bytes = b""):
        buf = C.create_string_buffer(uid, 128) if uid else None
        self._check(self._comm_init(int(rank), int(world), buf), None)

    def comm_destroy(self):
        self._check(self._comm_destroy(), None)

    # ---- helpers -------------------------------------------------------
    def _check(self, rc, h=None):
        if rc != 0:
            msg = self._last_error(h)
            raise FastBinlessError(rc, (msg or b"unknown error").decode())

    @staticmethod
    def _params(alpha, lam2, lam1, nouter=25, tol_val=0.2, bg_steps=5, free_decay=10000,
                compute_patchnos=True, csv_out="", maxdiag=0, verbose=False):
        lam2a = _f64(np.atleast_1d(lam2)); lam1a = _f64(np.atleast_1d(lam1))
        p = fb_params(alpha=float(alpha), lam2=_dp(lam2a), n_lam2=lam2a.size, lam1=_dp(lam1a),
                      n_lam1=lam1a.size, nouter=int(nouter), tol_val=float(tol_val), bg_steps=int(bg_steps),
                      free_decay=int(free_decay), compute_patchnos=int(bool(compute_patchnos)),
                      csv_out=(csv_out or "").encode(), maxdiag=int(maxdiag), verbose=int(bool(verbose)))
        p._keep = (lam2a, lam1a)
        return p

    def create_binned(self, reads, resolution, ndatasets, fetch_mat=True):
        """bin_data() of a CSnorm-like object: ``reads`` holds cut_name, cut_pos (obj@biases) and rd_name, rd_pos1,
        rd_pos2, contact_close / far / up / down (obj@counts). Returns a Handle whose matrix was built by the library."""
        a = {k: _i32(reads[k]) for k in ("cut_name", "cut_pos", "rd_name", "rd_pos1", "rd_pos2", "contact_close",
                                         "contact_far", "contact_up", "contact_down")}
        r = fb_reads(ndatasets=int(ndatasets), resolution=int(resolution), ncuts=a["cut_pos"].size, nreads=a["rd_pos1"].size,
                     **{k: _ip(v) for k, v in a.items()})
        h = C.c_void_p(); nb = C.c_int(0)
        self._check(self._create_binned(C.byref(r), C.byref(nb), C.byref(h)), None)
        N = int(self._mat_rows(h))
        if not fetch_mat:     # chromosome scale: the matrix stays in HBM (what on-device bin_data is for)
            hd = Handle(self, h, {"distance": np.empty(0)}, nb.value)
            hd.N = N; hd.ndatasets = N // hd.ncells
            return hd
        cols = {"distance": np.empty(N)}
        for k in ("name", "bin1", "pos1", "bin2", "pos2", "observed", "nobs"):
            cols[k] = np.empty(N, dtype=np.int32)
        self._check(self._get_mat(h, _ip(cols["name"]), _ip(cols["bin1"]), _ip(cols["pos1"]), _ip(cols["bin2"]), _ip(cols["pos2"]),
                                  _dp(cols["distance"]), _ip(cols["observed"]), _ip(cols["nobs"])), None)
        return Handle(self, h, cols, nb.value)

    def create(self, mat, nbins, ngpus=1):
        """fb_create, or fb_create_multi when ngpus > 1 (one process, one host thread and communicator per device)."""
        cols = {}
        for k in ("name", "bin1", "pos1", "bin2", "pos2", "observed", "nobs"):
            cols[k] = _i32(mat[k]) if mat.get(k) is not None else None
        cols["distance"] = _f64(mat["distance"])
        m = fb_mat(N=cols["distance"].size, distance=_dp(cols["distance"]),
                   **{k: _ip(v) for k, v in cols.items() if k != "distance"})
        h = C.c_void_p()
        rc = self._create(int(nbins), C.byref(m), C.byref(h)) if int(ngpus) <= 1 else \
            self._create_multi(int(nbins), C.byref(m), int(ngpus), C.byref(h))
        self._check(rc, None)
        return Handle(self, h, cols, int(nbins))


class Handle:
    """An fb_handle plus the host columns it was created from."""

    def __init__(self, lib: FastBinlessLib, h, cols, nbins):
        self.lib, self.h, self.cols, self.nbins = lib, h, cols, nbins
        self.N = cols["distance"].size
        self.ncells = nbins * (nbins + 1) // 2
        self.ndatasets = self.N // self.ncells

    def close(self):
        if self.h:
            self.lib._destroy(self.h)
            self.h = None

    def kernel_launches(self):
        return int(self.lib._kernel_launches(self.h))

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def run_signal(self, alpha, lam2, lam1=0.0, columns=True, buffers=None, **kw):
        """``columns``: True = all data-frame columns, False = only log_signal + leg states, "none" = leg states
        only (results stay on the device). ``buffers`` may supply preallocated (e.g. pinned) output arrays."""
        p = self.lib._params(alpha, lam2, lam1, **kw)
        N = self.N
        buffers = buffers or {}
        names = ("background", "biasmat", "decaymat", "signal", "binless", "phihat", "weight")
        arr = {k: buffers.get(k, None) if buffers.get(k, None) is not None else np.empty(N) for k in names} if columns is True else {}
        patch = (buffers.get("patchno") if buffers.get("patchno") is not None else np.empty(N, dtype=np.int32)) if columns is True else None
        log_signal = None if columns == "none" else (buffers.get("log_signal") if buffers.get("log_signal") is not None else np.empty(N))
        b_est = np.empty(self.nbins); e_est = np.empty(self.ndatasets)
        cap = max(1, self.ndatasets * int(p.nouter))
        iters = np.zeros(cap, dtype=np.int32)
        out = fb_signal_out(patchno=_ip(patch), log_signal=_dp(log_signal),
                            solve_iters=_ip(iters), solve_iters_cap=cap,
                            **{k: _dp(v) for k, v in arr.items()})
        out.biases.estimate = _dp(b_est); out.biases.n = self.nbins
        out.exposures.estimate = _dp(e_est); out.exposures.n = self.ndatasets
        rc = self.lib._run_signal(self.h, C.byref(p), C.byref(out))
        self.lib._check(rc, self.h)
        res = dict(arr)
        if columns is True:
            res["patchno"] = patch
        res.update(log_signal=log_signal,
                   biases=dict(estimate=b_est[:out.biases.n].copy(), mean=out.biases.mean),
                   decay=dict(beta=np.array(out.decay.beta[:]), lambda_=out.decay.lambda_, mean=out.decay.mean),
                   exposures=dict(estimate=e_est[:out.exposures.n].copy(), mean=out.exposures.mean),
                   nbins=self.nbins,
                   diagnostics=dict(nsteps=out.nsteps, converged=bool(out.converged),
                                    prec_abs=np.array(out.prec_abs[:min(out.nsteps, FB_MAX_OUTER)]),
                                    prec_rel=np.array(out.prec_rel[:min(out.nsteps, FB_MAX_OUTER)]),
                                    admm_iters=out.admm_iters, admm_cell_iters=out.admm_cell_iters,
                                    n_solves=out.n_solves, solve_iters=iters[:out.n_solves].copy(),
                                    ms_total=out.ms_total, ms_solver=out.ms_solver))
        return res

    def restore(self, log_signal, biases, decay, exposures):
        ls = _f64(log_signal)
        be = _f64(biases["estimate"]); ee = _f64(exposures["estimate"])
        bs = fb_mean_state(estimate=_dp(be), n=be.size, mean=float(biases["mean"]))
        es = fb_mean_state(estimate=_dp(ee), n=ee.size, mean=float(exposures["mean"]))
        gs = fb_gam_state(lambda_=float(decay["lambda_"]), mean=float(decay["mean"]))
        for i, v in enumerate(np.asarray(decay["beta"], dtype=np.float64)):
            gs.beta[i] = v
        rc = self.lib._restore(self.h, _dp(ls), C.byref(bs), C.byref(gs), C.byref(es))
        self.lib._check(rc, self.h)

    def run_difference(self, ref, alpha, lam2, lam1=0.0, tol_val=0.2, compute_patchnos=True, maxdiag=0, verbose=False):
        p = self.lib._params(alpha, lam2, lam1, tol_val=tol_val, compute_patchnos=compute_patchnos,
                             maxdiag=maxdiag, verbose=verbose)
        N = self.N
        arr = {k: np.empty(N) for k in ("difference", "deltahat", "weight", "phi_ref")}
        patch = np.empty(N, dtype=np.int32)
        cap = max(1, self.ndatasets)
        iters = np.zeros(cap, dtype=np.int32)
        out = fb_diff_out(patchno=_ip(patch), solve_iters=_ip(iters), solve_iters_cap=cap,
                          **{k: _dp(v) for k, v in arr.items()})
        rc = self.lib._run_difference(self.h, int(ref), C.byref(p), C.byref(out))
        self.lib._check(rc, self.h)
        res = dict(arr); res["patchno"] = patch
        res["diagnostics"] = dict(admm_iters=out.admm_iters, admm_cell_iters=out.admm_cell_iters,
                                  n_solves=out.n_solves, solve_iters=iters[:out.n_solves].copy(),
                                  ms_total=out.ms_total, ms_solver=out.ms_solver)
        return res

    def eval_cv(self, alpha, lam2, lam1=0.0, group=0, tol_val=0.1, maxdiag=0):
        lam2a = _f64(np.atleast_1d(lam2))
        p = self.lib._params(alpha, lam2a, lam1, tol_val=tol_val, maxdiag=maxdiag)
        N = self.N
        outs = (fb_cv_out * lam2a.size)()
        keep = []
        for o in outs:
            a = {k: np.empty(N) for k in ("signal", "binless", "phihat", "weight", "log_signal")}
            pt = np.empty(N, dtype=np.int32)
            for k, v in a.items():
                setattr(o, k, _dp(v))
            o.patchno = _ip(pt)
            a["patchno"] = pt
            keep.append(a)
        rc = self.lib._eval_cv(self.h, C.byref(p), lam2a.size, _dp(lam2a), int(group), outs)
        self.lib._check(rc, self.h)
        return keep

    def sweep_lambda(self, alpha, lam2, lam1=0.0, tol_val=0.2, maxdiag=0):
        """lambda2 / BIC sweep (fb_sweep_lambda): arrays of shape (len(lam2), ndatasets)."""
        lam2a = _f64(np.atleast_1d(lam2))
        p = self.lib._params(alpha, lam2a[:1] if lam2a.size else [0.0], lam1, tol_val=tol_val, maxdiag=maxdiag)
        shape = (lam2a.size, self.ndatasets)
        bic = np.empty(shape); chisq = np.empty(shape)
        dof = np.empty(shape, dtype=np.int32); steps = np.empty(shape, dtype=np.int32)
        o = fb_sweep_out(bic=_dp(bic), chisq=_dp(chisq), dof=_ip(dof), steps=_ip(steps))
        rc = self.lib._sweep_lambda(self.h, C.byref(p), lam2a.size, _dp(lam2a), C.byref(o))
        self.lib._check(rc, self.h)
        return {"lam2": lam2a, "bic": bic, "chisq": chisq, "dof": dof, "steps": steps, "ms_solver": o.ms_solver,
                "admm_cell_iters": int(o.admm_cell_iters)}


def _cv_score_impl(self, alpha, lam2, lam1=0.0, tol_val=0.2, maxdiag=0):
    """Checkerboard CV score per (lambda2, dataset) (fb_cv_score): cv, cv_sd of shape (len(lam2), ndatasets)."""
    lam2a = _f64(np.atleast_1d(lam2))
    p = self.lib._params(alpha, lam2a[:1] if lam2a.size else [0.0], lam1, tol_val=tol_val, maxdiag=maxdiag)
    shape = (lam2a.size, self.ndatasets)
    cv = np.empty(shape); sd = np.empty(shape); steps = np.empty(shape + (2,), dtype=np.int32)
    o = fb_cv_score_out(cv=_dp(cv), cv_sd=_dp(sd), steps=_ip(steps))
    rc = self.lib._cv_score(self.h, C.byref(p), lam2a.size, _dp(lam2a), C.byref(o))
    self.lib._check(rc, self.h)
    return {"lam2": lam2a, "cv": cv, "cv_sd": sd, "steps": steps, "ms_solver": o.ms_solver,
            "admm_cell_iters": int(o.admm_cell_iters)}


def _band_export_impl(self, far_cutoff, maxdiag, lam1=0.0, lam1_final=0.0):
    """chromosome_binless post-processing (fb_band_export): lam1_sig per dataset, the band triplets (bin1, bin2, observed,
    binless) and the sparse log10-signal triplets per dataset (R/chromosome.R:262-330)."""
    l1 = _f64(np.atleast_1d(lam1))
    p = fb_band_params(far_cutoff=float(far_cutoff), maxdiag=int(maxdiag), lam1_final=float(lam1_final), lam1=_dp(l1), n_lam1=l1.size)
    D = self.ndatasets
    sig = np.empty(D); nsig = np.zeros(D, dtype=np.int64)
    o = fb_band_out(lam1_sig=_dp(sig), nsig=nsig.ctypes.data_as(C.POINTER(C.c_int64)))
    self.lib._check(self.lib._band_export(self.h, C.byref(p), C.byref(o)), self.h)        # sizes
    nband, tot = int(o.nband), int(nsig.sum())
    b1 = np.empty(nband, dtype=np.int32); b2 = np.empty(nband, dtype=np.int32)
    bo = np.empty((D, nband)); bb = np.empty((D, nband))
    s1 = np.empty(max(tot, 1), dtype=np.int32); s2 = np.empty(max(tot, 1), dtype=np.int32); sv = np.empty(max(tot, 1))
    o = fb_band_out(lam1_sig=_dp(sig), nsig=nsig.ctypes.data_as(C.POINTER(C.c_int64)), band_bin1=_ip(b1), band_bin2=_ip(b2),
                    band_observed=_dp(bo), band_binless=_dp(bb), sig_cap=tot, sig_bin1=_ip(s1), sig_bin2=_ip(s2), sig_log10=_dp(sv))
    self.lib._check(self.lib._band_export(self.h, C.byref(p), C.byref(o)), self.h)
    return {"lam1_sig": sig, "band_bin1": b1, "band_bin2": b2, "band_observed": bo, "band_binless": bb, "nsig": nsig,
            "sig_bin1": s1[:tot], "sig_bin2": s2[:tot], "sig_log10": sv[:tot]}


Handle.cv_score = _cv_score_impl
Handle.band_export = _band_export_impl


def _cts_struct(cts, keep):
    cols = {k: _i32(cts[k]) if k in ("bin1", "bin2") else _f64(cts[k]) for k in ("bin1", "bin2", "count", "lmu_nosig", "nobs", "log_decay")}
    keep.append(cols)
    return fb_cts(N=cols["bin1"].size, bin1=_ip(cols["bin1"]), bin2=_ip(cols["bin2"]), count=_dp(cols["count"]),
                  lmu_nosig=_dp(cols["lmu_nosig"]), nobs=_dp(cols["nobs"]), log_decay=_dp(cols["log_decay"]))


def _meta_struct(md, keep):
    br = _i32(md.get("bad_rows", [])); bd = _i32(md.get("bad_diags", [])); dg = _i32(md["diag_grp"])
    keep.append((br, bd, dg))
    return fb_perf_meta(n_bad_rows=br.size, bad_rows=_ip(br), n_bad_diags=bd.size, bad_diags=_ip(bd), diag_grp=_ip(dg))


def wgfl_signal_bic(lib: FastBinlessLib, cts, dispersion, nouter, nbins, gfl_state, lam2, tol_val, metadata, beta,
                    constrained, fixed, positive=True, lambda1_fixed=False, lambda1_fix_value=0.0):
    """wgfl_signal_BIC (perf_iteration_signal.cpp:22-157) through the C ABI; arguments in the reference's order."""
    keep = []
    nc = nbins * (nbins + 1) // 2
    c = _cts_struct(cts, keep); m = _meta_struct(metadata, keep)
    p = fb_perf_params(dispersion=float(dispersion), nouter=int(nouter), lam2=float(lam2), tol_val=float(tol_val),
                       constrained=int(bool(constrained)), fixed=int(bool(fixed)), positive=int(bool(positive)),
                       lambda1_fixed=int(bool(lambda1_fixed)), lambda1_fix_value=float(lambda1_fix_value))
    b0 = _f64(beta)
    sb, sz, su = (np.zeros(nc), np.zeros(2 * nc - 2), np.zeros(2 * nc - 2))
    warm = bool(gfl_state)
    if warm:
        sb[:] = gfl_state["beta"]; sz[:] = gfl_state["z"]; su[:] = gfl_state["u"]
    st = fb_gfl_state(beta=_dp(sb), z=_dp(sz), u=_dp(su), alpha=float(gfl_state["alpha"]) if warm else 5.0, steps=0)
    cols = {k: np.empty(nc) for k in ("phi", "beta", "phihat", "nobs", "weight")}
    patch = np.empty(nc, dtype=np.int32)
    o = fb_perf_signal_out(patchno=_ip(patch), **{k: _dp(v) for k, v in cols.items()})
    rc = lib._wgfl_signal_bic(int(nbins), C.byref(c), C.byref(m), C.byref(p), _dp(b0), int(warm), C.byref(st), C.byref(o))
    lib._check(rc, None)
    res = dict(cols); res["patchno"] = patch
    res.update(dof=o.dof, converged=bool(o.converged), nouter_done=o.nouter_done, ncandidates=o.ncandidates, BIC=o.BIC,
               BIC_sd=o.BIC_sd, eCprime=o.eCprime, lambda1=o.lambda1, lambda2=float(lam2), admm_iters=int(o.admm_iters),
               GFLState={"beta": sb, "z": sz, "u": su, "alpha": st.alpha})
    return res


def wgfl_diff_bic(lib: FastBinlessLib, cts, ref, dispersion, nouter, nbins, gfl_state, lam2, tol_val, metadata, phi_ref,
                  beta, constrained, lambda1_fixed=False, lambda1_fix_value=0.0):
    """wgfl_diff_BIC (perf_iteration_diff.cpp:20-117) through the C ABI; arguments in the reference's order."""
    keep = []
    nc = nbins * (nbins + 1) // 2
    c = _cts_struct(cts, keep); r = _cts_struct(ref, keep); m = _meta_struct(metadata, keep)
    p = fb_perf_params(dispersion=float(dispersion), nouter=int(nouter), lam2=float(lam2), tol_val=float(tol_val),
                       constrained=int(bool(constrained)), fixed=1, positive=0,
                       lambda1_fixed=int(bool(lambda1_fixed)), lambda1_fix_value=float(lambda1_fix_value))
    b0 = _f64(beta); pr0 = _f64(phi_ref)
    sb, sz, su = (np.zeros(nc), np.zeros(2 * nc - 2), np.zeros(2 * nc - 2))
    warm = bool(gfl_state)
    if warm:
        sb[:] = gfl_state["beta"]; sz[:] = gfl_state["z"]; su[:] = gfl_state["u"]
    st = fb_gfl_state(beta=_dp(sb), z=_dp(sz), u=_dp(su), alpha=float(gfl_state["alpha"]) if warm else 5.0, steps=0)
    names = ("phi_ref", "delta", "beta", "phihat", "phihat_var", "phihat_ref", "phihat_var_ref", "nobs", "deltahat", "weight")
    cols = {k: np.empty(nc) for k in names}
    patch = np.empty(nc, dtype=np.int32)
    o = fb_perf_diff_out(patchno=_ip(patch), **{k: _dp(v) for k, v in cols.items()})
    rc = lib._wgfl_diff_bic(int(nbins), C.byref(c), C.byref(r), C.byref(m), C.byref(p), _dp(pr0), _dp(b0), int(warm), C.byref(st), C.byref(o))
    lib._check(rc, None)
    res = dict(cols); res["patchno"] = patch
    res.update(dof=o.dof, converged=bool(o.converged), nouter_done=o.nouter_done, ncandidates=o.ncandidates, BIC=o.BIC,
               BIC_sd=o.BIC_sd, eCprime=0.0, lambda1=o.lambda1, lambda2=float(lam2), admm_iters=int(o.admm_iters),
               GFLState={"beta": sb, "z": sz, "u": su, "alpha": st.alpha})
    return res


def gfl_solve(lib: FastBinlessLib, nbins, y, w, lam2, converge, maxsteps=30000, maxdiag=0, state=None, ngpus=1):
    """FusedLassoGaussianEstimator<GFLLibrary>::optimize on one packed triangle (ngpus > 1: shared by that many devices of
    this process, fb_gfl_solve_multi)."""
    y = _f64(y); w = _f64(w)
    nc = nbins * (nbins + 1) // 2
    md = min(int(maxdiag), nbins)
    if md > (nbins + 1) // 2 or md <= 1:
        n_solved = nc
        tsz = 2 * nc - 2
    else:
        n_solved = md * (nbins - md + 1) + (md - 1) * md // 2
        nrow_tr = sum(1 for i in range(nbins) if min(md, nbins - i) >= 2)
        tsz = (n_solved - (nbins - nrow_tr)) + (n_solved - 1)
    warm = state is not None
    if warm:
        beta = _f64(state["beta"]).copy(); z = _f64(state["z"]).copy(); u = _f64(state["u"]).copy()
        alpha = float(state["alpha"])
    else:
        beta = np.zeros(n_solved); z = np.zeros(tsz); u = np.zeros(tsz); alpha = 5.0
    st = fb_gfl_state(beta=_dp(beta), z=_dp(z), u=_dp(u), alpha=alpha, steps=0)
    full = np.empty(nc)
    if int(ngpus) > 1:
        rc = lib._gfl_solve_multi(int(ngpus), int(nbins), int(maxdiag), _dp(y), _dp(w), float(lam2), float(converge), int(maxsteps),
                                  int(warm), C.byref(st), _dp(full))
    else:
        rc = lib._gfl_solve(int(nbins), int(maxdiag), _dp(y), _dp(w), float(lam2), float(converge), int(maxsteps),
                            int(warm), C.byref(st), _dp(full))
    lib._check(rc, None)
    return dict(beta=beta, z=z, u=u, alpha=st.alpha, steps=st.steps, beta_full=full)
