/*
 * fastbinless.h -- C ABI of the B200-native fast-binless library
 * (libfastbinless_cuda.so).
 *
 * This is the drop-in boundary for the two R entry points that the reference
 * registers in RCPP_MODULE(binless_cpp):
 *
 *   fast_binless            -> binless::fast::binless()
 *                              /root/reference/binless/src/csnorm_cpp.cpp:18-22,
 *                              /root/reference/binless/src/fast_binless.cpp:85-184
 *   fast_binless_difference -> binless::fast::binless_difference()
 *                              csnorm_cpp.cpp:23-26, fast_binless.cpp:235-279
 *   (binless_eval_cv, fast_binless.cpp:186-233, exists in the reference's C++
 *    but is not registered with R; fb_eval_cv exposes it.)
 *
 * An Rcpp shim (r/fastbinless_shim.cpp) converts the R
 * DataFrame / List arguments to the plain buffers below and back, so the R
 * signatures and the returned list / data-frame layout stay unchanged.
 *
 * Conventions
 *   - Plain pointers and sizes only. The caller owns every host buffer passed
 *     in or out; the library owns all device memory behind fb_handle.
 *   - Every function returns 0 on success; non-zero means failure and
 *     fb_last_error() gives the message the shim hands to Rcpp::stop().
 *   - There is no CPU fallback: without a CUDA device every compute entry
 *     point fails with FB_ERR_CUDA.
 *   - Input layout is the one FastDataCore requires
 *     (/root/reference/binless/src/FastData.ipp:3-26): N = ndatasets *
 *     nbins*(nbins+1)/2 rows sorted by (name, bin1, bin2), i.e. dataset-major
 *     packed upper triangles in row-major order.
 */
#ifndef FASTBINLESS_H
#define FASTBINLESS_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#ifndef FB_API
#define FB_API __attribute__((visibility("default")))
#endif

/* Allow the CPU oracle (test infrastructure) to export the same ABI under a
 * different prefix (fbo_*) so tests drive both through identical structs. */
#ifndef FB_NAME
#define FB_NAME(x) fb_##x
#endif

enum {
    FB_OK = 0,
    FB_ERR_ARG = 1,     /* invalid argument: message mirrors the reference's Rcpp::stop text */
    FB_ERR_CUDA = 2,    /* CUDA runtime failure or no device */
    FB_ERR_STATE = 3,   /* call order violation (e.g. difference before signal/restore) */
    FB_ERR_NUMERIC = 4, /* e.g. "decomposition failed" (QuadProgGAMLibrary.cpp:21) */
    FB_ERR_INTERRUPT = 5,
    FB_ERR_COMM = 6     /* NCCL missing or a collective failed (multi-GPU only) */
};

#define FB_DECAY_K 50       /* Config<Decay,GAM>::K, fast_decay.hpp:25 */
#define FB_MAX_OUTER 256    /* capacity of the per-step diagnostic arrays */

typedef struct fb_handle fb_handle;

/* The `mat` data frame handed to fast_binless (columns of FastDataCore,
 * FastData.ipp:5-14). name/bin1/bin2 may be NULL, meaning the canonical order;
 * when given they are validated like the reference does. */
typedef struct {
    int64_t N;
    const int32_t *name, *bin1, *pos1, *bin2, *pos2;
    const int32_t *observed, *nobs;
    const double *distance;
} fb_mat;

/* Arguments of fast_binless / fast_binless_difference that are not data
 * (fast_binless.hpp:29-38; defaults csnorm_cpp.cpp:19-25). */
typedef struct {
    double alpha;
    const double *lam2; int n_lam2;   /* broadcast lam2[0] unless n_lam2 == number of solves */
    const double *lam1; int n_lam1;   /* NULL/0 means lam1 = 0 */
    unsigned nouter;                  /* default 25 */
    double tol_val;                   /* default 0.2 */
    unsigned bg_steps;                /* default 5 */
    unsigned free_decay;              /* default 10000 */
    int compute_patchnos;             /* default 1 */
    const char *csv_out;              /* default "" (NULL allowed) */
    unsigned maxdiag;                 /* default 0 = full triangle */
    int verbose;                      /* 1: print the reference's progress lines to stdout */
} fb_params;

/* State of one Mean leg (Params<Mean>, fast_fitter_mean.hpp:15-33) */
typedef struct { double *estimate; int n; double mean; } fb_mean_state;
/* State of the GAM decay leg (Params<GAM>, fast_fitter_gam.hpp:17-41) */
typedef struct { double beta[FB_DECAY_K]; double lambda; double mean; } fb_gam_state;

/* Outputs of fast_binless: the 16-column `mat` data frame minus the 8 input
 * columns (fast_dataframe.cpp:90-105) plus the list members biases / decay /
 * exposures / log_signal (fast_binless.cpp:179-183). All arrays are caller
 * allocated with N entries unless stated; any pointer may be NULL to skip
 * that column. */
typedef struct {
    double *background, *biasmat, *decaymat, *signal, *binless, *phihat, *weight;
    int32_t *patchno;            /* -1 everywhere when compute_patchnos == 0 */
    double *log_signal;
    fb_mean_state biases;        /* estimate has nbins entries */
    fb_gam_state decay;
    fb_mean_state exposures;     /* estimate has ndatasets entries */
    /* diagnostics (not part of the R return value) */
    int nsteps;                  /* outer steps performed */
    int converged;
    double prec_abs[FB_MAX_OUTER], prec_rel[FB_MAX_OUTER]; /* "reached precision" per step */
    int64_t admm_iters;          /* total ADMM iterations over all solves */
    int64_t admm_cell_iters;     /* sum over solves of cells_solved * iterations */
    int64_t n_solves;
    int32_t *solve_iters;        /* optional: iterations of each solve, in call order */
    int64_t solve_iters_cap;
    double ms_total, ms_solver;  /* device-timed, for benchmarking */
} fb_signal_out;

/* Outputs of fast_binless_difference: the 12-column data frame minus the 7
 * input columns (fast_dataframe.cpp:133-144). */
typedef struct {
    double *difference, *deltahat, *weight, *phi_ref;
    int32_t *patchno;
    int64_t admm_iters, admm_cell_iters, n_solves;
    int32_t *solve_iters; int64_t solve_iters_cap;
    double ms_total, ms_solver;
} fb_diff_out;

/* One element of the list binless_eval_cv returns (fast_binless.cpp:215-229):
 * the `mat` columns for one lambda2. Arrays have N entries. */
typedef struct {
    double *signal, *binless, *phihat, *weight, *log_signal;
    int32_t *patchno;
} fb_cv_out;

/* Input side: bin_data() for a CSnorm object, /root/reference/binless/R/preprocess.R:104-148
 * with add_bin_bounds_and_distance(), R/common.R:115-132, done on the device so that the dense matrix never exists on
 * the host. `cut_*` is obj@biases (one row per cut site: dataset 1..ndatasets and position), `rd_*` is obj@counts (one
 * row per pair of cut sites with its four contact counts). Bins: borders b0 + k*resolution from b0 = min(pos) - 1
 * while <= max(pos) + 1, the last incomplete bin discarded, intervals closed on the left, the last one also on the right;
 * nobs = 4 * cuts(bin1) * cuts(bin2), halved (integer) on the diagonal; observed = sum of the four counts; a cell without
 * a read row gets observed = nobs = 0 (a row whose four counts are zero keeps nobs); observed[(1, nbins)] += 1 in every dataset; pos = (begin + end) / 2 (integer),
 * distance = begin2 - begin1 + resolution / 2. */
typedef struct {
    int ndatasets, resolution;
    int64_t ncuts;  const int32_t *cut_name, *cut_pos;
    int64_t nreads; const int32_t *rd_name, *rd_pos1, *rd_pos2;
    const int32_t *contact_close, *contact_far, *contact_up, *contact_down;
} fb_reads;
/* the same handle fb_create returns, built from reads; *nbins receives the number of bins */
FB_API int FB_NAME(create_binned)(const fb_reads *reads, int *nbins, fb_handle **out);
/* the `mat` columns of a handle (what bin_data returns to R); arrays of N entries, any pointer may be NULL */
FB_API int FB_NAME(get_mat)(fb_handle *h, int32_t *name, int32_t *bin1, int32_t *pos1, int32_t *bin2, int32_t *pos2,
                            double *distance, int32_t *observed, int32_t *nobs);
FB_API int64_t FB_NAME(mat_rows)(const fb_handle *h);

/* FastDataCore constructor (FastData.ipp:3-26): validates and uploads. */
FB_API int FB_NAME(create)(int nbins, const fb_mat *mat, fb_handle **out);
FB_API void FB_NAME(destroy)(fb_handle *h);
FB_API const char *FB_NAME(last_error)(const fb_handle *h); /* h may be NULL: last create error */

/* fast_binless (fast_binless.cpp:85-184) */
FB_API int FB_NAME(run_signal)(fb_handle *h, const fb_params *p, fb_signal_out *out);

/* Rebuild estimators from a saved fast_binless result, as binless_difference
 * and binless_eval_cv do with set_state (fast_binless.cpp:244-260). */
FB_API int FB_NAME(restore)(fb_handle *h, const double *log_signal,
                            const fb_mean_state *biases, const fb_gam_state *decay,
                            const fb_mean_state *exposures);

/* fast_binless_difference (fast_binless.cpp:235-279); ref is 1-based. */
FB_API int FB_NAME(run_difference)(fb_handle *h, int ref, const fb_params *p, fb_diff_out *out);

/* binless_eval_cv (fast_binless.cpp:186-233): one dataset, nlam lambda2 values
 * evaluated in order with the warm-start chain of the reference; group is the
 * checkerboard hold-out group (0 = none). outs[nlam]. Every lambda starts from the log_signal given to fb_restore: the
 * reference re-seeds from mat$signal, the exp'd and thresholded column (fast_binless.cpp:199,222); pass that column to
 * fb_restore to reproduce it call for call. */
FB_API int FB_NAME(eval_cv)(fb_handle *h, const fb_params *p, int nlam, const double *lam2,
                            unsigned group, fb_cv_out *outs);

/* lambda2 / BIC sweep on a fitted or restored model (BASELINE.json config 3): for every lambda2 value and every
 * dataset one COLD-START fused-lasso solve of the signal step (fresh FusedLassoGaussianEstimator, alpha = 5, as
 * ScorePreparator<BIC> does per candidate: ScorePreparator.ipp:53-64), the solution centred like step_signal
 * (fast_binless.ipp:33-44), and the BIC of the optimized path evaluated on the fast path's IRLS quantities:
 *   chisq = sum_cells [ w (phihat - soft(beta, lam1))^2 + lam1^2 ]     Likelihoods.hpp:25-42 (eCprime = 0)
 *   dof   = number of patches (get_patch_numbers at tol_val on beta) holding a cell with |soft(beta, lam1)| > tol_val/2
 *                                                                      DOFComputer.hpp:16-24
 *   bic   = chisq + log(sum nobs of the dataset) * dof                 ScoreAssembler.hpp:41-55
 * The points are independent of each other, so a caller shards the lambda2 list over processes / GPUs.
 * Arrays hold nlam * ndatasets entries, lambda-major; any pointer may be NULL. */
typedef struct {
    double *bic, *chisq;
    int32_t *dof, *steps;        /* steps = return value of graph_fused_lasso_weight_warm per solve */
    double ms_solver;
    int64_t admm_cell_iters;
} fb_sweep_out;
FB_API int FB_NAME(sweep_lambda)(fb_handle *h, const fb_params *p, int nlam, const double *lam2, fb_sweep_out *out);

/* Checkerboard cross-validation score of lambda2 values on a fitted or restored model: what a
 * lambda2 search minimises. Per lambda2 value and dataset, for each of the two groups g = (bin1 + bin2) % 2 a COLD-START
 * solve with the cells of the group held out (y = -100, w = 0: fast_binless.ipp:14-22; ScorePreparator.ipp:20-49), the
 * solution centred like step_signal (fast_binless.ipp:33-44) and read at the held-out cells; then
 *   chisq_i = w_i (phihat_i - soft(beta_cv_i, lam1))^2 + lam1^2        Likelihoods.hpp:25-42 (eCprime = 0), w / phihat not held out
 *   CV_g = sum_{i in g} chisq_i / N_g,  CV = sum_g N_g CV_g / N,  CV_sd = sqrt(sum_g N_g (CV_g - CV)^2 / (N - 1))
 *                                                                      ScoreAssembler.hpp:15-39
 * cv / cv_sd hold nlam * ndatasets entries (lambda-major), steps 2 per entry (group 0, group 1); any pointer may be NULL. */
typedef struct {
    double *cv, *cv_sd;
    int32_t *steps;
    double ms_solver;
    int64_t admm_cell_iters;
} fb_cv_score_out;
FB_API int FB_NAME(cv_score)(fb_handle *h, const fb_params *p, int nlam, const double *lam2, fb_cv_score_out *out);

/* The csvfile writer of the reference (csvfile.hpp:30-117) for caller-owned columns: `header` (already quoted), then
 * nrows rows of ncols values printed with 4 decimals. gzip != 0: one gzip member, what the reference as built writes
 * (USE_BOOST = 1, binless/src/Makevars:1-2); 0: plain text (USE_BOOST = 0). Host only, no device needed. */
FB_API int FB_NAME(csv_write)(const char *path, const char *header, int64_t nrows, int ncols, const double *const *cols, int gzip);

/* Stand-alone fused-lasso solve on the triangle grid (the
 * FusedLassoGaussianEstimator<GFLLibrary> object of the reference,
 * FusedLassoGaussianEstimator.hpp:16-71, GFLLibraryTriangleImpl.cpp:60-71).
 * Used by the lambda2/BIC sweep and by the solver benchmarks: y, w, beta are
 * packed triangles of nbins*(nbins+1)/2 cells; state 0 = cold start, otherwise
 * z/u/alpha are read and written back (trail-slot order of the reference:
 * rows then columns, tsz = 2*ncells-2 entries, or the trapeze order). */
typedef struct {
    double *beta;       /* in (warm) / out, ncells of the solved grid */
    double *z, *u;      /* in/out, tsz entries, may be NULL for cold start + discard */
    double alpha;       /* in/out */
    int steps;          /* out: return value of graph_fused_lasso_weight_warm */
} fb_gfl_state;
FB_API int FB_NAME(gfl_solve)(int nbins, unsigned maxdiag, const double *y, const double *w,
                              double lam2, double converge, int maxsteps, int warm,
                              fb_gfl_state *st, double *beta_full);

/* ---- Optimized-path performance iteration on the same fused-lasso kernels ----
 * wgfl_signal_BIC (perf_iteration_signal.cpp:22-157) and wgfl_diff_BIC (perf_iteration_diff.cpp:20-117) of the reference:
 * the IRLS alternation of IRLSEstimator.hpp:28-60 between the scatter-add front end cts_to_signal_mat_core
 * (cts_core.c:5-47, through cts_to_mat.cpp:14-124 with its outlier rows / diagonals) and the triangle-grid fused lasso
 * (warm first, then a cold restart if it did not converge), the patch numbers of the result, and the lambda1 / eCprime
 * choice by BIC: EstimatedSparsity.ipp:11-30 over the candidates of candidate_generation.cpp / minimum_UB.cpp /
 * BoundsOptimizer.ipp, or FixedSparsity.hpp with BoundsRectifier.hpp; the likelihood is evaluated on a COLD-START
 * re-solve at lambda2 (ScorePreparator.ipp:53-64, Likelihoods.hpp:25-95), the degrees of freedom on the warm solution
 * (DOFComputer.hpp:16-24), BIC = sum chisq + log(sum nobs) dof (ScoreAssembler.hpp:41-55).
 * `cts` is the reference's cts data frame: one row per (bin pair, cut-site category) with 1-based bins, bin2 >= bin1. */
typedef struct {
    int64_t N;
    const int32_t *bin1, *bin2;
    const double *count, *lmu_nosig, *nobs, *log_decay;
} fb_cts;
typedef struct {                        /* the `metadata` list (cts_to_mat.cpp:14-32,86-90) */
    int n_bad_rows;  const int32_t *bad_rows;      /* bins (1-based) whose rows and columns are outliers */
    int n_bad_diags; const int32_t *bad_diags;     /* outlier diagonals, as bin2 - bin1 */
    const int32_t *diag_grp;                       /* nbins entries: group of diagonal bin2 - bin1 */
} fb_perf_meta;
typedef struct {
    double dispersion; int nouter; double lam2, tol_val;
    int constrained;                   /* ForbidDegeneracy / AllowDegeneracy */
    int fixed, positive;               /* signal only: eCprime fixed at 0 (ZeroOffset) / PositiveSign; fixed == 0 implies positive */
    int lambda1_fixed; double lambda1_fix_value;
} fb_perf_params;
/* Arrays have nbins(nbins+1)/2 entries; any may be NULL. `gfl` carries the GFLState list in (read when warm != 0; warm == 0
 * is the empty list: cold start with alpha = 5) and out (the state after the IRLS iterations, before the BIC re-solve
 * resets the solver; arrays that are NULL are not written). */
typedef struct {
    double *phi, *beta;                             /* soft-thresholded and raw solution */
    double *phihat, *nobs, *weight;                 /* mat columns (cts_to_mat.cpp:92-99) */
    int32_t *patchno;
    int dof, converged, nouter_done, ncandidates;
    double BIC, BIC_sd, eCprime, lambda1;
    int64_t admm_iters;                             /* ADMM iterations over all solves of the call (diagnostic) */
} fb_perf_signal_out;
typedef struct {
    double *phi_ref, *delta, *beta;
    double *phihat, *phihat_var, *phihat_ref, *phihat_var_ref, *nobs, *deltahat, *weight;
    int32_t *patchno;
    int dof, converged, nouter_done, ncandidates;
    double BIC, BIC_sd, lambda1;
    int64_t admm_iters;
} fb_perf_diff_out;
FB_API int FB_NAME(wgfl_signal_bic)(int nbins, const fb_cts *cts, const fb_perf_meta *meta, const fb_perf_params *p,
                                    const double *beta_init, int warm, fb_gfl_state *gfl, fb_perf_signal_out *out);
FB_API int FB_NAME(wgfl_diff_bic)(int nbins, const fb_cts *cts, const fb_cts *ref, const fb_perf_meta *meta,
                                  const fb_perf_params *p, const double *phi_ref_init, const double *beta_init,
                                  int warm, fb_gfl_state *gfl, fb_perf_diff_out *out);

/* ---- Output side of chromosome_binless (R/chromosome.R:262-330, signal case) on a fitted or
 * restored handle: lam1.sig = median(log(signal)) over the cells with distance > far_cutoff, per dataset (lam1_final when
 * there is none; :270-276), signal := pmax(signal / exp(lam1.sig), 1) (:277), and the triplets R's sparseMatrix() takes:
 * the band bin2 - bin1 <= maxdiag with `observed` and `binless` (:303-316), and the cells with signal != 1 with
 * log10(signal) (:318-323). `signal` / `binless` are the data frame columns of fast_binless at the given lam1
 * (fast_dataframe.cpp:71-76). Bins are 1-based; order is row-major (bin1, then bin2); the sparse arrays are dataset-major.
 * Call once with the array pointers NULL to learn nband and nsig, then with arrays of that size. */
typedef struct {
    double far_cutoff;
    unsigned maxdiag;
    double lam1_final;
    const double *lam1; int n_lam1;
} fb_band_params;
typedef struct {
    double *lam1_sig;                              /* [ndatasets] */
    int64_t nband;                                 /* out: band cells per dataset */
    int32_t *band_bin1, *band_bin2;                /* [nband] */
    double *band_observed, *band_binless;          /* [ndatasets * nband] */
    int64_t sig_cap;                               /* in: entries the three arrays below can hold */
    int64_t *nsig;                                 /* [ndatasets] out */
    int32_t *sig_bin1, *sig_bin2; double *sig_log10;
} fb_band_out;
FB_API int FB_NAME(band_export)(fb_handle *h, const fb_band_params *p, fb_band_out *out);

/* The far-cell median of fb_band_export (lam1_sig above) without any allocation. This entry point has no counterpart in
 * the reference; it was added for callers that cannot call cudaMalloc (code
 * under CUDA graph capture, tight serving loops, harnesses that forbid it).
 *   *median_dev = median over the cells with bin2 - bin1 >= far_from_diag of log(exp(soft_threshold(log_signal, lam1))),
 * NaN when there is no such cell (R's NA). log_signal_dev (one dataset: the packed row-major triangle of
 * nbins(nbins+1)/2 doubles, the log_signal column of fb_signal_out / fb_restore) and median_dev (one double) are DEVICE
 * pointers owned by the caller, on the current device; `stream` is a cudaStream_t (NULL: the default stream). The call
 * only enqueues kernels on that stream and returns: no cudaMalloc, cudaFree, copy or synchronisation; the median is in
 * *median_dev when the stream gets there. With a bin_data() layout far_from_diag is the first diagonal whose distance
 * exceeds far.cutoff, and the result equals lam1_sig of fb_band_export bit for bit.
 * Scratch (the 65536-bin digit histogram of the radix select and its state) is a static device array of
 * FB_NOALLOC_SCRATCH_BYTES that every GPU which loads the library holds until the process exits. Limit: nbins <=
 * FB_NOALLOC_MAX_NBINS (32-bit histogram counters, grid height), beyond it FB_ERR_ARG. All calls on a device share the
 * array: calls whose kernels can overlap on one GPU (different streams, several host threads) race and must be
 * serialised by the caller. */
#define FB_NOALLOC_MAX_NBINS 65535
#define FB_NOALLOC_SCRATCH_BYTES (65536 * 4 + 256)
FB_API int FB_NAME(far_median_noalloc)(int nbins, const double *log_signal_dev, double lam1, int far_from_diag,
                                       double *median_dev, void *stream);

/* Diagnostic: device time (CUDA events on the solver's stream, milliseconds) of the ADMM loop of the last fb_gfl_solve /
 * fb_gfl_solve_multi call made by this host thread: uploads, set-up and state export are outside it. */
FB_API double FB_NAME(gfl_last_solve_ms)(void);

/* Device and pinned-host blocks of destroyed handles are kept in a size-keyed cache (cudaMalloc / cudaFree cost
 * milliseconds per call and fast_binless is called once per locus); this returns them to the driver. */
FB_API void FB_NAME(release_cached_memory)(void);

/* Multi-GPU, one process per GPU ("trail blocks within one matrix"; BASELINE.json configs 4-5).
 * The reference has no counterpart: its solve, graph_fused_lasso_weight_warm (gfl_graph_fl.c:53-213), is one thread.
 * After fb_comm_init every fused-lasso solve of this process (inside fb_run_signal / fb_run_difference / fb_eval_cv /
 * fb_sweep_lambda / fb_gfl_solve) is shared by the `world` processes: the row chains and the column chains of the
 * trail decomposition (GFLLibraryTriangleImpl.cpp:7-35) are split into `world` row blocks and column blocks of equal
 * cell count (fb_shard_plan), rank r runs update_z / update_u (gfl_graph_fl.c:267-327) for the chains of block r, the
 * terms z - u of the cells where its blocks cross those of another rank are exchanged once per iteration
 * (ncclSend / ncclRecv) and the per-chain residual sums are all-reduced, so every rank takes the same convergence
 * and penalty decisions as the single-GPU run and the iterates are identical bit for bit. Every rank must make the
 * same calls with the same inputs and gets the complete outputs. The IRLS legs are replicated, not sharded.
 *   fb_comm_unique_id : rank 0 obtains the 128-byte NCCL id; the host (torch.distributed, MPI, a file) hands it round
 *   fb_comm_init      : joins the communicator on the CURRENT CUDA device; world == 1 switches sharding off
 *   fb_shard_plan     : host only (no device needed): block g owns rows [row_begin[g], row_begin[g+1]) and columns
 *                       [col_begin[g], col_begin[g+1]); arrays of world + 1 entries */
FB_API int FB_NAME(comm_unique_id)(void *id128);
FB_API int FB_NAME(comm_init)(int rank, int world, const void *id128);
FB_API int FB_NAME(comm_destroy)(void);
FB_API int FB_NAME(comm_rank)(void);
FB_API int FB_NAME(comm_world)(void);
FB_API int FB_NAME(shard_plan)(int nbins, unsigned maxdiag, int world, int32_t *row_begin, int32_t *col_begin);

/* Multi-GPU from ONE process (the reference entry point is one R call in one process,
 * csnorm_cpp.cpp:18-22): the handle holds one copy of the matrix per device (devices 0 .. ngpus-1 of the visible ones);
 * every fb_run_* / fb_restore / fb_eval_cv / fb_sweep_lambda / fb_cv_score call on it is executed by ngpus host threads,
 * one per device, each with its own NCCL communicator, sharing every fused-lasso solve exactly as ngpus processes
 * would after fb_comm_init (trail blocks; results identical bit for bit to one GPU). The caller sees rank 0's outputs.
 * ngpus == 1 is fb_create. fb_set_interrupt_cb is refused on such a handle. fb_gfl_solve_multi: the stand-alone solve. */
FB_API int FB_NAME(create_multi)(int nbins, const fb_mat *mat, int ngpus, fb_handle **out);
FB_API int FB_NAME(handle_ngpus)(const fb_handle *h);
FB_API int FB_NAME(gfl_solve_multi)(int ngpus, int nbins, unsigned maxdiag, const double *y, const double *w,
                                    double lam2, double converge, int maxsteps, int warm,
                                    fb_gfl_state *st, double *beta_full);

FB_API int FB_NAME(set_interrupt_cb)(fb_handle *h, int (*cb)(void *), void *arg);

/* Diagnostics: number of CUDA kernels launched by the last run_* / eval_cv call on this handle
 * (0 for the CPU oracle, which exports the same ABI under another prefix). */
FB_API int64_t FB_NAME(kernel_launches)(const fb_handle *h);

#ifdef __cplusplus
}
#endif
#endif /* FASTBINLESS_H */
