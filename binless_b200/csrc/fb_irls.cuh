// fb_irls.cuh -- IRLS passes over the binned upper-triangular matrix (device side).
//
// Kernels replacing, on the GPU, the O(N) passes of the reference driver
// (paths relative to /root/reference/binless/src):
//   get_log_expected              fast_expected.ipp:1-19
//   get_residuals (NB)            fast_residuals.ipp:47-67
//   Summarizer::update_summary    fast_summarizer.ipp:32-51   (binner * vec = segmented sums)
//   set_poisson_lsq_summary       fast_summarizer.ipp:3-30
//   step_signal glue              fast_binless.ipp:2-48
//   get_precision                 fast_binless.cpp:22-33
//   remove_signal_degeneracy      fast_binless.cpp:35-65
//   shift_signal                  fast_binless.cpp:67-83
//   step_difference glue          fast_binless.ipp:51-105
//   get_as_dataframe              fast_dataframe.cpp:9-145
//   get_patch_numbers             graph_helpers.cpp:12-59
//
// Layout: every per-cell column is one array of N = D * ncells entries, dataset
// major, each dataset a row-major packed upper triangle (FastData.ipp:3-26).
// With bin_data() inputs (R/preprocess.R:104-148) pos depends on the bin only and
// distance on the diagonal only; this is verified at ingest, after which the
// per-cell leg lookups are three small tables: brow[n] (bias of the bin that
// row/column r falls in), ddiag[n] (decay of diagonal k), expo[D].
#pragma once
#include "fb_util.cuh"
#include "fb_comm.cuh"

namespace fb {

struct Tri {
    int n = 0, D = 0;
    int64_t ncells = 0;
    __host__ __device__ int64_t N() const { return (int64_t)D * ncells; }
};

struct LegTables {
    const double *brow;   // [n]  binned bias estimate per row/column index
    const double *ddiag;  // [n]  binned decay estimate per diagonal
    const double *expo;   // [D]  binned exposure estimate per dataset
    // per-cell fallback: when cb1 is set, brow is indexed by BIAS BIN and ddiag by DECAY BIN, and every
    // cell carries the int32 bins of pos1 / pos2 / log(distance) (util.cpp:56-111 semantics); else null
    const int32_t *cb1 = nullptr, *cb2 = nullptr, *cdb = nullptr;
};

// The GPUs of a communicator share the REDUCTION passes of the outer loop (every rank holds the whole matrix): rank r reads
// the rows / tile rows / blocks r, r + world, ...; the per-rank results are combined exactly (integer min / max keys; sums
// whose other addends are exact zeros), so the outcome has the bits of the single-GPU pass. Default: one GPU.
struct Share {
    int rank = 0, world = 1;
    Comm *cm = nullptr;
    bool on() const { return cm != nullptr && world > 1; }
};

enum { LEG_EXPO = 0, LEG_BIAS = 1, LEG_DECAY = 2 };
enum { MODE_RESID = 0, MODE_POISSON = 1, MODE_NOBS = 2 };

constexpr int TILE_I = 32, TILE_J = 128, TILE_DIAGS = 160; // TILE_I + TILE_J - 1 = 159 slots used

inline int tiles_i(int n) { return (n + TILE_I - 1) / TILE_I; }
inline int tiles_j(int n) { return (n + TILE_J - 1) / TILE_J; }
// doubles of scratch needed by the partial sums of one leg pass
size_t leg_part_doubles(const Tri &t);

// One pass over all cells accumulating, per item of the leg (dataset / row-or-column index /
// diagonal), the pair (a, b):
//   MODE_RESID   a = w, b = z*w  with (z, w) the NB working residual and weight
//   MODE_POISSON a = observed,   b = nobs * exp(log_expected)
//   MODE_NOBS    a = nobs,       b = 0
// Result: items[2*item], items[2*item+1] for item in [0, D) / [0, n) / [0, n).
// Fixed-shape reduction: same bits on every run.
// The sums have a two-level shape independent of the number of GPUs: tile row tI is in group tI mod LEG_GROUPS; items_g
// (LEG_GROUPS x 2 x nitems doubles) receives the sums of the groups g with g mod grp_world == grp_rank (leg_pass_groups);
// after the other groups have been gathered from their owners, leg_pass_combine adds the groups in order. leg_pass does
// both on one GPU. nitems = D for the exposure leg, n otherwise.
constexpr int LEG_GROUPS = 8;
void leg_pass_groups(int leg, int mode, const Tri &t, LegTables L, double alpha, const int32_t *obs, const int32_t *nobs,
                     const double *phi, double *part, double *items_g, int grp_rank, int grp_world, cudaStream_t st);
void leg_pass_combine(int nitems, const double *items_g, double *items, int grp_world, cudaStream_t st);
void leg_pass(int leg, int mode, const Tri &t, LegTables L, double alpha, const int32_t *obs, const int32_t *nobs,
              const double *phi, double *part, double *items_g, double *items, cudaStream_t st);

// The same sums in the per-cell fallback, for the bias / decay leg: `off` / `list` are the CSR cell lists of the nb leg bins
// (bit 31 of an entry: the cell counts twice), ca / cb two N-double scratch arrays; items[2*bin], items[2*bin+1].
void leg_pass_cells(int mode, const Tri &t, LegTables L, double alpha, const int32_t *obs, const int32_t *nobs, const double *phi,
                    int nb, const int64_t *off, const uint32_t *list, double *ca, double *cb, double *items, cudaStream_t st);

// checkerboard CV: out2 = {sum of chisq, number of cells} over the cells of group g = (bin1 + bin2) % 2 of one dataset;
// partial: 2 * CV_BLOCKS doubles of scratch
constexpr int CV_BLOCKS = 592;
void cv_group_sum(int n, int g, const double *phihat, const double *w, const double *beta, double lam1, double *partial,
                  double *out2, cudaStream_t st);

// y = z + phi, w (step_signal, fast_binless.ipp:13-23); group = CV hold-out (0 = none). With a Share and a band width md the
// cells beyond the band are written only on the rows of this rank (the far-region sums of a band solve are all they feed)
void signal_yw(const Tri &t, LegTables L, double alpha, const int32_t *obs, const int32_t *nobs, const double *phi,
               unsigned group, double *y, double *w, cudaStream_t st, const Share &sh = Share(), int md = 1 << 30);

// beta[d] -= sum(beta*nobs)/sum(nobs) (fast_binless.ipp:33-44); scratch >= 2*1024+2 doubles
void center_dataset(int64_t ncells, const int32_t *nobs_d, double *beta_d, double *scratch, cudaStream_t st, const Share &sh = Share());

// abs = max|eta_new - eta_old|, max eta_new, min eta_new; out3 = device {abs, max, min}
void precision(const Tri &t, LegTables Lold, LegTables Lnew, const double *phi_old, const double *phi_new,
               double *scratch, double *out3, cudaStream_t st, const Share &sh = Share());

// remove_signal_degeneracy / shift_signal, in place on phi; keys: (2n+1)*D int64 scratch
void remove_degeneracy(const Tri &t, double *phi, long long *keys, cudaStream_t st, const Share &sh = Share());
void shift_signal(const Tri &t, double *phi, long long *keys, cudaStream_t st, const Share &sh = Share());

// Signal data frame columns (fast_dataframe.cpp:62-88); lam1 is a device array [D]. Any output may be null.
void output_signal(const Tri &t, LegTables L, const int32_t *nobs, const double *phi, const double *lam1,
                   double *background, double *biasmat, double *decaymat, double *signal, double *binless,
                   cudaStream_t st);

// step_difference glue (fast_binless.ipp:51-105)
//   phihat = z + phi, deltahat = phihat - phi_ref(cell), weights = w
void diff_prepare(const Tri &t, LegTables L, double alpha, const int32_t *obs, const int32_t *nobs, const double *phi,
                  int ref0, double *phihat, double *deltahat, double *w, cudaStream_t st);
//   phi_ref update + difference = exp(delta)
void diff_finish(const Tri &t, int ref0, const double *phihat, const double *w, const double *delta,
                 double *phi_ref, double *difference, cudaStream_t st);

// Patch labels of one dataset: connected components of the grid edges with |dvalue| < tol, numbered
// in order of their lowest cell (graph_helpers.cpp:12-59). parent: ncells int32 scratch, scan: ncells+1024 int32.
void patch_labels(int n, double tol, const double *value, int32_t *parent, int32_t *scan, int32_t *labels, cudaStream_t st);
void fill_i32(int32_t *p, int64_t n, int32_t v, cudaStream_t st);

// BIC pieces of one (lambda2, dataset) point of the sweep: out2 = {chisq, dof} (device doubles).
// flags: ncells int32 scratch, partial: SWEEP_BLOCKS doubles scratch.
constexpr int SWEEP_BLOCKS = 1024;
void sweep_score(int64_t ncells, const double *phihat, const double *w, const double *beta, const int32_t *labels,
                 double lam1, double tol, int32_t *flags, double *partial, double *out2, cudaStream_t st);

} // namespace fb
