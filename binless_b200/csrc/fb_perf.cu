// fb_perf.cu -- the optimized path's performance iteration on the fused-lasso kernels of this library (SURVEY.md 8f
// rank 2): fb_wgfl_signal_bic / fb_wgfl_diff_bic replace, on the GPU,
//   wgfl_signal_BIC             perf_iteration_signal.cpp:22-157
//   wgfl_diff_BIC               perf_iteration_diff.cpp:20-117
//   cts_to_mat<Signal / Diff>   cts_to_mat.cpp:14-124  +  cts_to_signal_mat_core  cts_core.c:5-47
//   IRLSEstimator::optimize     IRLSEstimator.hpp:28-60,  WeightsUpdater.hpp:17-92,  compute_phi_ref util.cpp:20-41
//   EstimatedSparsity / FixedSparsity, candidates, bounds, BIC:  EstimatedSparsity.ipp:11-30, FixedSparsity.hpp:35-45,
//       candidate_generation.cpp:8-44, minimum_UB.cpp:6-40, BoundsOptimizer.ipp:5-52, BoundsRectifier.hpp,
//       ScorePreparator.ipp:53-64, Likelihoods.hpp:25-95, DOFComputer.hpp:16-24, ScoreAssembler.hpp:41-55
// (paths relative to /root/reference/binless/src).
//
// Device side: the scatter-add front end as a gather (the rows of cts are sorted by cell once per call, stably, so a
// thread adds the rows of its cell in the order the reference's loop meets them: no atomics, same sums), the fused
// lasso (GflSolver, warm across the IRLS iterations), max |beta - beta_old|, the patch labels, and the chi square of
// ALL lambda1 / eCprime candidates in one pass over the cells per tile of candidates. Host side (O(ncells log ncells)
// once per call): sorting the patch values into candidate borders, the bounds of each candidate, the degrees of freedom.
#include "fb_gfl.cuh"
#include "fb_irls.cuh"
#include "fb_perf.cuh"
#include "../../include/fastbinless.h"
#include <algorithm>
#include <cmath>
#include <cstring>
#include <numeric>
#include <vector>

namespace fb {

namespace {

constexpr int PT = 256;            // threads per block of the per-cell kernels
constexpr int CAND_TILE = 8;       // candidates evaluated per pass over a cell's data
constexpr int CHI_BLOCKS = 296;    // 2 x 148 SMs: fixed shape => same bits on every run

// one cts table on the device: rows sorted by cell (stable), CSR offsets per cell
struct CtsDev {
    int64_t N = 0;
    DevBuf<int64_t> off;
    DevBuf<double> count, lmu, nobs, ldec;
};

__device__ __forceinline__ double soft_thr(double b, double eC, double l1) {   // soft_threshold, util.cpp:8-18
    const double v = b - eC;
    return v > 0 ? fmax(0., v - l1) : fmin(0., v + l1);
}

// cts_to_signal_mat_core (cts_core.c:5-47) + remove_outliers (cts_to_mat.cpp:14-32) + the weights of cts_to_mat.cpp:81-83.
// grid (ceil(n / PT), n): thread -> cell (i, i + off). phi = phi1 (+ phi2). betahat = phihat (- sub). nobs_out = ncounts (+ nobs_add).
__global__ void __launch_bounds__(PT)
k_cts_cell(int n, const int64_t *__restrict__ off, const double *__restrict__ count, const double *__restrict__ lmu,
           const double *__restrict__ nobs, const double *__restrict__ ldec, double dispersion,
           const double *__restrict__ phi1, const double *__restrict__ phi2, const double *__restrict__ sub,
           const double *__restrict__ nobs_add, const unsigned char *__restrict__ bad_row, const unsigned char *__restrict__ bad_diag,
           double *__restrict__ betahat, double *__restrict__ phihat_raw, double *__restrict__ weight,
           double *__restrict__ weight_nodecay, double *__restrict__ nobs_out) {
    const int i = blockIdx.y, o = blockIdx.x * PT + threadIdx.x;
    if (o >= n - i) return;
    const int j = i + o;
    const int64_t pos = rm_start(i, n) + o;
    const double ph = phi2 ? phi1[pos] + phi2[pos] : phi1[pos];
    double ncounts = 0.0, var = 0.0, var_nd = 0.0, phihat = 0.0;
    for (int64_t r = off[pos]; r < off[pos + 1]; ++r) {
        const double mu = exp(lmu[r] + ph);
        const double mu_nd = exp(lmu[r] - ldec[r] + ph);
        const double z = count[r] / mu - 1;
        const double v = 1 / mu + 1 / dispersion;
        const double v_nd = 1 / mu_nd + 1 / dispersion;
        const double wt = nobs[r] / v;
        const double wt_nd = nobs[r] / v_nd;
        ncounts += nobs[r];
        var += wt;
        var_nd += wt_nd;
        phihat += (z + ph) * wt;
    }
    var = 1 / var;
    var_nd = 1 / var_nd;
    phihat = ncounts > 0 ? phihat * var : 0.0;
    if (bad_row[i] || (j != i && bad_row[j]) || bad_diag[j - i]) var = INFINITY;
    if (phihat_raw) phihat_raw[pos] = phihat;
    betahat[pos] = sub ? phihat - sub[pos] : phihat;
    weight[pos] = 1 / var;
    weight_nodecay[pos] = 1 / var_nd;
    nobs_out[pos] = nobs_add ? ncounts + nobs_add[pos] : ncounts;
}

// compute_phi_ref (util.cpp:20-41), in place on phi_ref: phihat = phi_ref + deltahat (BinnedData.hpp:76)
__global__ void __launch_bounds__(PT)
k_phi_ref(int64_t nc, const double *__restrict__ deltahat, const double *__restrict__ weight, const double *__restrict__ phihat_ref,
          const double *__restrict__ weight_ref, const double *__restrict__ delta, double lam1, int soft,
          const double *__restrict__ phi_ref_in, double *__restrict__ phi_ref_out) {
    const int64_t i = (int64_t)blockIdx.x * PT + threadIdx.x;
    if (i >= nc) return;
    const double phihat = phi_ref_in[i] + deltahat[i];
    const double var = 1 / weight[i], var_ref = 1 / weight_ref[i];
    const double d = soft ? soft_thr(delta[i], 0., lam1) : delta[i];
    double val;
    if (var_ref == INFINITY && var == INFINITY) val = (phihat_ref[i] + phihat) / 2;
    else val = (phihat_ref[i] / var_ref + (phihat - d) / var) / (1 / var_ref + 1 / var);
    phi_ref_out[i] = fmax(val, 0.);
}

// max |a - b| (IRLSEstimator::get_precision, IRLSEstimator.hpp:74-81): exact in any order
__global__ void __launch_bounds__(PT)
k_maxabs_diff(int64_t nc, const double *__restrict__ a, const double *__restrict__ b, unsigned long long *__restrict__ key) {
    double m = 0.0;
    for (int64_t i = (int64_t)blockIdx.x * PT + threadIdx.x; i < nc; i += (int64_t)gridDim.x * PT) m = fmax(m, fabs(a[i] - b[i]));
    m = warp_max(m);
    if ((threadIdx.x & 31) == 0) atomicMax(key, (unsigned long long)__double_as_longlong(m));   // m >= 0: the bit pattern is monotone
}

__global__ void __launch_bounds__(PT)
k_abs(int64_t nc, const double *__restrict__ a, double *__restrict__ out) {
    const int64_t i = (int64_t)blockIdx.x * PT + threadIdx.x;
    if (i < nc) out[i] = fabs(a[i]);
}

// chi square of a tile of candidates (LB, UB) over one block's share of the cells: Likelihood<Signal> (Likelihoods.hpp:25-58)
// or Likelihood<Difference> (:61-95, with compute_phi_ref at the thresholded delta); partial[cand * CHI_BLOCKS + block]
template <bool DIFF>
__global__ void __launch_bounds__(PT)
k_chisq(int64_t nc, int ncand, const double *__restrict__ LB, const double *__restrict__ UB, const double *__restrict__ beta_cold,
        const double *__restrict__ betahat, const double *__restrict__ wnd, const double *__restrict__ phi_ref,
        const double *__restrict__ weight, const double *__restrict__ phihat_ref, const double *__restrict__ weight_ref,
        const double *__restrict__ wnd_ref, double *__restrict__ partial) {
    __shared__ double s_red[32];
    const int c0 = blockIdx.y * CAND_TILE;
    double l1[CAND_TILE], eC[CAND_TILE], acc[CAND_TILE];
#pragma unroll
    for (int k = 0; k < CAND_TILE; ++k) {
        const int c = min(c0 + k, ncand - 1);
        l1[k] = (UB[c] - LB[c]) / 2; eC[k] = (UB[c] + LB[c]) / 2.; acc[k] = 0.0;
    }
    const int64_t per = (nc + gridDim.x - 1) / gridDim.x;
    const int64_t lo = per * blockIdx.x, hi = lo + per < nc ? lo + per : nc;
    for (int64_t i = lo + threadIdx.x; i < hi; i += PT) {
        const double b = beta_cold[i], y = betahat[i], w = wnd[i];
        double pr = 0, wt = 0, yr = 0, wr = 0, wndr = 0;
        if (DIFF) { pr = phi_ref[i]; wt = weight[i]; yr = phihat_ref[i]; wr = weight_ref[i]; wndr = wnd_ref[i]; }
#pragma unroll
        for (int k = 0; k < CAND_TILE; ++k) {
            const double s = soft_thr(b, eC[k], l1[k]);
            const double logpri = l1[k] * l1[k];
            if (!DIFF) {
                const double d = y - (s + eC[k]);
                acc[k] += w * (d * d) + logpri;
            } else {
                const double phihat = pr + y;
                const double var = 1 / wt, var_ref = 1 / wr;
                double val;
                if (var_ref == INFINITY && var == INFINITY) val = (yr + phihat) / 2;
                else val = (yr / var_ref + (phihat - s) / var) / (1 / var_ref + 1 / var);
                val = fmax(val, 0.);
                const double d = phihat - (s + eC[k] + val), dr = yr - val;
                acc[k] += w * (d * d) + wndr * (dr * dr) + logpri;
            }
        }
    }
#pragma unroll
    for (int k = 0; k < CAND_TILE; ++k) {
        double v = warp_sum(acc[k]);
        __syncthreads();
        if ((threadIdx.x & 31) == 0) s_red[threadIdx.x >> 5] = v;
        __syncthreads();
        if (threadIdx.x < 32) {
            v = threadIdx.x < PT / 32 ? s_red[threadIdx.x] : 0.0;
            v = warp_sum(v);
            if (threadIdx.x == 0 && c0 + k < ncand) partial[(int64_t)(c0 + k) * gridDim.x + blockIdx.x] = v;
        }
    }
}
__global__ void __launch_bounds__(PT)
k_chisq_fin(int nblocks, const double *__restrict__ partial, double *__restrict__ out) {
    __shared__ double s_red[32];
    const int c = blockIdx.x;
    double v = 0.0;
    for (int k = threadIdx.x; k < nblocks; k += PT) v += partial[(int64_t)c * nblocks + k];
    v = warp_sum(v);
    if ((threadIdx.x & 31) == 0) s_red[threadIdx.x >> 5] = v;
    __syncthreads();
    if (threadIdx.x < 32) {
        v = threadIdx.x < PT / 32 ? s_red[threadIdx.x] : 0.0;
        v = warp_sum(v);
        if (threadIdx.x == 0) out[c] = v;
    }
}

// ---------------------------------------------------------------------------------------------------------------
struct Stream {
    cudaStream_t st = nullptr;
    Stream() { FB_CUDA(cudaStreamCreate(&st)); }
    ~Stream() { if (st) { cudaStreamSynchronize(st); cudaStreamDestroy(st); } }
};

inline int64_t cell_of(int b1, int b2, int n) { return (int64_t)(b1 - 1) * (n + 1) - (int64_t)(b1 - 1) * b1 / 2 + b2 - b1; }

// rows sorted by cell, stably (counting sort): the gather adds them in the order of the reference's loop
void upload_cts(CtsDev &d, const fb_cts *c, int n, cudaStream_t st) {
    if (!c || c->N < 0 || (c->N > 0 && (!c->bin1 || !c->bin2 || !c->count || !c->lmu_nosig || !c->nobs || !c->log_decay)))
        throw ArgError("missing cts column");
    const int64_t nc = (int64_t)n * (n + 1) / 2, N = c->N;
    std::vector<int64_t> off(nc + 1, 0);
    for (int64_t r = 0; r < N; ++r) {
        const int b1 = c->bin1[r], b2 = c->bin2[r];
        if (b1 < 1 || b2 < b1 || b2 > n) throw ArgError("cts: bins must satisfy 1 <= bin1 <= bin2 <= nbins");
        off[cell_of(b1, b2, n) + 1] += 1;
    }
    for (int64_t k = 0; k < nc; ++k) off[k + 1] += off[k];
    std::vector<int64_t> cur(off.begin(), off.end() - 1);
    std::vector<double> h((size_t)4 * std::max<int64_t>(1, N));
    double *hc = h.data(), *hl = hc + N, *hn = hl + N, *hd = hn + N;
    for (int64_t r = 0; r < N; ++r) {
        const int64_t q = cur[cell_of(c->bin1[r], c->bin2[r], n)]++;
        hc[q] = c->count[r]; hl[q] = c->lmu_nosig[r]; hn[q] = c->nobs[r]; hd[q] = c->log_decay[r];
    }
    d.N = N;
    d.off.alloc(nc + 1);
    const size_t na = (size_t)std::max<int64_t>(1, N);
    d.count.alloc(na); d.lmu.alloc(na); d.nobs.alloc(na); d.ldec.alloc(na);
    FB_CUDA(cudaMemcpyAsync(d.off.p, off.data(), (nc + 1) * sizeof(int64_t), cudaMemcpyHostToDevice, st));
    if (N) {
        FB_CUDA(cudaMemcpyAsync(d.count.p, hc, N * sizeof(double), cudaMemcpyHostToDevice, st));
        FB_CUDA(cudaMemcpyAsync(d.lmu.p, hl, N * sizeof(double), cudaMemcpyHostToDevice, st));
        FB_CUDA(cudaMemcpyAsync(d.nobs.p, hn, N * sizeof(double), cudaMemcpyHostToDevice, st));
        FB_CUDA(cudaMemcpyAsync(d.ldec.p, hd, N * sizeof(double), cudaMemcpyHostToDevice, st));
    }
    FB_CUDA(cudaStreamSynchronize(st));   // the host vectors go out of scope
}

struct Score { double eCprime = 0, lambda1 = 0, BIC = 0, BIC_sd = -1, LB = 0, UB = 0; int dof = 0; };

// everything wgfl_signal_BIC and wgfl_diff_BIC share
struct Perf {
    int n; int64_t nc; bool diff;
    const fb_perf_params *p;
    Stream stream; cudaStream_t st;
    GflSolver flo;
    CtsDev cts, ref;
    DevBuf<unsigned char> bad_row, bad_diag;
    std::vector<int> diag_grp;
    // BinnedData (BinnedData.hpp): beta, betahat (phihat | deltahat), weight, weight_nodecay, nobs (+ Difference members)
    DevBuf<double> beta, beta_new, betahat, weight, wnd, nobs;
    DevBuf<double> phi_ref, phihat_ref, weight_ref, wnd_ref, nobs_ref, tmp;
    DevBuf<int32_t> labels, parent, scan;
    DevBuf<unsigned long long> key;
    unsigned long long *key_h = nullptr;
    int64_t admm = 0, launches = 0;
    int counter = 0, ncandidates = 0;
    double converge;

    Perf(int nbins, bool difference, const fb_perf_params *pp, const fb_cts *c, const fb_cts *r, const fb_perf_meta *m)
        : n(nbins), nc((int64_t)nbins * (nbins + 1) / 2), diff(difference), p(pp), st(stream.st), flo(nbins, 0, stream.st) {
        if (!m || !m->diag_grp) throw ArgError("metadata$diag.grp is required");
        converge = p->tol_val / 20.;
        upload_cts(cts, c, n, st);
        if (diff) upload_cts(ref, r, n, st);
        std::vector<unsigned char> br(n, 0), bd(n, 0);
        for (int k = 0; k < m->n_bad_rows; ++k) if (m->bad_rows[k] >= 1 && m->bad_rows[k] <= n) br[m->bad_rows[k] - 1] = 1;
        for (int k = 0; k < m->n_bad_diags; ++k) if (m->bad_diags[k] >= 0 && m->bad_diags[k] < n) bd[m->bad_diags[k]] = 1;
        bad_row.alloc(n); bad_diag.alloc(n);
        FB_CUDA(cudaMemcpy(bad_row.p, br.data(), n, cudaMemcpyHostToDevice));
        FB_CUDA(cudaMemcpy(bad_diag.p, bd.data(), n, cudaMemcpyHostToDevice));
        diag_grp.assign(m->diag_grp, m->diag_grp + n);
        beta.alloc(nc); beta_new.alloc(nc); betahat.alloc(nc); weight.alloc(nc); wnd.alloc(nc); nobs.alloc(nc);
        if (diff) { phi_ref.alloc(nc); phihat_ref.alloc(nc); weight_ref.alloc(nc); wnd_ref.alloc(nc); nobs_ref.alloc(nc); }
        tmp.alloc(nc);
        labels.alloc(nc); parent.alloc(nc); scan.alloc(nc + nc / 4096 + 8);
        key.alloc(1);
        key_h = static_cast<unsigned long long *>(BlockCache::host().get(sizeof(unsigned long long)));
    }
    ~Perf() { cudaStreamSynchronize(st); if (key_h) BlockCache::host().put(key_h, sizeof(unsigned long long)); }

    dim3 cell_grid() const { return dim3((unsigned)((n + PT - 1) / PT), (unsigned)n); }
    unsigned lin_grid() const { return (unsigned)((nc + PT - 1) / PT); }

    // cts_to_mat (cts_to_mat.cpp:34-124) at the current beta (and phi_ref)
    void cts_to_mat() {
        if (!diff) {
            k_cts_cell<<<cell_grid(), PT, 0, st>>>(n, cts.off.p, cts.count.p, cts.lmu.p, cts.nobs.p, cts.ldec.p, p->dispersion, beta.p,
                                                   nullptr, nullptr, nullptr, bad_row.p, bad_diag.p, betahat.p, nullptr, weight.p, wnd.p, nobs.p);
            launches += 1;
        } else {
            k_cts_cell<<<cell_grid(), PT, 0, st>>>(n, ref.off.p, ref.count.p, ref.lmu.p, ref.nobs.p, ref.ldec.p, p->dispersion, phi_ref.p,
                                                   nullptr, nullptr, nullptr, bad_row.p, bad_diag.p, phihat_ref.p, nullptr, weight_ref.p,
                                                   wnd_ref.p, nobs_ref.p);
            k_cts_cell<<<cell_grid(), PT, 0, st>>>(n, cts.off.p, cts.count.p, cts.lmu.p, cts.nobs.p, cts.ldec.p, p->dispersion, phi_ref.p,
                                                   beta.p, phi_ref.p, nobs_ref.p, bad_row.p, bad_diag.p, betahat.p, nullptr, weight.p, wnd.p, nobs.p);
            launches += 2;
        }
        FB_LAUNCH_CHECK();
    }
    // WeightsUpdater<Difference>::set_beta_delta (WeightsUpdater.hpp:80-85): phi_ref from the current binned data and beta
    void update_phi_ref() {
        k_phi_ref<<<lin_grid(), PT, 0, st>>>(nc, betahat.p, weight.p, phihat_ref.p, weight_ref.p, beta.p, 0.0, 0, phi_ref.p, phi_ref.p);
        launches += 1;
    }
    // IRLSEstimator::optimize (IRLSEstimator.hpp:28-60): beta (device) in and out
    bool irls(unsigned nouter) {
        counter = 0;
        bool has_converged = false;
        // beta_old = beta: tmp holds it
        FB_CUDA(cudaMemcpyAsync(tmp.p, beta.p, nc * sizeof(double), cudaMemcpyDeviceToDevice, st));
        do {
            if (diff) update_phi_ref();
            cts_to_mat();
            const int steps = flo.optimize(betahat.p, weight.p, p->lam2, converge);
            admm += steps - 1;
            flo.get_beta_full(beta.p);                                   // flo.get(): clamped at +-50
            FB_CUDA(cudaMemsetAsync(key.p, 0, sizeof(unsigned long long), st));
            k_maxabs_diff<<<CHI_BLOCKS, PT, 0, st>>>(nc, beta.p, tmp.p, key.p);
            launches += 1;
            FB_CUDA(cudaMemcpyAsync(key_h, key.p, sizeof(unsigned long long), cudaMemcpyDeviceToHost, st));
            FB_CUDA(cudaMemcpyAsync(tmp.p, beta.p, nc * sizeof(double), cudaMemcpyDeviceToDevice, st));
            FB_CUDA(cudaStreamSynchronize(st));
            double precision;
            std::memcpy(&precision, key_h, sizeof precision);
            has_converged = precision < converge;
            ++counter;
        } while ((unsigned)counter < nouter && !has_converged);
        if (diff) update_phi_ref();                                       // wt_.set_beta(beta)
        return has_converged;
    }
    void cold_restart() {
        beta.zero(st);
        const double a = flo.alpha;                                       // reset() keeps alpha (GFLLibraryTriangleImpl.cpp:52-58)
        flo.reset();
        flo.alpha = a;
    }
    void import_state(int warm, const fb_gfl_state *g) {                  // set_state (GFLLibraryTriangleImpl.cpp:78-87)
        if (warm && g && g->z && g->u && g->beta) { flo.alpha = g->alpha; flo.import_state(g->beta, g->z, g->u); }
    }
    void export_state(fb_gfl_state *g) {
        if (!g) return;
        g->alpha = flo.alpha; g->steps = flo.last_steps;
        flo.export_state(g->beta, g->z, g->u);
    }

    // ---- lambda1 / eCprime by BIC ------------------------------------------------------------------------------
    struct Host { std::vector<double> beta, weight, nobs; std::vector<int32_t> patchno; } h;
    std::vector<double> sorted_vals;      // ZeroOffset: |beta| ascending; EstimatedOffset: beta ascending
    std::vector<int64_t> pos_weight_prefix;   // EstimatedOffset: cells with weight > 0 among the first k sorted
    std::vector<double> pmin, pmax;       // per patch of the warm solution: smallest / largest beta
    double lsnc = 0;
    DevBuf<double> beta_cold;
    bool zero_offset = true, positive = false, forbid = false;

    void prepare_sparsity() {
        // patches of the warm solution (perf_iteration_signal.cpp:60-63)
        patch_labels(n, p->tol_val, beta.p, parent.p, scan.p, labels.p, st);
        launches += 6;
        h.beta.resize(nc); h.weight.resize(nc); h.nobs.resize(nc); h.patchno.resize(nc);
        FB_CUDA(cudaMemcpyAsync(h.beta.data(), beta.p, nc * sizeof(double), cudaMemcpyDeviceToHost, st));
        FB_CUDA(cudaMemcpyAsync(h.weight.data(), weight.p, nc * sizeof(double), cudaMemcpyDeviceToHost, st));
        FB_CUDA(cudaMemcpyAsync(h.nobs.data(), nobs.p, nc * sizeof(double), cudaMemcpyDeviceToHost, st));
        FB_CUDA(cudaMemcpyAsync(h.patchno.data(), labels.p, nc * sizeof(int32_t), cudaMemcpyDeviceToHost, st));
        // ScorePreparator<BIC>::compute (ScorePreparator.ipp:53-64): cold-start re-solve on the last (betahat, weight)
        const double a = flo.alpha;
        flo.reset(); flo.alpha = a;
        const int steps = flo.optimize(betahat.p, weight.p, p->lam2, converge);
        admm += steps - 1;
        beta_cold.alloc(nc);
        flo.get_beta_full(beta_cold.p);
        FB_CUDA(cudaStreamSynchronize(st));
        double s = 0; for (double v : h.nobs) s += v;
        lsnc = std::log(s);
        int np = 0; for (int32_t q : h.patchno) np = std::max(np, q + 1);
        pmin.assign(np, INFINITY); pmax.assign(np, -INFINITY);
        for (int64_t i = 0; i < nc; ++i) { pmin[h.patchno[i]] = std::min(pmin[h.patchno[i]], h.beta[i]); pmax[h.patchno[i]] = std::max(pmax[h.patchno[i]], h.beta[i]); }
    }
    double value_of(double b) const { return zero_offset ? std::fabs(b) : b; }     // CandidatesValues.hpp
    double minimum_UB() const {                                                  // minimum_UB.hpp:25-33, minimum_UB.cpp:6-40
        const double minbeta = *std::min_element(h.beta.begin(), h.beta.end());
        double minUBo;
        if (zero_offset) minUBo = positive ? std::max(-minbeta, 0.) : 0.;
        else minUBo = minbeta;
        double vmin = INFINITY, vmax = -INFINITY;
        for (double b : h.beta) { const double v = value_of(b); vmin = std::min(vmin, v); vmax = std::max(vmax, v); }
        double minUBd;
        if (!forbid) minUBd = vmin - 1;
        else {
            const int ndiags = *std::max_element(diag_grp.begin(), diag_grp.end()) + 1;
            std::vector<char> touch(ndiags, 0);
            std::vector<double> dv(ndiags, vmax);
            int64_t i = 0;
            for (int r = 0; r < n; ++r)
                for (int c = r; c < n; ++c, ++i) {
                    const int g = diag_grp[c - r];
                    dv[g] = std::min(dv[g], value_of(h.beta[i])); touch[g] = 1;
                }
            double m = -INFINITY;
            for (int g = 0; g < ndiags; ++g) if (touch[g]) m = std::max(m, dv[g]);
            minUBd = m;
        }
        return std::max(minUBd, minUBo);
    }
    std::vector<double> candidates(double minUB) {                                // candidate_generation.hpp:27-40
        // patches of the candidate values (a second labelling: of |beta| when the offset is zero)
        if (zero_offset) { k_abs<<<lin_grid(), PT, 0, st>>>(nc, beta.p, tmp.p); launches += 1; }
        patch_labels(n, p->tol_val, zero_offset ? tmp.p : beta.p, parent.p, scan.p, labels.p, st);
        launches += 6;
        std::vector<int32_t> pv(nc);
        FB_CUDA(cudaMemcpyAsync(pv.data(), labels.p, nc * sizeof(int32_t), cudaMemcpyDeviceToHost, st));
        FB_CUDA(cudaStreamSynchronize(st));
        int np = 0; for (int32_t q : pv) np = std::max(np, q + 1);
        std::vector<double> borders(np);
        for (int64_t i = 0; i < nc; ++i) borders[pv[i]] = value_of(h.beta[i]);   // the last cell of a patch wins (:12)
        std::sort(borders.begin(), borders.end());
        std::vector<double> ret; ret.push_back(minUB);                            // cleanup_borders (:23-35)
        bool any = false; double last = 0;
        for (double b : borders) if (b >= minUB) { ret.push_back(b); any = true; last = b; }
        ret.push_back(any ? last + 1 : minUB + 1);
        std::vector<double> cand;
        for (size_t i = 0; i + 1 < ret.size(); ++i) cand.push_back((ret[i] + ret[i + 1]) / 2);
        return cand;
    }
    void prepare_bounds() {
        sorted_vals.resize(nc);
        if (zero_offset) {
            for (int64_t i = 0; i < nc; ++i) sorted_vals[i] = std::fabs(h.beta[i]);
            std::sort(sorted_vals.begin(), sorted_vals.end());
        } else {
            std::vector<int64_t> idx(nc);
            std::iota(idx.begin(), idx.end(), 0);
            std::sort(idx.begin(), idx.end(), [&](int64_t a, int64_t b) { return h.beta[a] < h.beta[b]; });
            pos_weight_prefix.assign(nc + 1, 0);
            for (int64_t k = 0; k < nc; ++k) {
                sorted_vals[k] = h.beta[idx[k]];
                pos_weight_prefix[k + 1] = pos_weight_prefix[k] + (h.weight[idx[k]] != 0 ? 1 : 0);
            }
        }
    }
    std::pair<double, double> optimize_bounds(double val) const {                 // BoundsOptimizer.ipp:5-52
        if (!zero_offset) {
            const double xmin = sorted_vals.front(), xmax = sorted_vals.back();
            const int64_t n1 = std::upper_bound(sorted_vals.begin(), sorted_vals.end(), val) - sorted_vals.begin();
            // weights are sums of non-negative terms: a group's sum is zero iff none of its weights is positive
            const bool grp1empty = n1 == 0 || pos_weight_prefix[n1] == 0;
            const bool grp2empty = n1 == nc || pos_weight_prefix[nc] - pos_weight_prefix[n1] == 0;
            double UB, LB;
            if (!grp2empty && !grp1empty) { UB = sorted_vals[n1 - 1]; LB = xmin; }
            else if (grp2empty) { UB = xmax; LB = xmin; }
            else { UB = xmin; LB = xmin; }
            if (UB < LB) throw NumError("Found LB > UB!");
            return {LB, UB};
        }
        const double mn = sorted_vals.front(), mx = sorted_vals.back();
        double xk;
        if (val > mx) xk = mx;
        else if (val < mn) xk = 0;
        else xk = *(std::upper_bound(sorted_vals.begin(), sorted_vals.end(), val) - 1);
        return {-xk, xk};
    }
    int dof_of(double LB, double UB) const {                                      // DOFComputer.hpp:16-24
        const double lambda1 = (UB - LB) / 2, eCprime = (UB + LB) / 2.;
        auto soft = [&](double b) { const double v = b - eCprime; return v > 0 ? std::max(0., v - lambda1) : std::min(0., v + lambda1); };
        int dof = 0;
        // soft() is monotone in beta: a patch holds a cell beyond tol/2 iff one of its two extreme values is
        for (size_t q = 0; q < pmin.size(); ++q)
            if (std::fabs(soft(pmax[q])) > p->tol_val / 2 || std::fabs(soft(pmin[q])) > p->tol_val / 2) ++dof;
        return dof;
    }
    // ScoreComputer::evaluate (ScoreComputer.ipp:24-41) for a list of bounds: chi squares on the device, dof on the host
    std::vector<Score> evaluate(const std::vector<std::pair<double, double>> &bounds) {
        const int nb = (int)bounds.size();
        std::vector<double> lb(nb), ub(nb), chi(nb);
        for (int k = 0; k < nb; ++k) { lb[k] = bounds[k].first; ub[k] = bounds[k].second; }
        DevBuf<double> dlb(nb), dub(nb), partial((size_t)nb * CHI_BLOCKS), dchi(nb);
        FB_CUDA(cudaMemcpyAsync(dlb.p, lb.data(), nb * sizeof(double), cudaMemcpyHostToDevice, st));
        FB_CUDA(cudaMemcpyAsync(dub.p, ub.data(), nb * sizeof(double), cudaMemcpyHostToDevice, st));
        const dim3 grid(CHI_BLOCKS, (unsigned)((nb + CAND_TILE - 1) / CAND_TILE));
        if (diff) k_chisq<true><<<grid, PT, 0, st>>>(nc, nb, dlb.p, dub.p, beta_cold.p, betahat.p, wnd.p, phi_ref.p, weight.p, phihat_ref.p,
                                                     weight_ref.p, wnd_ref.p, partial.p);
        else k_chisq<false><<<grid, PT, 0, st>>>(nc, nb, dlb.p, dub.p, beta_cold.p, betahat.p, wnd.p, nullptr, nullptr, nullptr, nullptr,
                                                 nullptr, partial.p);
        k_chisq_fin<<<nb, PT, 0, st>>>(CHI_BLOCKS, partial.p, dchi.p);
        launches += 2;
        FB_LAUNCH_CHECK();
        FB_CUDA(cudaMemcpyAsync(chi.data(), dchi.p, nb * sizeof(double), cudaMemcpyDeviceToHost, st));
        FB_CUDA(cudaStreamSynchronize(st));
        std::vector<Score> out(nb);
        for (int k = 0; k < nb; ++k) {
            Score &s = out[k];
            s.LB = lb[k]; s.UB = ub[k]; s.lambda1 = (ub[k] - lb[k]) / 2; s.eCprime = (ub[k] + lb[k]) / 2.;
            s.dof = dof_of(lb[k], ub[k]);
            s.BIC = chi[k] + lsnc * s.dof; s.BIC_sd = -1;
        }
        return out;
    }
    Score choose() {
        prepare_sparsity();
        const double minUB = minimum_UB();
        if (p->lambda1_fixed) {                                                   // FixedSparsity.hpp:35-45, BoundsRectifier.hpp
            ncandidates = 1;
            std::pair<double, double> b;
            if (!zero_offset) {
                const double LB = *std::min_element(h.beta.begin(), h.beta.end());
                double UB = LB + 2 * p->lambda1_fix_value;
                if (minUB > UB) UB = minUB;
                b = {LB, UB};
            } else {
                double UB = p->lambda1_fix_value;
                if (minUB > p->lambda1_fix_value) UB = minUB;
                b = {-UB, UB};
            }
            return evaluate({b})[0];
        }
        const std::vector<double> cand = candidates(minUB);                       // EstimatedSparsity.ipp:11-30
        ncandidates = (int)cand.size();
        prepare_bounds();
        std::vector<std::pair<double, double>> bounds;
        for (double c : cand) {
            const auto b = optimize_bounds(c);
            const double lambda1 = (b.second - b.first) / 2.;
            if (lambda1 < p->tol_val && c < cand.back() - 1e-12) continue;
            bounds.push_back(b);
        }
        if (bounds.empty()) throw NumError("Found 0 values!");
        const std::vector<Score> values = evaluate(bounds);
        size_t best = 0;                                                          // ScoreOptimizer<BIC> (k = 0): the first minimum
        for (size_t k = 1; k < values.size(); ++k) if (values[k].BIC < values[best].BIC) best = k;
        return values[best];
    }
    void download(double *h_dst, const double *d_src) { if (h_dst) FB_CUDA(cudaMemcpyAsync(h_dst, d_src, nc * sizeof(double), cudaMemcpyDeviceToHost, st)); }
};

void require_device() {
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) throw CudaError("no CUDA device available (there is no CPU fallback)");
}

} // namespace

void wgfl_signal_bic(int nbins, const fb_cts *cts, const fb_perf_meta *meta, const fb_perf_params *p, const double *beta_init,
                     int warm, fb_gfl_state *gfl, fb_perf_signal_out *out) {
    if (!p || !out || !beta_init || nbins < 2) throw ArgError("null argument");
    require_device();
    Perf P(nbins, false, p, cts, nullptr, meta);
    P.import_state(warm, gfl);
    FB_CUDA(cudaMemcpyAsync(P.beta.p, beta_init, P.nc * sizeof(double), cudaMemcpyHostToDevice, P.st));
    unsigned nwarm, ncold;
    if (p->nouter == 1) nwarm = 1; else nwarm = p->nouter / 2;
    ncold = p->nouter - nwarm;
    bool converged = P.irls(nwarm);
    if (ncold > 0 && !converged) { P.cold_restart(); converged = P.irls(ncold); }
    P.export_state(gfl);
    P.zero_offset = p->fixed != 0; P.positive = p->fixed ? p->positive != 0 : true; P.forbid = p->constrained != 0;
    const Score opt = P.choose();
    const int64_t nc = P.nc;
    if (out->phi) for (int64_t i = 0; i < nc; ++i) {
        const double v = P.h.beta[i] - opt.eCprime;
        out->phi[i] = v > 0 ? std::max(0., v - opt.lambda1) : std::min(0., v + opt.lambda1);
    }
    if (out->beta) std::memcpy(out->beta, P.h.beta.data(), nc * sizeof(double));
    if (out->weight) std::memcpy(out->weight, P.h.weight.data(), nc * sizeof(double));
    if (out->nobs) std::memcpy(out->nobs, P.h.nobs.data(), nc * sizeof(double));
    if (out->patchno) std::memcpy(out->patchno, P.h.patchno.data(), nc * sizeof(int32_t));
    P.download(out->phihat, P.betahat.p);
    FB_CUDA(cudaStreamSynchronize(P.st));
    out->dof = opt.dof; out->converged = converged; out->nouter_done = P.counter; out->ncandidates = P.ncandidates;
    out->BIC = opt.BIC; out->BIC_sd = opt.BIC_sd; out->eCprime = opt.eCprime; out->lambda1 = opt.lambda1;
    out->admm_iters = P.admm;
}

void wgfl_diff_bic(int nbins, const fb_cts *cts, const fb_cts *ref, const fb_perf_meta *meta, const fb_perf_params *p,
                   const double *phi_ref_init, const double *beta_init, int warm, fb_gfl_state *gfl, fb_perf_diff_out *out) {
    if (!p || !out || !beta_init || !phi_ref_init || nbins < 2) throw ArgError("null argument");
    require_device();
    Perf P(nbins, true, p, cts, ref, meta);
    P.import_state(warm, gfl);
    const int64_t nc = P.nc;
    // WeightsUpdater<Difference>::setUp (WeightsUpdater.hpp:57-62)
    FB_CUDA(cudaMemcpyAsync(P.beta.p, beta_init, nc * sizeof(double), cudaMemcpyHostToDevice, P.st));
    FB_CUDA(cudaMemcpyAsync(P.phi_ref.p, phi_ref_init, nc * sizeof(double), cudaMemcpyHostToDevice, P.st));
    P.cts_to_mat();
    bool converged = P.irls((unsigned)p->nouter);
    if (!converged) { P.cold_restart(); converged = P.irls((unsigned)p->nouter); }
    P.export_state(gfl);
    P.zero_offset = true; P.positive = false; P.forbid = p->constrained != 0;
    const Score opt = P.choose();
    if (out->delta) for (int64_t i = 0; i < nc; ++i) {
        const double v = P.h.beta[i];
        out->delta[i] = v > 0 ? std::max(0., v - opt.lambda1) : std::min(0., v + opt.lambda1);
    }
    if (out->beta) std::memcpy(out->beta, P.h.beta.data(), nc * sizeof(double));
    if (out->weight) std::memcpy(out->weight, P.h.weight.data(), nc * sizeof(double));
    if (out->nobs) std::memcpy(out->nobs, P.h.nobs.data(), nc * sizeof(double));
    if (out->patchno) std::memcpy(out->patchno, P.h.patchno.data(), nc * sizeof(int32_t));
    if (out->phi_ref) {   // compute_phi_ref(binned, soft(beta, 0, lambda1)) (perf_iteration_diff.cpp:85-91)
        k_phi_ref<<<P.lin_grid(), PT, 0, P.st>>>(nc, P.betahat.p, P.weight.p, P.phihat_ref.p, P.weight_ref.p, P.beta.p, opt.lambda1, 1,
                                                 P.phi_ref.p, P.tmp.p);
        P.download(out->phi_ref, P.tmp.p);
    }
    P.download(out->deltahat, P.betahat.p);
    P.download(out->phihat_ref, P.phihat_ref.p);
    FB_CUDA(cudaStreamSynchronize(P.st));
    if (out->phihat || out->phihat_var || out->phihat_var_ref) {
        std::vector<double> pr(nc), dh(nc), wr(nc);
        FB_CUDA(cudaMemcpy(pr.data(), P.phi_ref.p, nc * sizeof(double), cudaMemcpyDeviceToHost));
        FB_CUDA(cudaMemcpy(dh.data(), P.betahat.p, nc * sizeof(double), cudaMemcpyDeviceToHost));
        FB_CUDA(cudaMemcpy(wr.data(), P.weight_ref.p, nc * sizeof(double), cudaMemcpyDeviceToHost));
        for (int64_t i = 0; i < nc; ++i) {
            if (out->phihat) out->phihat[i] = pr[i] + dh[i];
            if (out->phihat_var) out->phihat_var[i] = 1 / P.h.weight[i];
            if (out->phihat_var_ref) out->phihat_var_ref[i] = 1 / wr[i];
        }
    }
    out->dof = opt.dof; out->converged = converged; out->nouter_done = P.counter; out->ncandidates = P.ncandidates;
    out->BIC = opt.BIC; out->BIC_sd = opt.BIC_sd; out->lambda1 = opt.lambda1;
    out->admm_iters = P.admm;
}

} // namespace fb
