// fb_gfl.cu -- ADMM kernels of the graph-fused-lasso solve (sm_100a).
//
// One ADMM iteration (gfl_graph_fl.c:143-183) is four launches on one stream:
//   k_beta    closed-form beta update over 32x32 tiles of the triangle, reading the
//             row duals in row-major and the column duals in column-major order
//             (shared-memory tile transpose), writing beta and the two chain
//             inputs yr = beta+u_row (row-major), yc = beta+u_col (column-major)
//   k_chains  the 1-D fused-lasso prox of every row / column chain: one LANE per
//             chain, 32 chains per warp, exact taut-string walk (same unique
//             minimiser as tf_dp_weight, gfl_tf.c:184-321); emits one
//             (position,value) marker per constant segment
//   k_expand  warp-cooperative, fully coalesced: expands the markers to z, applies
//             u += beta - z (gfl_graph_fl.c:306-327) and accumulates the primal /
//             dual residual sums (gfl_graph_fl.c:329-365) as per-block partials
//   k_ctl     single block: fixed-order reduction of the partials, convergence
//             test, varying-penalty rule every 50 steps (gfl_graph_fl.c:163-181)
// The host polls the control block every few iterations; once `done` is set the
// remaining queued launches return immediately.
#include "fb_gfl.cuh"
#include <cmath>
#include <vector>

namespace fb {

namespace {

constexpr int TILE = 32;
constexpr int EXPAND_WARPS = 4;
constexpr int CTL_THREADS = 1024;
constexpr int POLL_BATCH = 8;

// --------------------------------------------------------------------------
// beta update (update_beta_weight, gfl_graph_fl.c:241-264)
// --------------------------------------------------------------------------
__global__ void __launch_bounds__(TILE * 8)
k_beta(Grid g, const GflCtl *__restrict__ ctl, const double *__restrict__ y, const double *__restrict__ w,
       const double *__restrict__ zr, double *__restrict__ ur, const double *__restrict__ zc,
       double *__restrict__ uc, double *__restrict__ beta, double *__restrict__ yr, double *__restrict__ yc) {
    if (ctl->done) return;
    const int i0 = blockIdx.y * TILE, j0 = blockIdx.x * TILE;
    // tile must intersect the band 0 <= j-i < md
    if (j0 + TILE - 1 < i0 || j0 - (i0 + TILE - 1) >= g.md) return;
    __shared__ double s_tcol[TILE][TILE + 1]; // z_col - u_col, indexed [j][i]
    __shared__ double s_ucol[TILE][TILE + 1];
    __shared__ double s_beta[TILE][TILE + 1]; // indexed [i][j]
    const int tx = threadIdx.x, ty = threadIdx.y;
    const double alpha = ctl->alpha, inflate = ctl->inflate;
    const int mode = ctl->rescale;

    // phase A: column duals, coalesced along i in column-major storage
    for (int jj = ty; jj < TILE; jj += 8) {
        const int j = j0 + jj, i = i0 + tx;
        double t = 0.0, u = 0.0;
        if (j < g.n && i <= j && j - i < g.md && j >= 1) {
            const int64_t idx = g.cm(i, j);
            u = uc[idx];
            if (mode == 1) { u = u / inflate; uc[idx] = u; }
            else if (mode == 2) { u = u * inflate; uc[idx] = u; }
            t = zc[idx] - u;
        }
        s_tcol[jj][tx] = t;
        s_ucol[jj][tx] = u;
    }
    __syncthreads();
    // phase B: row duals + beta, coalesced along j in row-major storage
    for (int ii = ty; ii < TILE; ii += 8) {
        const int i = i0 + ii, j = j0 + tx;
        double b = 0.0;
        if (j < g.n && i <= j && j - i < g.md) {
            const int64_t idx = g.rm(i, j);
            const bool has_row = (i <= g.n - 2), has_col = (j >= 1);
            double r = 0.0, u = 0.0;
            if (has_row) {
                u = ur[idx];
                if (mode == 1) { u = u / inflate; ur[idx] = u; }
                else if (mode == 2) { u = u * inflate; ur[idx] = u; }
                r += zr[idx] - u;
            }
            if (has_col) r += s_tcol[tx][ii];
            const int nz = (has_row ? 1 : 0) + (has_col ? 1 : 0);
            const double wi = w[idx];
            b = (2.0 * y[idx] * wi + alpha * r) / (2.0 * wi + alpha * nz);
            beta[idx] = b;
            if (has_row) yr[idx] = b + u;
        }
        s_beta[ii][tx] = b;
    }
    __syncthreads();
    // phase C: column chain inputs, coalesced along i
    for (int jj = ty; jj < TILE; jj += 8) {
        const int j = j0 + jj, i = i0 + tx;
        if (j < g.n && i <= j && j - i < g.md && j >= 1) yc[g.cm(i, j)] = s_beta[tx][jj] + s_ucol[jj][tx];
    }
}

// --------------------------------------------------------------------------
// chain prox: lane-per-chain exact taut string. For the chain y_0..y_{L-1} and
// tube radius lam' = lam/(alpha/2) it finds the unique minimiser of
//   1/2 sum (x_k - y_k)^2 + lam' sum |x_{k+1} - x_k|
// i.e. exactly what tf_dp_weight(n, y, w = alpha/2, lam) returns
// (update_z, gfl_graph_fl.c:267-304). Slopes are kept as (numerator, length)
// pairs and compared by cross-multiplication, so the walk has no division.
// --------------------------------------------------------------------------
__global__ void __launch_bounds__(32)
k_chains(Grid g, const GflCtl *__restrict__ ctl, const double *__restrict__ yr, const double *__restrict__ yc,
         double *__restrict__ mr, double *__restrict__ mc, uint32_t *__restrict__ fr, uint32_t *__restrict__ fc) {
    if (ctl->done) return;
    const int q = blockIdx.x * 32 + threadIdx.x;
    bool active = q < g.nchains();
    int dir = 0, idx = 0, L = 0;
    int64_t start = 0;
    if (active) g.chain(q, dir, idx, L, start);
    const double *__restrict__ Y = (dir == 0 ? yr : yc) + start;
    double *__restrict__ M = (dir == 0 ? mr : mc) + start;
    uint32_t *__restrict__ F = (dir == 0 ? fr : fc);
    const double lam = ctl->lam / (ctl->alpha / 2.0);
    const double INF = __longlong_as_double(0x7ff0000000000000LL);

    if (lam == 0.0) { // tf_dp_weight copies y when lam == 0 (gfl_tf.c:210-215)
        if (active)
            for (int p = 0; p < L; ++p) { M[p] = Y[p]; atomicOr(&F[(start + p) >> 5], 1u << ((start + p) & 31)); }
        return;
    }

    int k0 = 0, j = 0, kmin = 0, kmax = 0;
    double T = 0.0, c = 0.0, a = 0.0, lenf = 0.0;
    double nmin = -INF, dmin = 1.0, nmax = INF, dmax = 1.0;
    double ycur = active ? Y[0] : 0.0;
    while (__any_sync(0xffffffffu, active)) {
        if (active) {
            const bool last = (j + 1 == L);
            const double ynext = last ? 0.0 : Y[j + 1]; // speculative: used unless a segment breaks
            const double yv = ycur;
            if (j == k0) c = yv;
            T += yv - c;
            lenf += 1.0;
            const double lj = last ? 0.0 : lam;
            const double base = T - a;
            const double nlo = base - lj, nhi = base + lj;
            const bool up = nlo * dmax > nmax * lenf;   // lower tube edge above the max-slope ray
            const bool dn = nhi * dmin < nmin * lenf;   // upper tube edge below the min-slope ray
            if (up || dn) {
                const int kend = up ? kmax : kmin;
                const double slope = up ? nmax / dmax : nmin / dmin;
                M[kend - 1] = c + slope;
                atomicOr(&F[(start + kend - 1) >> 5], 1u << ((start + kend - 1) & 31));
                a = up ? lam : -lam;
                k0 = kend; j = kend; kmin = kend; kmax = kend;
                T = 0.0; lenf = 0.0;
                nmin = -INF; dmin = 1.0; nmax = INF; dmax = 1.0;
                ycur = Y[j];
            } else {
                if (nlo * dmin >= nmin * lenf) { nmin = nlo; dmin = lenf; kmin = j + 1; }
                if (nhi * dmax <= nmax * lenf) { nmax = nhi; dmax = lenf; kmax = j + 1; }
                if (last) {
                    M[L - 1] = c + nlo / lenf;
                    atomicOr(&F[(start + L - 1) >> 5], 1u << ((start + L - 1) & 31));
                    active = false;
                } else { ++j; ycur = ynext; }
            }
        }
    }
}

// --------------------------------------------------------------------------
// expand markers -> z, u update, residual partial sums. Block = EXPAND_WARPS
// warps, each handling one 32-bit flag word (32 consecutive chain slots) of
// chain blockIdx.y. Also clears the other parity's flag words.
// --------------------------------------------------------------------------
__global__ void __launch_bounds__(32 * EXPAND_WARPS)
k_expand(Grid g, const GflCtl *__restrict__ ctl, const double *__restrict__ yr, const double *__restrict__ yc,
         const double *__restrict__ mr, const double *__restrict__ mc, const uint32_t *__restrict__ fr,
         const uint32_t *__restrict__ fc, uint32_t *__restrict__ fr_next, uint32_t *__restrict__ fc_next,
         double *__restrict__ zr, double *__restrict__ ur, double *__restrict__ zc, double *__restrict__ uc,
         double *__restrict__ partials) {
    if (ctl->done) return;
    __shared__ double s_p[EXPAND_WARPS], s_d[EXPAND_WARPS];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int q = blockIdx.y;
    int dir, idx, L; int64_t start;
    g.chain(q, dir, idx, L, start);
    const int64_t end = start + L;
    const int64_t w_first = start >> 5, w_last = (end - 1) >> 5;
    const int64_t wd = w_first + (int64_t)blockIdx.x * EXPAND_WARPS + wid;
    double accp = 0.0, accd = 0.0;
    if (wd <= w_last) {
        const double *__restrict__ Y = dir == 0 ? yr : yc;
        const double *__restrict__ M = dir == 0 ? mr : mc;
        const uint32_t *__restrict__ F = dir == 0 ? fr : fc;
        double *__restrict__ Z = dir == 0 ? zr : zc;
        double *__restrict__ U = dir == 0 ? ur : uc;
        const double alpha = ctl->alpha;
        const int64_t flat = wd * 32 + lane;
        const bool valid = flat >= start && flat < end;
        const uint32_t own = __ballot_sync(0xffffffffu, valid);
        const uint32_t bits = F[wd] & own;
        const bool marked = valid && ((bits >> lane) & 1u);
        const double mval = marked ? M[flat] : 0.0;
        // value of the first marker beyond this word (needed by slots after the last marker here)
        double carry = 0.0;
        const uint32_t top = bits ? (31 - __clz(bits)) : 0;
        const bool need_carry = (own >> 1 >> top) != 0 || bits == 0; // some valid lane above the highest marker
        if (need_carry) {
            int64_t ws = wd + 1;
            for (;;) { // scan following words of this chain, 32 at a time
                const int64_t wq = ws + lane;
                uint32_t b = 0;
                if (wq <= w_last) {
                    b = F[wq];
                    if (wq == w_last) { const int hi = (int)((end - 1) & 31); b &= (hi == 31) ? 0xffffffffu : ((2u << hi) - 1u); }
                }
                const uint32_t nz = __ballot_sync(0xffffffffu, b != 0);
                if (nz) {
                    const int src = __ffs(nz) - 1;
                    const uint32_t bw = __shfl_sync(0xffffffffu, b, src);
                    carry = M[(ws + src) * 32 + (__ffs(bw) - 1)];
                    break;
                }
                ws += 32; // the chain's last slot always carries a marker, so this terminates
            }
        }
        const uint32_t higher = bits >> lane;
        const int src = higher ? (lane + __ffs(higher) - 1) : lane;
        double z = __shfl_sync(0xffffffffu, mval, src);
        if (!higher) z = carry;
        if (valid) {
            const double yv = Y[flat], uo = U[flat], zo = Z[flat];
            const double un = yv - z;         // u + (beta - z) with beta + u = yv
            const double rp = (yv - uo) - z;  // beta - z
            accp = rp * rp;
            const double dd = alpha * (z - zo);
            accd = dd * dd;
            Z[flat] = z;
            U[flat] = un;
        }
        if (lane == 0) { // clear the next iteration's flag word (own bits only; words are shared at chain ends)
            uint32_t *Fn = dir == 0 ? fr_next : fc_next;
            if (own == 0xffffffffu) Fn[wd] = 0u; else atomicAnd(&Fn[wd], ~own);
        }
    }
    accp = warp_sum(accp); accd = warp_sum(accd);
    if (lane == 0) { s_p[wid] = accp; s_d[wid] = accd; }
    __syncthreads();
    if (threadIdx.x == 0) {
        double p = 0.0, d = 0.0;
        for (int k = 0; k < EXPAND_WARPS; ++k) { p += s_p[k]; d += s_d[k]; }
        const int64_t pid = (int64_t)blockIdx.y * gridDim.x + blockIdx.x;
        partials[2 * pid] = p;
        partials[2 * pid + 1] = d;
    }
}

// --------------------------------------------------------------------------
// residual norms, convergence, varying penalty (gfl_graph_fl.c:159-181)
// --------------------------------------------------------------------------
__global__ void __launch_bounds__(CTL_THREADS)
k_ctl(GflCtl *ctl, const double *__restrict__ partials, int npart, int64_t ncells, int64_t nz) {
    if (ctl->done) return;
    __shared__ double s_p[CTL_THREADS], s_d[CTL_THREADS];
    double p = 0.0, d = 0.0;
    for (int k = threadIdx.x; k < npart; k += CTL_THREADS) { p += partials[2 * k]; d += partials[2 * k + 1]; }
    s_p[threadIdx.x] = p; s_d[threadIdx.x] = d;
    __syncthreads();
    for (int o = CTL_THREADS / 2; o > 0; o >>= 1) {
        if ((int)threadIdx.x < o) { s_p[threadIdx.x] += s_p[threadIdx.x + o]; s_d[threadIdx.x] += s_d[threadIdx.x + o]; }
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        const double rp = sqrt(s_p[0] / (double)ncells);
        const double rd = sqrt(s_d[0] / (double)nz);
        const double cur = rp > rd ? rp : rd;
        ctl->presnorm = rp; ctl->dresnorm = rd;
        int step = ctl->step;
        int rescale = 0;
        if (step % 50 == 0 && rp > 10 * rd) { ctl->alpha *= ctl->inflate; rescale = 1; }
        else if (step % 50 == 0 && rd > 10 * rp) { ctl->alpha /= ctl->inflate; rescale = 2; }
        ctl->rescale = rescale;
        ++step;
        ctl->step = step;
        ctl->done = !(step < ctl->maxsteps && cur > ctl->converge);
    }
}

// applies a pending dual rescale after the last iteration of a solve
__global__ void k_rescale_u(double *__restrict__ u, int64_t n, int mode, double inflate) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) u[i] = (mode == 1) ? u[i] / inflate : u[i] * inflate;
}

// band gather / scatter (extract_trapeze, fill_triangle: GFLLibraryTrapezeImpl.cpp:93-131)
__global__ void k_band_gather(Grid g, int nbins, const double *__restrict__ tri, double *__restrict__ band) {
    const int i = blockIdx.y;
    const int len = g.row_len(i);
    const int64_t src = rm_start(i, nbins), dst = g.rm_row_start(i);
    for (int d = blockIdx.x * blockDim.x + threadIdx.x; d < len; d += gridDim.x * blockDim.x) band[dst + d] = tri[src + d];
}
// sum of w*y and w over the cells outside the band, per row partials
__global__ void k_band_outside_sums(Grid g, int nbins, const double *__restrict__ y, const double *__restrict__ w,
                                    double *__restrict__ partial) {
    const int i = blockIdx.x;
    __shared__ double s_a[256], s_b[256];
    double a = 0.0, b = 0.0;
    const int64_t src = rm_start(i, nbins);
    for (int d = g.md + threadIdx.x; i + d < nbins; d += blockDim.x) { a += w[src + d] * y[src + d]; b += w[src + d]; }
    s_a[threadIdx.x] = a; s_b[threadIdx.x] = b;
    __syncthreads();
    for (int o = 128; o > 0; o >>= 1) {
        if ((int)threadIdx.x < o) { s_a[threadIdx.x] += s_a[threadIdx.x + o]; s_b[threadIdx.x] += s_b[threadIdx.x + o]; }
        __syncthreads();
    }
    if (threadIdx.x == 0) { partial[2 * i] = s_a[0]; partial[2 * i + 1] = s_b[0]; }
}
__global__ void k_band_avg(int nrows, const double *__restrict__ partial, double *__restrict__ avg) {
    if (threadIdx.x == 0 && blockIdx.x == 0) {
        double a = 0.0, b = 0.0;
        for (int i = 0; i < nrows; ++i) { a += partial[2 * i]; b += partial[2 * i + 1]; }
        avg[0] = a / b;
    }
}
// beta (clamped at +-50) scattered to the full triangle; outside the band = avg
__global__ void k_beta_full(Grid g, int nbins, const double *__restrict__ beta, const double *__restrict__ avg,
                            double clampv, double *__restrict__ out) {
    const int i = blockIdx.y;
    const int len = g.row_len(i);
    const int64_t dst = rm_start(i, nbins), src = g.rm_row_start(i);
    for (int d = blockIdx.x * blockDim.x + threadIdx.x; i + d < nbins; d += gridDim.x * blockDim.x) {
        double v = (d < len) ? beta[src + d] : avg[0];
        v = fmin(clampv, fmax(-clampv, v));
        out[dst + d] = v;
    }
}

} // namespace

// ==========================================================================
GflSolver::GflSolver(int nbins, unsigned maxdiag, cudaStream_t stream) : nbins_(nbins), st_(stream) {
    unsigned md = maxdiag < (unsigned)nbins ? maxdiag : (unsigned)nbins; // FusedLassoGaussianEstimator.hpp:33
    if (md > (unsigned)(nbins + 1) / 2 || md <= 1) { trapeze_ = false; md = nbins; } // GFLLibrary.hpp:21-27
    else trapeze_ = true;
    if (nbins < 2) throw ArgError("fused lasso needs at least 2 bins");
    g_ = Grid::make(nbins, (int)md);
    const int64_t N = g_.ncells;
    beta_.alloc(N); yr_.alloc(N); yc_.alloc(N); zr_.alloc(N); ur_.alloc(N); zc_.alloc(N); uc_.alloc(N);
    mr_.alloc(N); mc_.alloc(N);
    nwords_ = (N + 31) / 32 + 1;
    flags_.alloc(4 * nwords_);
    if (trapeze_) { yband_.alloc(N); wband_.alloc(N); fillavg_.alloc(1); }
    const int maxwords = (g_.row_len(0) + 31) / 32 + 1;
    const int gx = (maxwords + EXPAND_WARPS - 1) / EXPAND_WARPS;
    n_partials_ = gx * g_.nchains();
    partials_.alloc(2 * (size_t)std::max(n_partials_, nbins));
    ctl_.alloc(1);
    FB_CUDA(cudaMallocHost((void **)&h_ctl_, sizeof(GflCtl)));
    FB_CUDA(cudaEventCreate(&ev0_)); FB_CUDA(cudaEventCreate(&ev1_));
    reset();
}

GflSolver::~GflSolver() {
    if (h_ctl_) cudaFreeHost(h_ctl_);
    cudaEventDestroy(ev0_); cudaEventDestroy(ev1_);
}

void GflSolver::reset() {
    counter = 0;
    alpha = 5.0;
    beta_.zero(st_); zr_.zero(st_); ur_.zero(st_); zc_.zero(st_); uc_.zero(st_);
    yr_.zero(st_); yc_.zero(st_); mr_.zero(st_); mc_.zero(st_);
}

void GflSolver::iteration(const double *y, const double *w, int parity) {
    uint32_t *f_cur = flags_.p + (size_t)parity * 2 * nwords_;
    uint32_t *f_nxt = flags_.p + (size_t)(parity ^ 1) * 2 * nwords_;
    dim3 tb(TILE, 8), tg((g_.n + TILE - 1) / TILE, (g_.n + TILE - 1) / TILE);
    k_beta<<<tg, tb, 0, st_>>>(g_, ctl_.p, y, w, zr_.p, ur_.p, zc_.p, uc_.p, beta_.p, yr_.p, yc_.p);
    k_chains<<<(g_.nchains() + 31) / 32, 32, 0, st_>>>(g_, ctl_.p, yr_.p, yc_.p, mr_.p, mc_.p, f_cur, f_cur + nwords_);
    const int maxwords = (g_.row_len(0) + 31) / 32 + 1;
    dim3 eg((maxwords + EXPAND_WARPS - 1) / EXPAND_WARPS, g_.nchains());
    k_expand<<<eg, 32 * EXPAND_WARPS, 0, st_>>>(g_, ctl_.p, yr_.p, yc_.p, mr_.p, mc_.p, f_cur, f_cur + nwords_,
                                               f_nxt, f_nxt + nwords_, zr_.p, ur_.p, zc_.p, uc_.p, partials_.p);
    k_ctl<<<1, CTL_THREADS, 0, st_>>>(ctl_.p, partials_.p, n_partials_, g_.ncells, 2 * g_.ncells - 2);
    kernel_launches += 4;
}

int GflSolver::optimize(const double *d_y_tri, const double *d_w_tri, double lam2, double converge) {
    FB_CUDA(cudaEventRecord(ev0_, st_));
    const double *y = d_y_tri, *w = d_w_tri;
    if (trapeze_) {
        dim3 gg(4, nbins_);
        k_band_gather<<<gg, 256, 0, st_>>>(g_, nbins_, d_y_tri, yband_.p);
        k_band_gather<<<gg, 256, 0, st_>>>(g_, nbins_, d_w_tri, wband_.p);
        y = yband_.p; w = wband_.p;
        kernel_launches += 2;
    }
    y_tri_ = d_y_tri; w_tri_ = d_w_tri;
    GflCtl c;
    c.alpha = alpha; c.lam = lam2; c.converge = converge; c.inflate = inflate;
    c.presnorm = c.dresnorm = 0; c.step = 1; c.maxsteps = ninner_max;
    c.done = !(c.step < c.maxsteps && (converge + 1) > converge);
    c.rescale = 0;
    *h_ctl_ = c;
    FB_CUDA(cudaMemcpyAsync(ctl_.p, h_ctl_, sizeof(GflCtl), cudaMemcpyHostToDevice, st_));
    flags_.zero(st_);
    int parity = 0;
    bool done = c.done != 0;
    while (!done) {
        for (int b = 0; b < POLL_BATCH; ++b) { iteration(y, w, parity); parity ^= 1; }
        FB_LAUNCH_CHECK();
        FB_CUDA(cudaMemcpyAsync(h_ctl_, ctl_.p, sizeof(GflCtl), cudaMemcpyDeviceToHost, st_));
        FB_CUDA(cudaStreamSynchronize(st_));
        done = h_ctl_->done != 0;
    }
    if (!c.done) {
        if (h_ctl_->rescale) { // rescale decided on the very last step still applies to u
            const int64_t N = g_.ncells;
            k_rescale_u<<<(unsigned)((N + 255) / 256), 256, 0, st_>>>(ur_.p, N, h_ctl_->rescale, inflate);
            k_rescale_u<<<(unsigned)((N + 255) / 256), 256, 0, st_>>>(uc_.p, N, h_ctl_->rescale, inflate);
            kernel_launches += 2;
        }
        alpha = h_ctl_->alpha;
        last_steps = h_ctl_->step;
    } else last_steps = 1;
    counter += last_steps;
    if (trapeze_) { // average of the far region (fill_triangle)
        k_band_outside_sums<<<nbins_, 256, 0, st_>>>(g_, nbins_, d_y_tri, d_w_tri, partials_.p);
        k_band_avg<<<1, 32, 0, st_>>>(nbins_, partials_.p, fillavg_.p);
        kernel_launches += 2;
    }
    FB_CUDA(cudaEventRecord(ev1_, st_));
    FB_CUDA(cudaEventSynchronize(ev1_));
    FB_CUDA(cudaEventElapsedTime(&ms_last, ev0_, ev1_));
    return last_steps;
}

void GflSolver::get_beta_full(double *d_out_tri) {
    dim3 gg(4, nbins_);
    k_beta_full<<<gg, 256, 0, st_>>>(g_, nbins_, beta_.p, trapeze_ ? fillavg_.p : beta_.p, 50.0, d_out_tri);
    kernel_launches += 1;
    FB_LAUNCH_CHECK();
}

// reference slot order: row chains (rm index p < N-1), then column chains
// (cm index >= 1 maps to slot N-1 + cm-1)
void GflSolver::export_state(double *h_beta, double *h_z, double *h_u) {
    const int64_t N = g_.ncells;
    FB_CUDA(cudaStreamSynchronize(st_));
    if (h_beta) FB_CUDA(cudaMemcpy(h_beta, beta_.p, N * sizeof(double), cudaMemcpyDeviceToHost));
    if (h_z) {
        FB_CUDA(cudaMemcpy(h_z, zr_.p, (N - 1) * sizeof(double), cudaMemcpyDeviceToHost));
        FB_CUDA(cudaMemcpy(h_z + (N - 1), zc_.p + 1, (N - 1) * sizeof(double), cudaMemcpyDeviceToHost));
    }
    if (h_u) {
        FB_CUDA(cudaMemcpy(h_u, ur_.p, (N - 1) * sizeof(double), cudaMemcpyDeviceToHost));
        FB_CUDA(cudaMemcpy(h_u + (N - 1), uc_.p + 1, (N - 1) * sizeof(double), cudaMemcpyDeviceToHost));
    }
}
void GflSolver::import_state(const double *h_beta, const double *h_z, const double *h_u) {
    const int64_t N = g_.ncells;
    FB_CUDA(cudaStreamSynchronize(st_));
    if (h_beta) FB_CUDA(cudaMemcpy(beta_.p, h_beta, N * sizeof(double), cudaMemcpyHostToDevice));
    if (h_z) {
        FB_CUDA(cudaMemcpy(zr_.p, h_z, (N - 1) * sizeof(double), cudaMemcpyHostToDevice));
        FB_CUDA(cudaMemcpy(zc_.p + 1, h_z + (N - 1), (N - 1) * sizeof(double), cudaMemcpyHostToDevice));
    }
    if (h_u) {
        FB_CUDA(cudaMemcpy(ur_.p, h_u, (N - 1) * sizeof(double), cudaMemcpyHostToDevice));
        FB_CUDA(cudaMemcpy(uc_.p + 1, h_u + (N - 1), (N - 1) * sizeof(double), cudaMemcpyHostToDevice));
    }
}

} // namespace fb
