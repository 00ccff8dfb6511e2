// fb_irls.cu -- IRLS / glue kernels of the fast-binless outer loop (sm_100a). See fb_irls.cuh.
#include "fb_irls.cuh"
#include <cmath>

namespace fb {

namespace {

constexpr int EW_THREADS = 256, EW_PER_THREAD = 4, EW_SPAN = EW_THREADS * EW_PER_THREAD;

// get_log_expected for one cell (fast_expected.ipp:1-19): ((bias + decay) + signal) + exposure, with the
// bias term the sum of the two selector entries (fast_bias_mean.hpp:50-52). bias_decay(): the two table terms, from the
// closed-form tables (bin_data layout) or, in the per-cell fallback (L.cb1 set), through the int32 bin indices of the cell.
__device__ __forceinline__ void bias_decay(const LegTables &L, int i, int j, int64_t idx, double &b, double &dc) {
    if (L.cb1) {
        const int p = L.cb1[idx], q = L.cb2[idx];
        b = (p == q) ? 2. * L.brow[p] : (p < q ? L.brow[p] + L.brow[q] : L.brow[q] + L.brow[p]);
        dc = L.ddiag[L.cdb[idx]];
    } else {
        const double bi = L.brow[i];
        b = (i == j) ? 2. * bi : bi + L.brow[j];
        dc = L.ddiag[j - i];
    }
}
// the closed-form tables only (k_leg_pass: the per-cell fallback sums its legs elsewhere, leg_pass_cells)
__device__ __forceinline__ double log_expected_tab(const LegTables &L, int i, int j, int d, double phi) {
    const double bi = L.brow[i];
    const double b = (i == j) ? 2. * bi : bi + L.brow[j];
    return ((b + L.ddiag[j - i]) + phi) + L.expo[d];
}
__device__ __forceinline__ double log_expected(const LegTables &L, int i, int j, int d, int64_t idx, double phi) {
    double b, dc;
    bias_decay(L, i, j, idx, b, dc);
    return ((b + dc) + phi) + L.expo[d];
}

// NB working residual and weight (fast_residuals.ipp:47-67)
__device__ __forceinline__ void nb_residual(double eta, int obs, int nobs, double alpha, double &z, double &w) {
    if (nobs > 0) {
        const double mu = exp(eta);
        z = (double)obs / ((double)nobs * mu) - 1;
        w = (double)nobs / (1 / mu + 1 / alpha);
    } else { z = 0; w = 0; }
}

// order-preserving map double -> int64 so that min/max can use integer atomics (exact, order independent)
__device__ __forceinline__ long long dkey(double x) {
    const long long b = __double_as_longlong(x);
    return b >= 0 ? b : (b ^ 0x7fffffffffffffffLL);
}
__device__ __forceinline__ double dkey_inv(long long k) {
    return __longlong_as_double(k >= 0 ? k : (k ^ 0x7fffffffffffffffLL));
}

__device__ __forceinline__ double block_sum(double v, double *s /*[32]*/) {
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
    v = warp_sum(v);
    __syncthreads();
    if (lane == 0) s[wid] = v;
    __syncthreads();
    double r = 0.0;
    if (wid == 0) {
        r = lane < nw ? s[lane] : 0.0;
        r = warp_sum(r);
    }
    return r; // valid in warp 0
}

// --------------------------------------------------------------------------
// leg pass: 32 x 128 tiles of one dataset's triangle
// --------------------------------------------------------------------------
template <int LEG, int MODE>
__global__ void __launch_bounds__(TILE_J)
k_leg_pass(Tri t, LegTables L, double alpha, const int32_t *__restrict__ obs, const int32_t *__restrict__ nobs,
           const double *__restrict__ phi, double *__restrict__ part, const int grp_rank, const int grp_world) {
    const int tJ = blockIdx.x, tI = blockIdx.y, d = blockIdx.z;
    const int ntJ = gridDim.x, ntI = gridDim.y;
    const int i0 = tI * TILE_I, j0 = tJ * TILE_J;
    if (j0 + TILE_J - 1 < i0) return; // tile entirely below the diagonal
    if (((tI % LEG_GROUPS) % grp_world) != grp_rank) return;   // a tile row of a group another GPU sums (IRLS legs sharded)
    const int tx = threadIdx.x, lane = tx & 31, wid = tx >> 5;
    const int j = j0 + tx;
    const int n = t.n;
    __shared__ double s_row[TILE_J / 32][TILE_I][2];
    __shared__ double s_diag[TILE_J / 32][TILE_DIAGS][2];
    __shared__ double s_red[2][32];
    if (LEG == LEG_DECAY) {
        for (int k = tx; k < (TILE_J / 32) * TILE_DIAGS * 2; k += TILE_J) (&s_diag[0][0][0])[k] = 0.0;
        __syncthreads();
    }
    double ca = 0.0, cb = 0.0;
    const int64_t base = (int64_t)d * t.ncells;
#pragma unroll 1
    for (int g = 0; g < TILE_I; g += 4) {
        int ob[4], no[4];
        double ph[4];
        bool ok[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const int i = i0 + g + q;
            ok[q] = (i < n) && (j < n) && (j >= i);
            ob[q] = 0; no[q] = 0; ph[q] = 0.0;
            if (ok[q]) {
                const int64_t idx = base + rm_start(i, n) + (j - i);
                no[q] = nobs[idx];
                if (MODE != MODE_NOBS) { ob[q] = obs[idx]; ph[q] = phi[idx]; }
            }
        }
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const int ii = g + q, i = i0 + ii;
            double a = 0.0, b = 0.0;
            if (ok[q]) {
                if (MODE == MODE_NOBS) { a = (double)no[q]; }
                else {
                    const double eta = log_expected_tab(L, i, j, d, ph[q]);
                    if (MODE == MODE_RESID) {
                        double z, w;
                        nb_residual(eta, ob[q], no[q], alpha, z, w);
                        a = w; b = z * w;
                    } else { a = (double)ob[q]; b = exp(eta) * (double)no[q]; }
                }
            }
            if (LEG == LEG_EXPO) { ca += a; cb += b; }
            if (LEG == LEG_BIAS) {
                ca += a; cb += b;
                const double ra = warp_sum(a), rb = warp_sum(b);
                if (lane == 0) { s_row[wid][ii][0] = ra; s_row[wid][ii][1] = rb; }
            }
            if (LEG == LEG_DECAY) {
                const int slot = tx + (TILE_I - 1) - ii; // (j - i) - (j0 - i0 - 31)
                s_diag[wid][slot][0] += a;
                s_diag[wid][slot][1] += b;
                __syncwarp();
            }
        }
    }
    if (LEG == LEG_EXPO) {
        const double ta = block_sum(ca, s_red[0]), tb = block_sum(cb, s_red[1]);
        if (tx == 0) {
            const int64_t p = ((int64_t)(d * ntI + tI) * ntJ + tJ) * 2;
            part[p] = ta; part[p + 1] = tb;
        }
    }
    if (LEG == LEG_BIAS) {
        __syncthreads();
        if (tx < TILE_I && i0 + tx < n) {
            double ra = 0.0, rb = 0.0;
#pragma unroll
            for (int w = 0; w < TILE_J / 32; ++w) { ra += s_row[w][tx][0]; rb += s_row[w][tx][1]; }
            const int64_t p = ((int64_t)(d * ntJ + tJ) * n + i0 + tx) * 2;
            part[p] = ra; part[p + 1] = rb;
        }
        if (j < n) {
            const int64_t p = (int64_t)t.D * ntJ * n * 2 + ((int64_t)(d * ntI + tI) * n + j) * 2;
            part[p] = ca; part[p + 1] = cb;
        }
    }
    if (LEG == LEG_DECAY) {
        __syncthreads();
        for (int slot = tx; slot < TILE_DIAGS; slot += TILE_J) {
            double ra = 0.0, rb = 0.0;
#pragma unroll
            for (int w = 0; w < TILE_J / 32; ++w) { ra += s_diag[w][slot][0]; rb += s_diag[w][slot][1]; }
            const int64_t p = (((int64_t)(d * ntI + tI) * ntJ + tJ) * TILE_DIAGS + slot) * 2;
            part[p] = ra; part[p + 1] = rb;
        }
    }
}

// finalize: fixed-order sums of the tile partials into per-item pairs. The sums have a two-level shape that does not
// depend on how many GPUs share the call: tile row tI belongs to group tI mod LEG_GROUPS; level 1 adds the partials of one
// group (one warp or block per (item, group): lane / thread strided, then a butterfly), level 2 adds the LEG_GROUPS group
// sums in order. One GPU computes every group itself; `world` GPUs (world divides LEG_GROUPS) compute the groups
// g mod world == rank each, all-gather the group sums and form the same level-2 sum: bit-identical results.
__device__ __forceinline__ bool group_is_mine(int g, int grp_rank, int grp_world) { return (g % grp_world) == grp_rank; }
// where the sums of group g live: the groups of one rank side by side, so that ONE all-gather collects everything
__host__ __device__ __forceinline__ int group_slot(int g, int grp_world) { return (g % grp_world) * (LEG_GROUPS / grp_world) + g / grp_world; }

__global__ void __launch_bounds__(256)
k_fin_expo(Tri t, int ntI, int ntJ, const double *__restrict__ part, double *__restrict__ items_g, int grp_rank, int grp_world) {
    __shared__ double s_red[2][32];
    const int d = blockIdx.x, g = blockIdx.y;
    if (!group_is_mine(g, grp_rank, grp_world)) return;
    const int nti_g = ntI > g ? (ntI - g + LEG_GROUPS - 1) / LEG_GROUPS : 0;
    double a = 0.0, b = 0.0;
    for (int k = threadIdx.x; k < nti_g * ntJ; k += blockDim.x) {
        const int tI = g + LEG_GROUPS * (k / ntJ), tJ = k % ntJ;
        if (tJ * TILE_J + TILE_J - 1 < tI * TILE_I) continue;
        const int64_t p = ((int64_t)(d * ntI + tI) * ntJ + tJ) * 2;
        a += part[p]; b += part[p + 1];
    }
    a = block_sum(a, s_red[0]); b = block_sum(b, s_red[1]);
    const int64_t o = ((int64_t)group_slot(g, grp_world) * t.D + d) * 2;
    if (threadIdx.x == 0) { items_g[o] = a; items_g[o + 1] = b; }
}

// one warp per (bin, group): the row partials (by tile column; they all come from the tile row of the bin) and the
// column partials (by tile row) of every dataset
__global__ void __launch_bounds__(128)
k_fin_bias(Tri t, int ntI, int ntJ, const double *__restrict__ part, double *__restrict__ items_g, int grp_rank, int grp_world) {
    const int r = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const int g = blockIdx.y;
    const int lane = threadIdx.x & 31;
    const int n = t.n;
    if (r >= n || !group_is_mine(g, grp_rank, grp_world)) return;
    const double2 *rp = reinterpret_cast<const double2 *>(part);
    const double2 *cp = reinterpret_cast<const double2 *>(part + (int64_t)t.D * ntJ * n * 2);
    const int tj_min = ((r / TILE_I) * TILE_I) / TILE_J;
    int ti_max = ((r / TILE_J) * TILE_J + TILE_J - 1) / TILE_I;
    if (ti_max > ntI - 1) ti_max = ntI - 1;
    const int nrow = (((r / TILE_I) % LEG_GROUPS) == g) ? ntJ - tj_min : 0;
    const int ncol = ti_max >= g ? (ti_max - g) / LEG_GROUPS + 1 : 0, per = nrow + ncol;
    double a = 0.0, b = 0.0;
    for (int q = lane; q < t.D * per; q += 32) {
        const int d = q / per, k = q - d * per;
        const double2 v = k < nrow ? __ldg(rp + ((int64_t)(d * ntJ + tj_min + k) * n + r))
                                   : __ldg(cp + ((int64_t)(d * ntI + g + LEG_GROUPS * (k - nrow)) * n + r));
        a += v.x; b += v.y;
    }
    a = warp_sum(a); b = warp_sum(b);
    const int64_t o = ((int64_t)group_slot(g, grp_world) * n + r) * 2;
    if (lane == 0) { items_g[o] = a; items_g[o + 1] = b; }
}

// one warp per (diagonal, group)
__global__ void __launch_bounds__(128)
k_fin_decay(Tri t, int ntI, int ntJ, const double *__restrict__ part, double *__restrict__ items_g, int grp_rank, int grp_world) {
    const int k = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const int g = blockIdx.y;
    const int lane = threadIdx.x & 31;
    if (k >= t.n || !group_is_mine(g, grp_rank, grp_world)) return;
    const int nti_g = ntI > g ? (ntI - g + LEG_GROUPS - 1) / LEG_GROUPS : 0;
    double a = 0.0, b = 0.0;
    for (int q = lane; q < t.D * nti_g; q += 32) {
        const int d = q / nti_g, tI = g + LEG_GROUPS * (q % nti_g);
        const int i0 = tI * TILE_I;
        const int hi = k + i0 + (TILE_I - 1);          // j0 <= hi
        const int lo = hi - (TILE_DIAGS - 2);          // j0 >= lo  (slot <= 158)
        int tj_lo = lo <= 0 ? 0 : (lo + TILE_J - 1) / TILE_J;
        int tj_hi = hi / TILE_J;
        if (tj_hi > ntJ - 1) tj_hi = ntJ - 1;
        for (int tJ = tj_lo; tJ <= tj_hi; ++tJ) {
            const int j0 = tJ * TILE_J;
            if (j0 + TILE_J - 1 < i0) continue;
            const int slot = k - (j0 - i0 - (TILE_I - 1));
            const int64_t p = (((int64_t)(d * ntI + tI) * ntJ + tJ) * TILE_DIAGS + slot) * 2;
            a += part[p]; b += part[p + 1];
        }
    }
    a = warp_sum(a); b = warp_sum(b);
    const int64_t o = ((int64_t)group_slot(g, grp_world) * t.n + k) * 2;
    if (lane == 0) { items_g[o] = a; items_g[o + 1] = b; }
}

// level 2: the group sums in order
__global__ void __launch_bounds__(256)
k_sum_groups(int nitems, const double *__restrict__ items_g, double *__restrict__ items, int grp_world) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= 2 * nitems) return;
    double v = 0.0;
#pragma unroll
    for (int g = 0; g < LEG_GROUPS; ++g) v += items_g[(int64_t)group_slot(g, grp_world) * 2 * nitems + i];
    items[i] = v;
}

// --------------------------------------------------------------------------
// elementwise kernels: grid (ceil(n / EW_SPAN), n, D), thread -> cells of row blockIdx.y
// --------------------------------------------------------------------------
#define EW_FOREACH_CELL(t)                                                                    \
    const int i = blockIdx.y, d = blockIdx.z, n = (t).n;                                      \
    const int64_t rowbase = (int64_t)d * (t).ncells + rm_start(i, n);                          \
    _Pragma("unroll") for (int q_ = 0; q_ < EW_PER_THREAD; ++q_)                              \
        for (int off = blockIdx.x * EW_SPAN + q_ * EW_THREADS + threadIdx.x, j = i + off,     \
                 once_ = 1; once_ && off < n - i; once_ = 0)

__global__ void __launch_bounds__(EW_THREADS)
k_signal_yw(Tri t, LegTables L, double alpha, const int32_t *__restrict__ obs, const int32_t *__restrict__ nobs,
            const double *__restrict__ phi, unsigned group, double *__restrict__ y, double *__restrict__ w,
            const int sh_rank, const int sh_world, const int md) {
    EW_FOREACH_CELL(t) {
        // band solves shared between GPUs: beyond the band only the far-region sums need (y, w), and their rows are dealt
        // to the ranks (k_band_outside_sums reads the same rows)
        if (off >= md && (i % sh_world) != sh_rank) continue;
        const int64_t idx = rowbase + off;
        const double ph = phi[idx];
        double z, wt;
        nb_residual(log_expected(L, i, j, d, idx, ph), obs[idx], nobs[idx], alpha, z, wt);
        // checkerboard hold-out on 1-based bins (fast_binless.ipp:14-22)
        const unsigned cvgroup = 1 + ((unsigned)((j + 1) + (i + 1)) % 2);
        if (cvgroup == group) { y[idx] = -100; w[idx] = 0; }
        else { y[idx] = z + ph; w[idx] = wt; }
    }
}

__global__ void __launch_bounds__(EW_THREADS)
k_precision(Tri t, LegTables Lo, LegTables Ln, const double *__restrict__ phi_old, const double *__restrict__ phi_new,
            long long *__restrict__ keys, const int sh_rank, const int sh_world) {
    const double INF = __longlong_as_double(0x7ff0000000000000LL);
    double dabs = -INF, mx = -INF, mn = INF;
    // a block walks rows blockIdx.y, blockIdx.y + gridDim.y, ... (a mix of long and short rows) and reduces once:
    // one block per row spent most of its time in the reduction (28 800 blocks at 2400 bins x 4)
    const int d = blockIdx.z, n = t.n;
    for (int i = blockIdx.y; i < n; i += gridDim.y) {
        if ((i % sh_world) != sh_rank) continue;      // rows of another GPU of the communicator
        const int64_t rowbase = (int64_t)d * t.ncells + rm_start(i, n);
        for (int off = threadIdx.x; off < n - i; off += EW_THREADS) {
            const int j = i + off;
            const int64_t idx = rowbase + off;
            const double eo = log_expected(Lo, i, j, d, idx, phi_old[idx]);
            const double en = log_expected(Ln, i, j, d, idx, phi_new[idx]);
            dabs = fmax(dabs, fabs(en - eo));
            mx = fmax(mx, en); mn = fmin(mn, en);
        }
    }
    __shared__ double s[3][EW_THREADS / 32];
    dabs = warp_max(dabs); mx = warp_max(mx); mn = warp_min(mn);
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    if (lane == 0) { s[0][wid] = dabs; s[1][wid] = mx; s[2][wid] = mn; }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int k = 1; k < EW_THREADS / 32; ++k) { dabs = fmax(dabs, s[0][k]); mx = fmax(mx, s[1][k]); mn = fmin(mn, s[2][k]); }
        if (mx > -INF) {
            // ~29 000 blocks update three addresses: look first (a stale value only costs an unneeded atomic)
            const long long ka = dkey(dabs), kx = dkey(mx), kn = dkey(mn);
            if (ka > __ldcg(&keys[0])) atomicMax(&keys[0], ka);
            if (kx > __ldcg(&keys[1])) atomicMax(&keys[1], kx);
            if (kn < __ldcg(&keys[2])) atomicMin(&keys[2], kn);
        }
    }
}
__global__ void k_precision_init(long long *keys) {
    const double INF = __longlong_as_double(0x7ff0000000000000LL);
    keys[0] = dkey(-INF); keys[1] = dkey(-INF); keys[2] = dkey(INF);
}
__global__ void k_precision_fin(const long long *keys, double *out3) {
    out3[0] = dkey_inv(keys[0]); out3[1] = dkey_inv(keys[1]); out3[2] = dkey_inv(keys[2]);
}

// minima per (dataset, row-or-column index) and (dataset, diagonal): tile kernel like the leg pass
__global__ void __launch_bounds__(TILE_J)
k_minima(Tri t, const double *__restrict__ phi, long long *__restrict__ rowkey, long long *__restrict__ diagkey,
         const int sh_rank, const int sh_world) {
    const int tJ = blockIdx.x, tI = blockIdx.y, d = blockIdx.z;
    const int i0 = tI * TILE_I, j0 = tJ * TILE_J;
    if (j0 + TILE_J - 1 < i0) return;
    if ((tI % sh_world) != sh_rank) return;
    const int tx = threadIdx.x, lane = tx & 31, wid = tx >> 5, n = t.n;
    const int j = j0 + tx;
    const double INF = __longlong_as_double(0x7ff0000000000000LL);
    __shared__ double s_row[TILE_J / 32][TILE_I];
    __shared__ double s_diag[TILE_J / 32][TILE_DIAGS];
    for (int k = tx; k < (TILE_J / 32) * TILE_DIAGS; k += TILE_J) (&s_diag[0][0])[k] = INF;
    __syncthreads();
    double cmin = INF;
    const int64_t base = (int64_t)d * t.ncells;
#pragma unroll 4
    for (int ii = 0; ii < TILE_I; ++ii) {
        const int i = i0 + ii;
        double v = INF;
        if (i < n && j < n && j >= i) v = phi[base + rm_start(i, n) + (j - i)];
        cmin = fmin(cmin, v);
        const double r = warp_min(v);
        if (lane == 0) s_row[wid][ii] = r;
        const int slot = tx + (TILE_I - 1) - ii;
        s_diag[wid][slot] = fmin(s_diag[wid][slot], v);
        __syncwarp();
    }
    __syncthreads();
    if (tx < TILE_I && i0 + tx < n) {
        double r = s_row[0][tx];
#pragma unroll
        for (int w = 1; w < TILE_J / 32; ++w) r = fmin(r, s_row[w][tx]);
        if (r < INF) atomicMin(&rowkey[(int64_t)d * n + i0 + tx], dkey(r));
    }
    if (j < n && cmin < INF) atomicMin(&rowkey[(int64_t)d * n + j], dkey(cmin));
    for (int slot = tx; slot < TILE_DIAGS - 1; slot += TILE_J) {
        double r = s_diag[0][slot];
#pragma unroll
        for (int w = 1; w < TILE_J / 32; ++w) r = fmin(r, s_diag[w][slot]);
        const int k = slot + (j0 - i0 - (TILE_I - 1));
        if (r < INF && k >= 0 && k < n) atomicMin(&diagkey[(int64_t)d * n + k], dkey(r));
    }
}
__global__ void k_fill_keys(long long *keys, int64_t n, double v) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) keys[i] = dkey(v);
}
// signal -= max(min_row[b1], min_row[b2], min_diag[b2-b1]); clamp at 0 (fast_binless.cpp:56-63)
__global__ void __launch_bounds__(EW_THREADS)
k_degen_apply(Tri t, double *__restrict__ phi, const long long *__restrict__ rowkey, const long long *__restrict__ diagkey) {
    EW_FOREACH_CELL(t) {
        const int64_t idx = rowbase + off;
        const double adj = fmax(fmax(dkey_inv(rowkey[(int64_t)d * n + i]), dkey_inv(rowkey[(int64_t)d * n + j])),
                                dkey_inv(diagkey[(int64_t)d * n + (j - i)]));
        phi[idx] = fmax(phi[idx] - adj, 0.);
    }
}
__global__ void __launch_bounds__(EW_THREADS)
k_dataset_min(Tri t, const double *__restrict__ phi, long long *__restrict__ key, const int sh_rank, const int sh_world) {
    const double INF = __longlong_as_double(0x7ff0000000000000LL);
    if (((int)blockIdx.y % sh_world) != sh_rank) return;
    double mn = INF;
    EW_FOREACH_CELL(t) { (void)j; mn = fmin(mn, phi[rowbase + off]); }
    mn = warp_min(mn);
    __shared__ double s[EW_THREADS / 32];
    if ((threadIdx.x & 31) == 0) s[threadIdx.x >> 5] = mn;
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int k = 1; k < EW_THREADS / 32; ++k) mn = fmin(mn, s[k]);
        if (mn < INF) atomicMin(&key[blockIdx.z], dkey(mn));
    }
}
__global__ void __launch_bounds__(EW_THREADS)
k_shift_apply(Tri t, double *__restrict__ phi, const long long *__restrict__ key) {
    EW_FOREACH_CELL(t) {
        (void)j;
        const int64_t idx = rowbase + off;
        phi[idx] = fmax(phi[idx] - dkey_inv(key[d]), 0.);
    }
}

// centring: fixed grid of partial dot products, then one block
constexpr int CENTER_BLOCKS = 1024;
__global__ void __launch_bounds__(256)
k_center_partial(int64_t ncells, const int32_t *__restrict__ nobs, const double *__restrict__ beta, double *__restrict__ scratch,
                 const int sh_rank, const int sh_world) {
    __shared__ double s_red[2][32];
    if (((int)blockIdx.x % sh_world) != sh_rank) {      // a block of another GPU: an exact zero in the all-reduce
        if (threadIdx.x == 0) { scratch[2 * blockIdx.x] = 0.0; scratch[2 * blockIdx.x + 1] = 0.0; }
        return;
    }
    double a = 0.0, b = 0.0;
    for (int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; c < ncells; c += (int64_t)gridDim.x * blockDim.x) {
        const double no = (double)nobs[c];
        a += beta[c] * no; b += no;
    }
    a = block_sum(a, s_red[0]); b = block_sum(b, s_red[1]);
    if (threadIdx.x == 0) { scratch[2 * blockIdx.x] = a; scratch[2 * blockIdx.x + 1] = b; }
}
__global__ void __launch_bounds__(CENTER_BLOCKS)
k_center_final(double *__restrict__ scratch) {
    __shared__ double s_red[2][32];
    double a = scratch[2 * threadIdx.x], b = scratch[2 * threadIdx.x + 1];
    a = block_sum(a, s_red[0]); b = block_sum(b, s_red[1]);
    if (threadIdx.x == 0) scratch[2 * CENTER_BLOCKS] = a / b;
}
__global__ void __launch_bounds__(256)
k_center_apply(int64_t ncells, double *__restrict__ beta, const double *__restrict__ scratch) {
    const double avg = scratch[2 * CENTER_BLOCKS];
    for (int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; c < ncells; c += (int64_t)gridDim.x * blockDim.x)
        beta[c] = beta[c] - avg;
}

__global__ void __launch_bounds__(EW_THREADS)
k_output_signal(Tri t, LegTables L, const int32_t *__restrict__ nobs, const double *__restrict__ phi,
                const double *__restrict__ lam1, double *__restrict__ background, double *__restrict__ biasmat,
                double *__restrict__ decaymat, double *__restrict__ signal, double *__restrict__ binless) {
    EW_FOREACH_CELL(t) {
        const int64_t idx = rowbase + off;
        double bipbj, ldec;
        bias_decay(L, i, j, idx, bipbj, ldec);
        const double l1 = lam1[d];
        double lsig = phi[idx];
        lsig = (lsig > 0) ? fmax(0., lsig - l1) : fmin(0., lsig + l1); // soft_threshold, util.cpp:8-18
        if (biasmat) biasmat[idx] = exp(bipbj);
        if (decaymat) decaymat[idx] = exp(ldec);
        if (signal) signal[idx] = exp(lsig);
        if (binless) binless[idx] = exp(ldec + lsig);
        if (background) background[idx] = (double)nobs[idx] * exp(bipbj + ldec + L.expo[d]);
    }
}

__global__ void __launch_bounds__(EW_THREADS)
k_diff_prepare(Tri t, LegTables L, double alpha, const int32_t *__restrict__ obs, const int32_t *__restrict__ nobs,
               const double *__restrict__ phi, int ref0, double *__restrict__ phihat, double *__restrict__ deltahat,
               double *__restrict__ w) {
    EW_FOREACH_CELL(t) {
        const int64_t idx = rowbase + off;
        const double pref = phi[(int64_t)ref0 * t.ncells + (idx - (int64_t)d * t.ncells)];
        // FastData<Difference>: log_difference = log_signal - phi_ref, log_signal = phi_ref + log_difference (FastData.cpp:6-31)
        const double dsig = pref + (phi[idx] - pref);
        double z, wt;
        nb_residual(log_expected(L, i, j, d, idx, dsig), obs[idx], nobs[idx], alpha, z, wt);
        const double ph = z + dsig;
        phihat[idx] = ph; deltahat[idx] = ph - pref; w[idx] = wt;
    }
}
__global__ void __launch_bounds__(EW_THREADS)
k_diff_finish(Tri t, int ref0, const double *__restrict__ phihat, const double *__restrict__ w,
              const double *__restrict__ delta, double *__restrict__ phi_ref, double *__restrict__ difference) {
    EW_FOREACH_CELL(t) {
        (void)j;
        const int64_t idx = rowbase + off;
        const int64_t ridx = (int64_t)ref0 * t.ncells + (idx - (int64_t)d * t.ncells);
        const double pr = phihat[ridx], wr = w[ridx], ph = phihat[idx], wt = w[idx], dl = delta[idx];
        double val;
        if (wr == 0 && wt == 0) val = (pr + ph) / 2;
        else val = (pr * wr + (ph - dl) * wt) / (wr + wt);
        if (phi_ref) phi_ref[idx] = fmax(val, 0.);
        if (difference) difference[idx] = exp(dl);
    }
}

// --------------------------------------------------------------------------
// patch labels: lock-free union-find (larger root hooks onto smaller root, so the
// final root of a component is its lowest cell), then rank of the root among roots
// --------------------------------------------------------------------------
__device__ __forceinline__ int uf_find(int32_t *parent, int a) {
    int p = parent[a];
    while (p != a) {
        const int gp = parent[p];
        if (gp != p) parent[a] = gp; // path halving; parents only ever decrease
        a = p; p = gp;
    }
    return a;
}
__device__ __forceinline__ void uf_union(int32_t *parent, int a, int b) {
    for (;;) {
        a = uf_find(parent, a); b = uf_find(parent, b);
        if (a == b) return;
        if (a < b) { const int tmp = a; a = b; b = tmp; }
        if (atomicCAS(&parent[a], a, b) == a) return;
    }
}
__global__ void k_cc_init(int64_t ncells, int32_t *parent) {
    const int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (c < ncells) parent[c] = (int32_t)c;
}
__global__ void __launch_bounds__(EW_THREADS)
k_cc_hook(int n, double tol, const double *__restrict__ value, int32_t *parent) {
    const int i = blockIdx.y;
    const int64_t rowbase = rm_start(i, n);
#pragma unroll
    for (int q = 0; q < EW_PER_THREAD; ++q) {
        const int off = blockIdx.x * EW_SPAN + q * EW_THREADS + threadIdx.x;
        if (off >= n - i) continue;
        const int64_t v = rowbase + off;
        const double x = value[v];
        if (off + 1 < n - i && fabs(x - value[v + 1]) < tol) uf_union(parent, (int)v, (int)(v + 1));
        if (off >= 1) { // cell (i+1, j) exists iff j >= i+1
            const int64_t dn = v + (n - i - 1);
            if (fabs(x - value[dn]) < tol) uf_union(parent, (int)v, (int)dn);
        }
    }
}
// Only the thread owning cell c writes parent[c] here (no path halving: a concurrent halving store could
// overwrite an already flattened entry with a non-root ancestor).
__global__ void k_cc_flatten(int64_t ncells, int32_t *parent) {
    const int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= ncells) return;
    int a = (int)c, p = __ldcg(&parent[a]);
    while (p != a) { a = p; p = __ldcg(&parent[a]); }
    parent[c] = a;
}
// exclusive scan of is_root in blocks of 1024 threads x 4 items
constexpr int SCAN_ITEMS = 4, SCAN_THREADS = 1024, SCAN_SPAN = SCAN_ITEMS * SCAN_THREADS;
__global__ void __launch_bounds__(SCAN_THREADS)
k_scan_local(int64_t ncells, const int32_t *__restrict__ parent, int32_t *__restrict__ scan, int32_t *__restrict__ blocktot) {
    __shared__ int s_w[32];
    const int64_t c0 = (int64_t)blockIdx.x * SCAN_SPAN + (int64_t)threadIdx.x * SCAN_ITEMS;
    int f[SCAN_ITEMS], tot = 0;
#pragma unroll
    for (int q = 0; q < SCAN_ITEMS; ++q) { const int64_t c = c0 + q; f[q] = (c < ncells && parent[c] == (int32_t)c) ? 1 : 0; tot += f[q]; }
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    int inc = tot;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const int v = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += v; }
    if (lane == 31) s_w[wid] = inc;
    __syncthreads();
    if (wid == 0) {
        int v = s_w[lane], w = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const int u = __shfl_up_sync(0xffffffffu, w, o); if (lane >= o) w += u; }
        s_w[lane] = w - v; // exclusive
        if (lane == 31) blocktot[blockIdx.x] = w;
    }
    __syncthreads();
    int run = s_w[wid] + inc - tot;
#pragma unroll
    for (int q = 0; q < SCAN_ITEMS; ++q) { const int64_t c = c0 + q; if (c < ncells) scan[c] = run; run += f[q]; }
}
__global__ void __launch_bounds__(1024)
k_scan_blocktot(int nblocks, int32_t *blocktot) { // in-place exclusive scan by one block
    __shared__ int s_w[32];
    __shared__ int s_carry;
    if (threadIdx.x == 0) s_carry = 0;
    __syncthreads();
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    for (int base = 0; base < nblocks; base += 1024) {
        const int k = base + threadIdx.x;
        const int v = k < nblocks ? blocktot[k] : 0;
        int inc = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const int u = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += u; }
        if (lane == 31) s_w[wid] = inc;
        __syncthreads();
        if (wid == 0) {
            int w = s_w[lane], w0 = w;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) { const int u = __shfl_up_sync(0xffffffffu, w, o); if (lane >= o) w += u; }
            s_w[lane] = w - w0;
        }
        __syncthreads();
        const int excl = s_carry + s_w[wid] + inc - v;
        if (k < nblocks) blocktot[k] = excl;
        __syncthreads();
        if (threadIdx.x == 1023) s_carry = excl + v;
        __syncthreads();
    }
}
__global__ void k_cc_label(int64_t ncells, const int32_t *__restrict__ parent, const int32_t *__restrict__ scan,
                           const int32_t *__restrict__ blocktot, int32_t *__restrict__ labels) {
    const int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (c < ncells) { const int r = parent[c]; labels[c] = scan[r] + blocktot[r / SCAN_SPAN]; }
}
__global__ void k_fill_i32(int32_t *p, int64_t n, int32_t v) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) p[i] = v;
}


// --------------------------------------------------------------------------
// per-cell fallback of the bias / decay leg passes (inputs whose pos / distance are not functions of the
// bin / diagonal, binned cell by cell like util.cpp:56-111): the pair (a, b) of every cell, then one warp per leg bin
// sums the cells of its list (ascending cell order, lane-strided, butterfly: a fixed shape, same bits on every run)
// --------------------------------------------------------------------------
template <int MODE>
__global__ void __launch_bounds__(EW_THREADS)
k_cell_ab(Tri t, LegTables L, double alpha, const int32_t *__restrict__ obs, const int32_t *__restrict__ nobs,
          const double *__restrict__ phi, double *__restrict__ ca, double *__restrict__ cb) {
    EW_FOREACH_CELL(t) {
        const int64_t idx = rowbase + off;
        double a = 0.0, b = 0.0;
        const int no = nobs[idx];
        if (MODE == MODE_NOBS) a = (double)no;
        else {
            const double eta = log_expected(L, i, j, d, idx, phi[idx]);
            if (MODE == MODE_RESID) { double z, w; nb_residual(eta, obs[idx], no, alpha, z, w); a = w; b = z * w; }
            else { a = (double)obs[idx]; b = exp(eta) * (double)no; }
        }
        ca[idx] = a; cb[idx] = b;
    }
}
// list entry = cell index, bit 31 set when the cell counts twice (both selector entries of the bias binner in this bin)
__global__ void __launch_bounds__(128)
k_bin_sum(int nb, const int64_t *__restrict__ off, const uint32_t *__restrict__ list, const double *__restrict__ ca,
          const double *__restrict__ cb, double *__restrict__ items) {
    const int bin = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (bin >= nb) return;
    double a = 0.0, b = 0.0;
    for (int64_t k = off[bin] + lane; k < off[bin + 1]; k += 32) {
        const uint32_t e = list[k];
        const uint32_t c = e & 0x7fffffffu;
        const double m = (e >> 31) ? 2.0 : 1.0;
        a += m * ca[c]; b += m * cb[c];
    }
    a = warp_sum(a); b = warp_sum(b);
    if (lane == 0) { items[2 * bin] = a; items[2 * bin + 1] = b; }
}

// exposure leg of the per-cell fallback: the cells of a dataset are one contiguous range; fixed-shape sum
__global__ void __launch_bounds__(1024)
k_range_sum(int64_t ncells, const double *__restrict__ ca, const double *__restrict__ cb, double *__restrict__ items) {
    __shared__ double s[32];
    const int d = blockIdx.x;
    double a = 0.0, b = 0.0;
    for (int64_t i = threadIdx.x; i < ncells; i += blockDim.x) { a += ca[(int64_t)d * ncells + i]; b += cb[(int64_t)d * ncells + i]; }
    const double ra = block_sum(a, s);
    const double rb = block_sum(b, s);
    if (threadIdx.x == 0) { items[2 * d] = ra; items[2 * d + 1] = rb; }
}

// checkerboard cross-validation: sum and count of chisq = w (phihat - soft(beta, lam1))^2 + lam1^2 over the cells of
// group g = (bin1 + bin2) % 2 (ScorePreparator.ipp:8-16, Likelihoods.hpp:25-42 with eCprime = 0); one partial per block
__global__ void __launch_bounds__(EW_THREADS)
k_cv_group_sum(int n, int g, const double *__restrict__ phihat, const double *__restrict__ w, const double *__restrict__ beta,
               double lam1, double *__restrict__ partial) {
    __shared__ double s[32];
    double acc = 0.0, cnt = 0.0;
    for (int i = blockIdx.x; i < n; i += gridDim.x) {
        const int64_t rowbase = rm_start(i, n);
        for (int off = threadIdx.x; off < n - i; off += EW_THREADS) {
            const int j = i + off;
            if ((((i + 1) + (j + 1)) & 1) != g) continue;
            const int64_t idx = rowbase + off;
            const double b = beta[idx];
            const double soft = b > 0 ? fmax(0., b - lam1) : fmin(0., b + lam1);
            const double dd = phihat[idx] - soft;
            acc += w[idx] * (dd * dd) + lam1 * lam1;
            cnt += 1.0;
        }
    }
    const double ra = block_sum(acc, s);
    const double rc = block_sum(cnt, s);
    if (threadIdx.x == 0) { partial[2 * blockIdx.x] = ra; partial[2 * blockIdx.x + 1] = rc; }
}
__global__ void __launch_bounds__(1024)
k_cv_fin(int nblocks, const double *__restrict__ partial, double *__restrict__ out2) {
    __shared__ double s[32];
    double a = 0.0, c = 0.0;
    for (int k = threadIdx.x; k < nblocks; k += blockDim.x) { a += partial[2 * k]; c += partial[2 * k + 1]; }
    const double ra = block_sum(a, s);
    const double rc = block_sum(c, s);
    if (threadIdx.x == 0) { out2[0] = ra; out2[1] = rc; }
}

inline dim3 ew_grid(const Tri &t) { return dim3((unsigned)((t.n + EW_SPAN - 1) / EW_SPAN), (unsigned)t.n, (unsigned)t.D); }
inline dim3 tile_grid(const Tri &t) { return dim3((unsigned)tiles_j(t.n), (unsigned)tiles_i(t.n), (unsigned)t.D); }

template <int LEG>
void leg_pass_mode(int mode, const Tri &t, LegTables L, double alpha, const int32_t *obs, const int32_t *nobs,
                   const double *phi, double *part, int grp_rank, int grp_world, cudaStream_t st) {
    const dim3 g = tile_grid(t);
    switch (mode) {
    case MODE_RESID: k_leg_pass<LEG, MODE_RESID><<<g, TILE_J, 0, st>>>(t, L, alpha, obs, nobs, phi, part, grp_rank, grp_world); break;
    case MODE_POISSON: k_leg_pass<LEG, MODE_POISSON><<<g, TILE_J, 0, st>>>(t, L, alpha, obs, nobs, phi, part, grp_rank, grp_world); break;
    default: k_leg_pass<LEG, MODE_NOBS><<<g, TILE_J, 0, st>>>(t, L, alpha, obs, nobs, phi, part, grp_rank, grp_world); break;
    }
}

} // namespace

size_t leg_part_doubles(const Tri &t) {
    const size_t ntI = tiles_i(t.n), ntJ = tiles_j(t.n), D = t.D, n = t.n;
    const size_t e = D * ntI * ntJ * 2, b = D * n * (ntI + ntJ) * 2, dc = D * ntI * ntJ * TILE_DIAGS * 2;
    return std::max(e, std::max(b, dc));
}

void leg_pass_groups(int leg, int mode, const Tri &t, LegTables L, double alpha, const int32_t *obs, const int32_t *nobs,
                     const double *phi, double *part, double *items_g, int grp_rank, int grp_world, cudaStream_t st) {
    const int ntI = tiles_i(t.n), ntJ = tiles_j(t.n);
    if (leg == LEG_EXPO) {
        leg_pass_mode<LEG_EXPO>(mode, t, L, alpha, obs, nobs, phi, part, grp_rank, grp_world, st);
        k_fin_expo<<<dim3(t.D, LEG_GROUPS), 256, 0, st>>>(t, ntI, ntJ, part, items_g, grp_rank, grp_world);
    } else if (leg == LEG_BIAS) {
        leg_pass_mode<LEG_BIAS>(mode, t, L, alpha, obs, nobs, phi, part, grp_rank, grp_world, st);
        k_fin_bias<<<dim3((t.n + 3) / 4, LEG_GROUPS), 128, 0, st>>>(t, ntI, ntJ, part, items_g, grp_rank, grp_world);
    } else {
        leg_pass_mode<LEG_DECAY>(mode, t, L, alpha, obs, nobs, phi, part, grp_rank, grp_world, st);
        k_fin_decay<<<dim3((t.n + 3) / 4, LEG_GROUPS), 128, 0, st>>>(t, ntI, ntJ, part, items_g, grp_rank, grp_world);
    }
    FB_LAUNCH_CHECK();
}

void leg_pass_combine(int nitems, const double *items_g, double *items, int grp_world, cudaStream_t st) {
    k_sum_groups<<<(2 * nitems + 255) / 256, 256, 0, st>>>(nitems, items_g, items, grp_world);
    FB_LAUNCH_CHECK();
}

void leg_pass(int leg, int mode, const Tri &t, LegTables L, double alpha, const int32_t *obs, const int32_t *nobs,
              const double *phi, double *part, double *items_g, double *items, cudaStream_t st) {
    leg_pass_groups(leg, mode, t, L, alpha, obs, nobs, phi, part, items_g, 0, 1, st);
    leg_pass_combine(leg == LEG_EXPO ? t.D : t.n, items_g, items, 1, st);
}

void leg_pass_cells(int mode, const Tri &t, LegTables L, double alpha, const int32_t *obs, const int32_t *nobs, const double *phi,
                    int nb, const int64_t *off, const uint32_t *list, double *ca, double *cb, double *items, cudaStream_t st) {
    switch (mode) {
    case MODE_RESID: k_cell_ab<MODE_RESID><<<ew_grid(t), EW_THREADS, 0, st>>>(t, L, alpha, obs, nobs, phi, ca, cb); break;
    case MODE_POISSON: k_cell_ab<MODE_POISSON><<<ew_grid(t), EW_THREADS, 0, st>>>(t, L, alpha, obs, nobs, phi, ca, cb); break;
    default: k_cell_ab<MODE_NOBS><<<ew_grid(t), EW_THREADS, 0, st>>>(t, L, alpha, obs, nobs, phi, ca, cb); break;
    }
    if (off) k_bin_sum<<<(nb + 3) / 4, 128, 0, st>>>(nb, off, list, ca, cb, items);
    else k_range_sum<<<nb, 1024, 0, st>>>(t.ncells, ca, cb, items);      // exposure leg: bin = dataset
    FB_LAUNCH_CHECK();
}

void cv_group_sum(int n, int g, const double *phihat, const double *w, const double *beta, double lam1, double *partial,
                  double *out2, cudaStream_t st) {
    k_cv_group_sum<<<CV_BLOCKS, EW_THREADS, 0, st>>>(n, g, phihat, w, beta, lam1, partial);
    k_cv_fin<<<1, 1024, 0, st>>>(CV_BLOCKS, partial, out2);
    FB_LAUNCH_CHECK();
}

void signal_yw(const Tri &t, LegTables L, double alpha, const int32_t *obs, const int32_t *nobs, const double *phi,
               unsigned group, double *y, double *w, cudaStream_t st, const Share &sh, int md) {
    k_signal_yw<<<ew_grid(t), EW_THREADS, 0, st>>>(t, L, alpha, obs, nobs, phi, group, y, w, sh.rank, sh.world, sh.on() ? md : t.n);
    FB_LAUNCH_CHECK();
}

void center_dataset(int64_t ncells, const int32_t *nobs_d, double *beta_d, double *scratch, cudaStream_t st, const Share &sh) {
    k_center_partial<<<CENTER_BLOCKS, 256, 0, st>>>(ncells, nobs_d, beta_d, scratch, sh.rank, sh.world);
    if (sh.on()) sh.cm->allreduce_sum(scratch, scratch, 2 * CENTER_BLOCKS, st);
    k_center_final<<<1, CENTER_BLOCKS, 0, st>>>(scratch);
    k_center_apply<<<CENTER_BLOCKS, 256, 0, st>>>(ncells, beta_d, scratch);
    FB_LAUNCH_CHECK();
}

void precision(const Tri &t, LegTables Lold, LegTables Lnew, const double *phi_old, const double *phi_new,
               double *scratch, double *out3, cudaStream_t st, const Share &sh) {
    long long *keys = reinterpret_cast<long long *>(scratch);
    k_precision_init<<<1, 1, 0, st>>>(keys);
    k_precision<<<dim3(1, (unsigned)std::min(t.n, 592), (unsigned)t.D), EW_THREADS, 0, st>>>(t, Lold, Lnew, phi_old, phi_new, keys,
                                                                                         sh.rank, sh.world);
    if (sh.on()) { sh.cm->allreduce_i64(keys, 2, true, st); sh.cm->allreduce_i64(keys + 2, 1, false, st); }
    k_precision_fin<<<1, 1, 0, st>>>(keys, out3);
    FB_LAUNCH_CHECK();
}

void remove_degeneracy(const Tri &t, double *phi, long long *keys, cudaStream_t st, const Share &sh) {
    const double INF = INFINITY;
    const int64_t nk = (int64_t)2 * t.n * t.D;
    k_fill_keys<<<(unsigned)((nk + 255) / 256), 256, 0, st>>>(keys, nk, INF);
    long long *rowkey = keys, *diagkey = keys + (int64_t)t.n * t.D;
    k_minima<<<tile_grid(t), TILE_J, 0, st>>>(t, phi, rowkey, diagkey, sh.rank, sh.world);
    if (sh.on()) sh.cm->allreduce_i64(keys, (size_t)nk, false, st);
    k_degen_apply<<<ew_grid(t), EW_THREADS, 0, st>>>(t, phi, rowkey, diagkey);
    FB_LAUNCH_CHECK();
}

void shift_signal(const Tri &t, double *phi, long long *keys, cudaStream_t st, const Share &sh) {
    k_fill_keys<<<(t.D + 255) / 256, 256, 0, st>>>(keys, t.D, INFINITY);
    k_dataset_min<<<ew_grid(t), EW_THREADS, 0, st>>>(t, phi, keys, sh.rank, sh.world);
    if (sh.on()) sh.cm->allreduce_i64(keys, (size_t)t.D, false, st);
    k_shift_apply<<<ew_grid(t), EW_THREADS, 0, st>>>(t, phi, keys);
    FB_LAUNCH_CHECK();
}

void output_signal(const Tri &t, LegTables L, const int32_t *nobs, const double *phi, const double *lam1,
                   double *background, double *biasmat, double *decaymat, double *signal, double *binless,
                   cudaStream_t st) {
    k_output_signal<<<ew_grid(t), EW_THREADS, 0, st>>>(t, L, nobs, phi, lam1, background, biasmat, decaymat, signal, binless);
    FB_LAUNCH_CHECK();
}

void diff_prepare(const Tri &t, LegTables L, double alpha, const int32_t *obs, const int32_t *nobs, const double *phi,
                  int ref0, double *phihat, double *deltahat, double *w, cudaStream_t st) {
    k_diff_prepare<<<ew_grid(t), EW_THREADS, 0, st>>>(t, L, alpha, obs, nobs, phi, ref0, phihat, deltahat, w);
    FB_LAUNCH_CHECK();
}

void diff_finish(const Tri &t, int ref0, const double *phihat, const double *w, const double *delta,
                 double *phi_ref, double *difference, cudaStream_t st) {
    k_diff_finish<<<ew_grid(t), EW_THREADS, 0, st>>>(t, ref0, phihat, w, delta, phi_ref, difference);
    FB_LAUNCH_CHECK();
}

void patch_labels(int n, double tol, const double *value, int32_t *parent, int32_t *scan, int32_t *labels, cudaStream_t st) {
    const int64_t ncells = (int64_t)n * (n + 1) / 2;
    const unsigned g1 = (unsigned)((ncells + 255) / 256);
    const int nblk = (int)((ncells + SCAN_SPAN - 1) / SCAN_SPAN);
    int32_t *blocktot = scan + ncells;
    k_cc_init<<<g1, 256, 0, st>>>(ncells, parent);
    k_cc_hook<<<dim3((unsigned)((n + EW_SPAN - 1) / EW_SPAN), (unsigned)n), EW_THREADS, 0, st>>>(n, tol, value, parent);
    k_cc_flatten<<<g1, 256, 0, st>>>(ncells, parent);
    k_scan_local<<<nblk, SCAN_THREADS, 0, st>>>(ncells, parent, scan, blocktot);
    k_scan_blocktot<<<1, 1024, 0, st>>>(nblk, blocktot);
    k_cc_label<<<g1, 256, 0, st>>>(ncells, parent, scan, blocktot, labels);
    FB_LAUNCH_CHECK();
}

namespace {
// chisq partial sums (fixed grid => same bits on every run) and per-patch selection flags
__global__ void __launch_bounds__(256)
k_sweep_score(int64_t ncells, const double *__restrict__ phihat, const double *__restrict__ w, const double *__restrict__ beta,
              const int32_t *__restrict__ labels, double lam1, double half_tol, int32_t *__restrict__ flags,
              double *__restrict__ partial) {
    __shared__ double s[32];
    double acc = 0.0;
    const int64_t per = (ncells + gridDim.x - 1) / gridDim.x;
    const int64_t lo = per * blockIdx.x, hi = lo + per < ncells ? lo + per : ncells;
    for (int64_t i = lo + threadIdx.x; i < hi; i += blockDim.x) {
        const double b = beta[i];
        const double soft = b > 0 ? fmax(0., b - lam1) : fmin(0., b + lam1); // soft_threshold(beta, 0, lam1), util.cpp:8-18
        const double d = phihat[i] - soft;
        acc += w[i] * (d * d) + lam1 * lam1;
        if (fabs(soft) > half_tol) flags[labels[i]] = 1;
    }
    const double r = block_sum(acc, s);
    if (threadIdx.x == 0) partial[blockIdx.x] = r;
}
__global__ void __launch_bounds__(1024)
k_sweep_fin(int nblocks, const double *__restrict__ partial, int64_t ncells, const int32_t *__restrict__ flags,
            double *__restrict__ out2) {
    __shared__ double s[32];
    double a = 0.0, c = 0.0;
    for (int k = threadIdx.x; k < nblocks; k += blockDim.x) a += partial[k];
    for (int64_t i = threadIdx.x; i < ncells; i += blockDim.x) c += flags[i] ? 1.0 : 0.0;
    const double ra = block_sum(a, s);
    const double rc = block_sum(c, s);
    if (threadIdx.x == 0) { out2[0] = ra; out2[1] = rc; }
}
} // namespace

void sweep_score(int64_t ncells, const double *phihat, const double *w, const double *beta, const int32_t *labels,
                 double lam1, double tol, int32_t *flags, double *partial, double *out2, cudaStream_t st) {
    FB_CUDA(cudaMemsetAsync(flags, 0, ncells * sizeof(int32_t), st));
    k_sweep_score<<<SWEEP_BLOCKS, 256, 0, st>>>(ncells, phihat, w, beta, labels, lam1, tol / 2, flags, partial);
    k_sweep_fin<<<1, 1024, 0, st>>>(SWEEP_BLOCKS, partial, ncells, flags, out2);
    FB_LAUNCH_CHECK();
}

void fill_i32(int32_t *p, int64_t n, int32_t v, cudaStream_t st) {
    k_fill_i32<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(p, n, v);
    FB_LAUNCH_CHECK();
}

} // namespace fb
