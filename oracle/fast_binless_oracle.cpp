/*
 * oracle/fast_binless_oracle.cpp -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * CPU restatement (plain C++17, single thread, no Rcpp / Eigen / Boost) of the
 * reference's fast-binless driver, used as the parity oracle for the CUDA
 * library and as the timed CPU baseline. Every function cites the reference
 * file:line it follows (paths relative to /root/reference/binless/src).
 *
 * PARITY PINNING
 *   - The GFL ADMM solver + 1-D DP (gfl_graph_fl.c, gfl_tf.c): PINNED. This
 *     file is linked either against the reference's own C sources compiled in
 *     place (oracle/_ref/libfbo_ref.so, -DORACLE_USE_REFERENCE_GFL) or against
 *     oracle/gfl_port.c (oracle/_build/libfbo_port.so); tests require the two
 *     to agree bit for bit.
 *   - The C++ driver restated here (fast_binless.cpp & friends): PARITY
 *     UNPINNED. The reference ships no tests / golden vectors, its example
 *     inputs are git-LFS stubs, and its C++ cannot be built here (needs Rcpp,
 *     RcppEigen, Boost; none installed, no network). Third-party arithmetic
 *     restated from its published algorithm: Eigen (CRAN RcppEigen, version
 *     unpinned by binless/DESCRIPTION) SimplicialLLT / sparse products -> dense
 *     Cholesky and ordered sums below; QuadProg++ (Goldfarb-Idnani dual active
 *     set, vendored in the reference but Eigen-based) -> gi_solve() below;
 *     Boost connected_components -> union-find + lowest-vertex numbering.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline/reference
 * legs may load the libraries built from this file.
 */
#define FB_NAME(x) fbo_##x
#include "../include/fastbinless.h"

#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <limits>
#include <map>
#include <set>
#include <memory>
#include <numeric>
#include <stdexcept>
#include <string>
#include <vector>

extern "C" {
#ifdef ORACLE_USE_REFERENCE_GFL
/* the reference's own symbol (gfl_graph_fl.h:39-43), compiled from /root/reference */
int graph_fused_lasso_weight_warm(int n, double *y, double *w, int ntrails, unsigned *trails,
                                  unsigned *breakpoints, double lam, double *alpha, double inflate,
                                  int maxsteps, double converge, double *beta, double *z, double *u);
#define ORACLE_GFL graph_fused_lasso_weight_warm
static const char *kGflKind = "reference";
#else
int gflp_graph_fused_lasso_weight_warm(int n, const double *y, const double *w, int ntrails,
                                       const unsigned *trails, const unsigned *breakpoints,
                                       double lam, double *alpha, double inflate, int maxsteps,
                                       double converge, double *beta, double *z, double *u);
#define ORACLE_GFL gflp_graph_fused_lasso_weight_warm
static const char *kGflKind = "port";
#endif
}

namespace {

typedef std::vector<double> vec;
typedef std::vector<int> ivec;

struct ArgError : std::runtime_error { using std::runtime_error::runtime_error; };
struct NumError : std::runtime_error { using std::runtime_error::runtime_error; };

std::string g_create_error;

/* ------------------------------------------------------------------------
 * Input holder: FastDataCore (FastData.hpp:13-64, FastData.ipp:3-26)
 * ---------------------------------------------------------------------- */
struct Data {
    unsigned nbins = 0, ncells = 0, ndatasets = 0, N = 0;
    ivec name, bin1, pos1, bin2, pos2, observed, nobs;
    vec distance;
    vec beta, betahat, weights; /* log_signal|log_difference, phihat|deltahat, weights */
    vec phi_ref;                /* Difference only */
    int ref0 = -1;              /* Difference only (0-based) */
};

void build_data(Data &d, int nbins, const fb_mat *m) {
    if (!m || m->N <= 0 || nbins <= 0) throw ArgError("empty input");
    if (!m->pos1 || !m->pos2 || !m->observed || !m->nobs || !m->distance)
        throw ArgError("missing input column");
    const int64_t n = m->N;
    d.nbins = nbins;
    d.ncells = d.nbins * (d.nbins + 1) / 2;
    auto cp = [n](const int32_t *p) { return ivec(p, p + n); };
    if (m->name) d.name = cp(m->name);
    if (m->bin1) d.bin1 = cp(m->bin1);
    if (m->bin2) d.bin2 = cp(m->bin2);
    if (!m->name || !m->bin1 || !m->bin2) {
        /* canonical order: what create_empty_matrix_cpp produces (util.cpp:113-146) */
        if (n % d.ncells != 0)
            throw ArgError("Input size should be N(N+1)/2 dense matrix of observed counts for each dataset");
        ivec nm(n), b1(n), b2(n);
        int64_t i = 0;
        for (int ds = 1; ds <= (int)(n / d.ncells); ++ds)
            for (int a = 1; a <= nbins; ++a)
                for (int b = a; b <= nbins; ++b, ++i) { nm[i] = ds; b1[i] = a; b2[i] = b; }
        if (!m->name) d.name = nm;
        if (!m->bin1) d.bin1 = b1;
        if (!m->bin2) d.bin2 = b2;
    }
    d.pos1 = cp(m->pos1); d.pos2 = cp(m->pos2);
    d.observed = cp(m->observed); d.nobs = cp(m->nobs);
    d.distance.assign(m->distance, m->distance + n);
    d.ndatasets = *std::max_element(d.name.begin(), d.name.end());
    d.N = d.ndatasets * d.ncells;
    /* FastData.ipp:20-25 */
    if ((int64_t)d.N != n)
        throw ArgError("Input size should be N(N+1)/2 dense matrix of observed counts for each dataset");
    if (*std::min_element(d.name.begin(), d.name.end()) != 1)
        throw ArgError("Name column values should start at 1");
    if (*std::min_element(d.bin1.begin(), d.bin1.end()) != 1 ||
        *std::min_element(d.bin2.begin(), d.bin2.end()) != 1)
        throw ArgError("bins should range from 1 to the number of bins");
    if (*std::max_element(d.bin1.begin(), d.bin1.end()) != (int)d.nbins ||
        *std::max_element(d.bin2.begin(), d.bin2.end()) != (int)d.nbins)
        throw ArgError("bins should range from 1 to the number of bins");
    d.beta.assign(d.N, 0); d.betahat.assign(d.N, 0); d.weights.assign(d.N, 0);
}

/* FastData<Difference>::get_log_signal / set_log_signal (FastData.cpp:6-31) */
vec diff_get_log_signal(const Data &d) {
    vec r(d.N);
    for (unsigned i = 0; i < d.N; ++i) r[i] = d.phi_ref[i] + d.beta[i];
    return r;
}
void diff_set_log_signal(Data &d, const vec &log_signal) {
    vec phi_ref; phi_ref.reserve(d.N);
    for (unsigned ds = 0; ds < d.ndatasets; ++ds)
        phi_ref.insert(phi_ref.end(), log_signal.begin() + (size_t)d.ref0 * d.ncells,
                       log_signal.begin() + (size_t)(d.ref0 + 1) * d.ncells);
    for (unsigned i = 0; i < d.N; ++i) d.beta[i] = log_signal[i] - phi_ref[i];
    d.phi_ref.swap(phi_ref);
}

/* ------------------------------------------------------------------------
 * Binner: make_even_bins / bin_data / bin_data_evenly (util.cpp:56-111)
 * The 0/1 sparse selector matrix is kept as one bin index per datum.
 * ---------------------------------------------------------------------- */
struct BinDesign { unsigned Nbins; std::map<double, unsigned> map; };

BinDesign make_even_bins(double lower, double upper, unsigned Nbins) {
    double size = upper - lower;
    BinDesign bd; bd.Nbins = Nbins;
    for (unsigned i = 0; i <= Nbins; ++i) {
        double begin = lower + i * size / double(Nbins);
        bd.map[begin] = i;
    }
    return bd;
}

/* returns per-datum bin index after dropping unused bins; nrows = bins kept */
ivec bin_data(const vec &data, const BinDesign &design, bool drop, unsigned &nrows) {
    const size_t n = data.size();
    ivec idx(n);
    for (size_t i = 0; i < n; ++i) {
        auto it = design.map.upper_bound(data[i]);
        unsigned j;
        if (it == design.map.end()) j = design.map.rbegin()->second - 1;
        else j = it->second - 1;
        idx[i] = (int)j;
    }
    nrows = design.Nbins;
    if (drop) {
        std::vector<char> has(design.Nbins, 0);
        for (size_t i = 0; i < n; ++i) has[idx[i]] = 1;
        ivec remap(design.Nbins, -1);
        unsigned k = 0;
        for (unsigned o = 0; o < design.Nbins; ++o) if (has[o]) remap[o] = (int)k++;
        for (size_t i = 0; i < n; ++i) idx[i] = remap[idx[i]];
        nrows = k;
    }
    return idx;
}

/* ------------------------------------------------------------------------
 * Leg = Summarizer settings + summary + fitter params
 * (fast_summarizer.hpp:36-50, fast_estimator.hpp:19-46)
 * A cell has one selector entry (exposure, decay) or two (bias: bin1 and bin2,
 * fast_bias_mean.hpp:50-52; the diagonal gets coefficient 2).
 * ---------------------------------------------------------------------- */
struct Leg {
    unsigned nb = 0;
    ivec i1, i2;          /* per cell; i2 empty unless bias */
    vec nobs, support;
    double support_min = 0, support_max = 0;
    vec phihat, weight;   /* Summary */

    /* binner * vec, accumulated per bin in ascending cell order like Eigen's
     * row-major sparse * dense product (fast_summarizer.hpp:74) */
    vec summarize(const vec &v) const {
        vec acc(nb, 0.0);
        const size_t n = i1.size();
        if (i2.empty()) {
            for (size_t c = 0; c < n; ++c) acc[i1[c]] += v[c];
        } else {
            for (size_t c = 0; c < n; ++c) {
                if (i1[c] == i2[c]) acc[i1[c]] += 2. * v[c];
                else { acc[i1[c]] += v[c]; acc[i2[c]] += v[c]; }
            }
        }
        return acc;
    }
    /* binner^T * est (fast_estimator.hpp:40-42) */
    vec spread(const vec &est) const {
        const size_t n = i1.size();
        vec r(n);
        if (i2.empty()) for (size_t c = 0; c < n; ++c) r[c] = est[i1[c]];
        else for (size_t c = 0; c < n; ++c)
            r[c] = (i1[c] == i2[c]) ? 2. * est[i1[c]]
                 : (i1[c] < i2[c] ? est[i1[c]] + est[i2[c]] : est[i2[c]] + est[i1[c]]);
        return r;
    }
};

vec to_double(const ivec &v) { return vec(v.begin(), v.end()); }

/* SummarizerSettingsImpl<Exposure,Mean> (fast_exposure.hpp:35-51) */
void build_exposure_leg(Leg &L, const Data &d) {
    vec nm = to_double(d.name);
    L.support_min = *std::min_element(nm.begin(), nm.end());
    L.support_max = *std::max_element(nm.begin(), nm.end());
    unsigned Nb = (unsigned)(L.support_max - L.support_min + 1);
    L.i1 = bin_data(nm, make_even_bins(L.support_min, L.support_max, Nb), true, L.nb);
    vec nobs = to_double(d.nobs);
    L.nobs = L.summarize(nobs);
    vec t(d.N);
    for (unsigned i = 0; i < d.N; ++i) t[i] = nm[i] * nobs[i];
    L.support = L.summarize(t);
    for (unsigned k = 0; k < L.nb; ++k) L.support[k] /= L.nobs[k];
}

/* SummarizerSettingsImpl<Bias,Mean> (fast_bias_mean.hpp:37-68) */
void build_bias_leg(Leg &L, const Data &d, unsigned nbins) {
    vec p1 = to_double(d.pos1), p2 = to_double(d.pos2);
    double mn = std::min(*std::min_element(p1.begin(), p1.end()), *std::min_element(p2.begin(), p2.end()));
    double mx = std::max(*std::max_element(p1.begin(), p1.end()), *std::max_element(p2.begin(), p2.end()));
    L.support_min = mn; L.support_max = mx;
    BinDesign design = make_even_bins(mn, mx, nbins);
    unsigned n1, n2;
    L.i1 = bin_data(p1, design, true, n1);
    L.i2 = bin_data(p2, design, true, n2);
    if (n1 != n2) throw ArgError("bias binner: pos1 and pos2 do not populate the same bins");
    L.nb = n1;
    vec nobs = to_double(d.nobs);
    L.nobs = L.summarize(nobs);
    for (auto &x : L.nobs) x = x / 2.;
    /* support = (mean pos1 per bin + mean pos2 per bin)/2 */
    vec s1(L.nb, 0), c1(L.nb, 0), s2(L.nb, 0), c2(L.nb, 0);
    for (unsigned i = 0; i < d.N; ++i) {
        s1[L.i1[i]] += p1[i]; c1[L.i1[i]] += 1.;
        s2[L.i2[i]] += p2[i]; c2[L.i2[i]] += 1.;
    }
    L.support.resize(L.nb);
    for (unsigned k = 0; k < L.nb; ++k) L.support[k] = (s1[k] / c1[k] + s2[k] / c2[k]) / 2.;
}

/* SummarizerSettingsImpl<Decay,GAM> (fast_decay.hpp:48-68) */
void build_decay_leg(Leg &L, const Data &d, unsigned K, unsigned bins_per_bf) {
    vec ld(d.N);
    for (unsigned i = 0; i < d.N; ++i) ld[i] = std::log(d.distance[i]);
    L.support_min = *std::min_element(ld.begin(), ld.end());
    L.support_max = *std::max_element(ld.begin(), ld.end());
    L.i1 = bin_data(ld, make_even_bins(L.support_min, L.support_max, K * bins_per_bf), true, L.nb);
    L.nobs = L.summarize(to_double(d.nobs));
    vec s = L.summarize(ld), c = L.summarize(vec(d.N, 1.0));
    L.support.resize(L.nb);
    for (unsigned k = 0; k < L.nb; ++k) L.support[k] = s[k] / c[k];
}

/* ------------------------------------------------------------------------
 * Dense helpers for the K=50 GAM
 * ---------------------------------------------------------------------- */
struct DMat {
    int r = 0, c = 0; vec a;
    DMat() {}
    DMat(int r_, int c_) : r(r_), c(c_), a((size_t)r_ * c_, 0.0) {}
    double &operator()(int i, int j) { return a[(size_t)i * c + j]; }
    double operator()(int i, int j) const { return a[(size_t)i * c + j]; }
};

/* cardinal_cubic_bspline_design + generate_spline_base (spline.cpp:4-26).
 * Column shift by the permutation R is B(:,i) <- B(:,i+1) (wrapping). */
DMat generate_spline_base(const vec &x, double xmin, double xmax, unsigned K) {
    const unsigned degree = 3;
    const double dx = 1.01 * (xmax - xmin) / (K - degree);
    const double lower = xmin - dx * 0.01;
    const unsigned nk = K + degree + 1;
    /* Eigen LinSpaced(size, low, high): low + i*(high-low)/(size-1), last = high */
    vec knots(nk);
    {
        const double low = lower - degree * dx, high = lower + K * dx;
        const double step = (high - low) / double(nk - 1);
        for (unsigned i = 0; i < nk; ++i) knots[i] = (i == nk - 1) ? high : low + i * step;
    }
    const unsigned N = x.size();
    DMat out(N, K);
    vec B(nk), Bn(nk), P(nk);
    for (unsigned r = 0; r < N; ++r) {
        for (unsigned j = 0; j < nk; ++j) {
            P[j] = (x[r] - knots[j]) / dx;
            B[j] = (knots[j] <= x[r] && x[r] < knots[j] + dx) ? 1.0 : 0.0;
        }
        for (unsigned k = 1; k <= degree; ++k) {
            for (unsigned j = 0; j < nk; ++j) {
                double shifted = B[(j + 1) % nk];
                Bn[j] = (P[j] * B[j] + (k + 1 - P[j]) * shifted) / k;
            }
            B.swap(Bn);
        }
        for (unsigned j = 0; j < K; ++j) out(r, j) = B[j];
    }
    return out;
}

/* In-place dense Cholesky A = L L^T (lower). Stands in for Eigen::SimplicialLLT
 * with NaturalOrdering (QuadProgGAMLibrary.cpp:15-22). */
bool cholesky(const DMat &A, DMat &L) {
    const int n = A.r;
    L = DMat(n, n);
    for (int j = 0; j < n; ++j) {
        double s = A(j, j);
        for (int k = 0; k < j; ++k) s -= L(j, k) * L(j, k);
        if (!(s > 0)) return false;
        L(j, j) = std::sqrt(s);
        for (int i = j + 1; i < n; ++i) {
            double t = A(i, j);
            for (int k = 0; k < j; ++k) t -= L(i, k) * L(j, k);
            L(i, j) = t / L(j, j);
        }
    }
    return true;
}
void solve_lower(const DMat &L, vec &b) { /* L x = b */
    const int n = L.r;
    for (int i = 0; i < n; ++i) {
        double s = b[i];
        for (int k = 0; k < i; ++k) s -= L(i, k) * b[k];
        b[i] = s / L(i, i);
    }
}
void solve_lower_T(const DMat &L, vec &b) { /* L^T x = b */
    const int n = L.r;
    for (int i = n - 1; i >= 0; --i) {
        double s = b[i];
        for (int k = i + 1; k < n; ++k) s -= L(k, i) * b[k];
        b[i] = s / L(i, i);
    }
}

/* Goldfarb-Idnani dual active-set method for
 *     min 1/2 x'Gx + g0'x   s.t.  CI' x >= 0      (ci0 = 0, no equalities)
 * given G = L L' and x = unconstrained minimiser on entry; this is what
 * QuadProgPP::solve_quadprog_with_guess computes for the decay leg
 * (QuadProg++.cpp:165-, called at QuadProgGAMLibrary.cpp:56-59). The problem is
 * strictly convex so the minimiser is unique; any exact solver agrees to
 * rounding. CI is n x m (one constraint per column).                          */
void gi_solve(const DMat &L, const vec &g0, const DMat &CI, vec &x) {
    const int n = L.r, m = CI.c;
    const double eps = std::numeric_limits<double>::epsilon();
    const double inf = std::numeric_limits<double>::infinity();
    (void)g0;
    DMat J(n, n), R(n, n);
    vec d(n, 0.0), zv(n), np(n), s(m), rv(m + 1, 0.0), u(m + 1, 0.0);
    ivec A(m + 1, 0);
    std::vector<char> active(m, 0);
    double c1 = 0, c2 = 0, Rnorm = 1.0;
    int iq = 0;
    for (int i = 0; i < n; ++i) c1 += L(i, i) * L(i, i);
    for (int i = 0; i < n; ++i) { /* J = L^-T */
        vec e(n, 0.0); e[i] = 1.0;
        solve_lower_T(L, e);
        for (int k = 0; k < n; ++k) J(k, i) = e[k];
        c2 += J(i, i);
    }
    auto slack = [&](int i) { double sum = 0; for (int k = 0; k < n; ++k) sum += CI(k, i) * x[k]; return sum; };
    /* plane rotation that maps (a,b) -> (h,0); applied to a pair of vectors */
    struct Rot { double cc, sn, xny, h; bool skip; };
    auto make_rot = [](double a, double b) {
        Rot r; r.h = std::hypot(a, b); r.skip = (r.h == 0.0);
        if (r.skip) return r;
        r.cc = a / r.h; r.sn = b / r.h;
        if (r.cc < 0.0) { r.cc = -r.cc; r.sn = -r.sn; r.h = -r.h; }
        r.xny = r.sn / (1.0 + r.cc);
        return r;
    };
    auto add_constraint = [&]() { /* appends np (d = J'np) to the active set */
        for (int j = n - 1; j >= iq + 1; --j) {
            Rot r = make_rot(d[j - 1], d[j]);
            if (r.skip) continue;
            d[j] = 0.0; d[j - 1] = r.h;
            for (int k = 0; k < n; ++k) {
                double a = J(k, j - 1), b = J(k, j);
                J(k, j - 1) = a * r.cc + b * r.sn;
                J(k, j) = r.xny * (a + J(k, j - 1)) - b;
            }
        }
        ++iq;
        for (int i = 0; i < iq; ++i) R(i, iq - 1) = d[i];
        if (std::fabs(d[iq - 1]) <= eps * Rnorm) throw NumError("quadratic program: dependent constraints");
        Rnorm = std::max(Rnorm, std::fabs(d[iq - 1]));
    };
    auto drop_constraint = [&](int l) { /* removes constraint l; slot iq (candidate) moves down */
        int qq = -1;
        for (int i = 0; i < iq; ++i) if (A[i] == l) { qq = i; break; }
        for (int i = qq; i < iq - 1; ++i) {
            A[i] = A[i + 1]; u[i] = u[i + 1];
            for (int j = 0; j < n; ++j) R(j, i) = R(j, i + 1);
        }
        A[iq - 1] = A[iq]; u[iq - 1] = u[iq];
        A[iq] = 0; u[iq] = 0.0;
        for (int j = 0; j < iq; ++j) R(j, iq - 1) = 0.0;
        --iq;
        active[l] = 0;
        for (int j = qq; j < iq; ++j) {
            Rot r = make_rot(R(j, j), R(j + 1, j));
            if (r.skip) continue;
            R(j + 1, j) = 0.0; R(j, j) = r.h;
            for (int k = j + 1; k < iq; ++k) {
                double a = R(j, k), b = R(j + 1, k);
                R(j, k) = a * r.cc + b * r.sn;
                R(j + 1, k) = r.xny * (a + R(j, k)) - b;
            }
            for (int k = 0; k < n; ++k) {
                double a = J(k, j), b = J(k, j + 1);
                J(k, j) = a * r.cc + b * r.sn;
                J(k, j + 1) = r.xny * (J(k, j) + a) - b;
            }
        }
    };

    for (int iter = 0; iter <= 10 * m; ++iter) { /* QuadProg++.cpp:289 */
        /* step 1: most violated inactive constraint; psi sums the inactive set only (:314-318),
         * tolerance (m - iq) eps c1 c2 100 (:328) */
        double psi = 0;
        for (int i = 0; i < m; ++i) { s[i] = slack(i); if (!active[i]) psi += std::min(0.0, s[i]); }
        if (std::fabs(psi) <= (m - iq) * eps * c1 * c2 * 100.0) return;
        int ip = -1; double ss = 0.0;
        for (int i = 0; i < m; ++i) if (!active[i] && s[i] < ss) { ss = s[i]; ip = i; }
        if (ip < 0) return;
        for (int k = 0; k < n; ++k) np[k] = CI(k, ip);
        u[iq] = 0.0; A[iq] = ip;
        for (;;) { /* step 2: move towards satisfying ip, dropping blockers */
            for (int i = 0; i < n; ++i) { double sum = 0; for (int k = 0; k < n; ++k) sum += J(k, i) * np[k]; d[i] = sum; }
            for (int i = 0; i < n; ++i) { double sum = 0; for (int j = iq; j < n; ++j) sum += J(i, j) * d[j]; zv[i] = sum; }
            for (int i = iq - 1; i >= 0; --i) {
                double sum = d[i];
                for (int j = i + 1; j < iq; ++j) sum -= R(i, j) * rv[j];
                rv[i] = sum / R(i, i);
            }
            int l = -1; double t1 = inf;
            for (int k = 0; k < iq; ++k) if (rv[k] > 0.0 && u[k] / rv[k] < t1) { t1 = u[k] / rv[k]; l = A[k]; }
            double zz = 0, t2 = inf;
            for (int k = 0; k < n; ++k) zz += zv[k] * zv[k];
            if (std::fabs(zz) > eps) { double zn = 0; for (int k = 0; k < n; ++k) zn += zv[k] * np[k]; t2 = -s[ip] / zn; }
            const double t = std::min(t1, t2);
            if (t >= inf) throw NumError("quadratic program is infeasible");
            if (t2 >= inf) { /* step in the dual space only */
                for (int k = 0; k < iq; ++k) u[k] -= t * rv[k];
                u[iq] += t;
                drop_constraint(l);
                continue;
            }
            for (int k = 0; k < n; ++k) x[k] += t * zv[k];
            for (int k = 0; k < iq; ++k) u[k] -= t * rv[k];
            u[iq] += t;
            if (t == t2) { add_constraint(); active[ip] = 1; break; } /* full step */
            drop_constraint(l);                                        /* partial step */
            s[ip] = slack(ip);
        }
    }
    /* iteration cap reached: the reference leaves the loop and returns the current iterate */
}

/* ------------------------------------------------------------------------
 * Decay GAM: FitterSettingsImpl<Decay,GAM> (fast_decay.hpp:71-89) +
 * QuadProgGAMLibrary (QuadProgGAMLibrary.cpp:7-73) + Params<GAM>
 * ---------------------------------------------------------------------- */
struct DecayGam {
    unsigned K = FB_DECAY_K, max_iter = 100;
    double sigma = 1, tol_val = 0.2;
    DMat X;          /* nb x K */
    DMat CI;         /* K x nin, one constraint per column: CI' beta >= 0 */
    int nin = 0;
    DMat K2DtD;      /* K x K */
    /* library state (persists across optimize calls; QuadProgGAMLibrary.cpp:12) */
    vec qp_beta; double qp_lambda = 1;
    /* Params<GAM> (fast_fitter_gam.hpp:20) */
    vec beta; double lambda = -1, mean = 0;

    void setup(const Leg &L, double tol, unsigned free_decay) {
        tol_val = tol;
        X = generate_spline_base(L.support, L.support_min, L.support_max, K);
        /* second_order_difference_matrix (gam.cpp:15-21): rows e_i - 2e_{i+1} + e_{i+2} */
        DMat D(K - 2, K);
        for (unsigned i = 0; i + 2 < K; ++i) { D(i, i) = 1; D(i, i + 1) = -2; D(i, i + 2) = 1; }
        /* K2DtD = (K*D')(K*D')' (QuadProgGAMLibrary.cpp:30-31) */
        K2DtD = DMat(K, K);
        for (unsigned a = 0; a < K; ++a)
            for (unsigned b = 0; b < K; ++b) {
                double s = 0;
                for (unsigned r = 0; r + 2 < K; ++r) s += (K * D(r, a)) * (K * D(r, b));
                K2DtD(a, b) = s;
            }
        /* decreasing_constraint(K, free_first) (gam.cpp:6-13, fast_decay.hpp:85-87):
         * rows Kfree..K-2 of -(first difference) => beta_i - beta_{i+1} >= 0 */
        unsigned free_first = (unsigned)((double)K * (std::log((double)free_decay) - L.support_min)
                                         / (L.support_max - L.support_min));
        if (K <= 1 || free_first >= K) { nin = 0; CI = DMat(K, 0); }
        else {
            nin = (int)(K - 1 - free_first);
            CI = DMat(K, nin);
            for (int c = 0; c < nin; ++c) { unsigned i = free_first + c; CI(i, c) = 1; CI(i + 1, c) = -1; }
        }
        qp_beta.assign(K, 1.0); qp_lambda = 1;
        beta.assign(K, 0.0); lambda = -1; mean = 0;
    }

    vec X_times(const vec &b) const {
        vec r(X.r, 0.0);
        for (int i = 0; i < X.r; ++i) {
            double s = 0;
            for (unsigned k = 0; k < K; ++k) { double xv = X(i, k); if (xv != 0) s += xv * b[k]; }
            r[i] = s;
        }
        return r;
    }
    vec get_estimate() const { return X_times(beta); }
    vec get_binned_estimate() const { /* fast_fitter.hpp:31-33 */
        vec e = get_estimate();
        for (auto &v : e) v -= mean;
        return e;
    }

    /* QuadProgGAMLibrary::optimize (QuadProgGAMLibrary.cpp:24-73) */
    void optimize(const vec &y, const vec &Sm1) {
        const int nb = X.r;
        DMat XtSm2X(K, K); vec XtSm2y(K, 0.0);
        for (int r = 0; r < nb; ++r) {
            const double s = Sm1[r], sy = s * y[r];
            for (unsigned a = 0; a < K; ++a) {
                double ta = s * X(r, a);
                if (ta == 0) continue;
                XtSm2y[a] += ta * sy;
                for (unsigned b = 0; b <= a; ++b) XtSm2X(a, b) += ta * (s * X(r, b));
            }
        }
        for (unsigned a = 0; a < K; ++a) for (unsigned b = a + 1; b < K; ++b) XtSm2X(a, b) = XtSm2X(b, a);
        vec beta_old = qp_beta;
        unsigned iter = 0; bool converged = false;
        do {
            DMat A(K, K);
            for (unsigned a = 0; a < K; ++a)
                for (unsigned b = 0; b < K; ++b)
                    A(a, b) = XtSm2X(a, b) + qp_lambda * qp_lambda * K2DtD(a, b) + (a == b ? 1e-10 : 0.0);
            DMat L;
            if (!cholesky(A, L)) throw NumError("decomposition failed");
            vec g0(K); for (unsigned a = 0; a < K; ++a) g0[a] = -XtSm2y[a];
            qp_beta = g0;
            solve_lower(L, qp_beta); solve_lower_T(L, qp_beta);
            for (auto &v : qp_beta) v = -v;
            if (nin > 0) gi_solve(L, g0, CI, qp_beta);
            /* lambda (QuadProgGAMLibrary.cpp:62) */
            double sq = 0;
            for (unsigned r = 0; r + 2 < K; ++r) {
                double t = K * (qp_beta[r] - 2 * qp_beta[r + 1] + qp_beta[r + 2]);
                sq += t * t;
            }
            qp_lambda = std::sqrt((K - 2) / (sq + 1. / (sigma * sigma)));
            double mx = *std::max_element(qp_beta.begin(), qp_beta.end());
            double mn = *std::min_element(qp_beta.begin(), qp_beta.end());
            double precision = mx - mn;
            if (precision > 0) {
                double dm = 0;
                for (unsigned a = 0; a < K; ++a) dm = std::max(dm, std::fabs(qp_beta[a] - beta_old[a]));
                precision = dm / precision;
            }
            converged = precision < tol_val;
            beta_old = qp_beta;
            iter++;
        } while (iter < max_iter && !converged);
    }

    /* FitterImpl<Leg,GAM>::update_params (fast_fitter_gam.hpp:75-92,108-111) */
    void update_params(const vec &phihat, const vec &weight, const vec &nobs) {
        vec Sm1(weight.size());
        for (size_t i = 0; i < weight.size(); ++i) Sm1[i] = std::sqrt(weight[i]);
        optimize(phihat, Sm1);
        lambda = qp_lambda; beta = qp_beta;
        vec e = get_estimate();
        double num = 0, den = 0;
        for (size_t i = 0; i < e.size(); ++i) { num += nobs[i] * e[i]; den += nobs[i]; }
        mean = num / den;
    }
};

/* FitterImpl<Leg,Mean>::update_params (fast_fitter_mean.hpp:55-79) */
struct MeanFit {
    vec estimate; double mean = 0;
    void update_params(const vec &phihat, const vec &nobs, bool center, bool cap) {
        vec est = phihat;
        if (center) {
            double num = 0, den = 0;
            for (size_t i = 0; i < est.size(); ++i) { num += nobs[i] * est[i]; den += nobs[i]; }
            double avg = num / den;
            for (auto &v : est) v = v - avg;
        }
        if (cap) {
            double sq = 0; for (double v : est) sq += v * v;
            double stdev = std::sqrt(sq / est.size());
            for (auto &v : est) v = std::max(std::min(v, 3 * stdev), -3 * stdev);
        }
        estimate = est;
    }
    vec get_binned_estimate() const { vec e = estimate; for (auto &v : e) v -= mean; return e; }
};

/* ------------------------------------------------------------------------
 * GFL solver object: GFLLibrary + Triangle/Trapeze impl
 * (GFLLibrary.hpp:14-60, GFLLibraryTriangleImpl.cpp, GFLLibraryTrapezeImpl.cpp,
 *  Settings.hpp:13-20)
 * ---------------------------------------------------------------------- */
std::vector<ivec> triangle_grid_chain(int nrows) { /* GFLLibraryTriangleImpl.cpp:7-35 */
    std::vector<ivec> chains;
    int idx = 0;
    for (int i = 0; i < nrows - 1; ++i) { /* rows of consecutive numbers, last cell excluded */
        ivec c(nrows - i);
        std::iota(c.begin(), c.end(), idx);
        idx += nrows - i;
        chains.push_back(c);
    }
    for (int col = 1; col < nrows; ++col) { /* columns: next = cur + (nrows-1-row) */
        ivec c; int cell = col;
        c.push_back(cell);
        for (int row = 0; row < col; ++row) { cell += nrows - 1 - row; c.push_back(cell); }
        chains.push_back(c);
    }
    return chains;
}

std::vector<ivec> trapezoidal_grid_chain(int nrows, int maxdiag) { /* GFLLibraryTrapezeImpl.cpp:8-69 */
    if (maxdiag > (nrows + 1) / 2) throw ArgError("maxdiag is too large for this implementation!\n");
    if (maxdiag <= 1) throw ArgError("maxdiag must be at least 2!\n");
    const int nfull = nrows - maxdiag + 1;
    /* band index of (i,j), 0-based, j-i < maxdiag */
    std::vector<int> base(nrows);
    for (int i = 0, acc = 0; i < nrows; ++i) { base[i] = acc; acc += std::min(maxdiag, nrows - i); }
    std::vector<ivec> chains;
    for (int i = 0; i < nrows; ++i) { /* row chains of length >= 2 */
        int len = std::min(maxdiag, nrows - i);
        if (len < 2) continue;
        ivec c(len); std::iota(c.begin(), c.end(), base[i]);
        chains.push_back(c);
    }
    (void)nfull;
    for (int col = 1; col < nrows; ++col) { /* column chains restricted to the band */
        ivec c;
        for (int row = std::max(0, col - maxdiag + 1); row <= col; ++row) c.push_back(base[row] + (col - row));
        chains.push_back(c);
    }
    return chains;
}

struct GflSolver {
    int nrows = 0, maxdiag = 0; bool trapeze = false;
    unsigned N = 0; int ntrails = 0; unsigned tsz = 0;
    std::vector<unsigned> trails, breakpoints;
    vec beta_impl, z, u, beta_full;
    double alpha = 5.0, inflate = 2.0; int ninner_max = 30000; unsigned counter = 0;
    int last_steps = 0;

    GflSolver(unsigned nr, unsigned md) : nrows(nr) {
        md = std::min(md, nr); /* FusedLassoGaussianEstimator.hpp:33 */
        std::vector<ivec> chains;
        if (md > (nr + 1) / 2 || md <= 1) { /* GFLLibrary.hpp:21-27 */
            trapeze = false; N = nr * (nr + 1) / 2; chains = triangle_grid_chain(nr);
        } else {
            trapeze = true; maxdiag = md;
            N = md * (nr - md + 1) + (md - 1) * md / 2; chains = trapezoidal_grid_chain(nr, md);
        }
        for (auto &c : chains) {
            if (!trails.empty()) breakpoints.push_back(trails.size());
            trails.insert(trails.end(), c.begin(), c.end());
        }
        if (!trails.empty()) breakpoints.push_back(trails.size());
        ntrails = breakpoints.size(); tsz = trails.size();
        reset();
    }
    void reset() { counter = 0; beta_impl.assign(N, 0); z.assign(tsz, 0); u.assign(tsz, 0); }

    unsigned row_base(unsigned i) const { return i * (i + 1) / 2 + i * (nrows - i); }

    /* GFLLibraryTriangleImpl::optimize (:60-71) / GFLLibraryTrapezeImpl::optimize (:133-146) */
    void optimize(const vec &y, const vec &w, double lam2, double converge) {
        if (!trapeze) {
            last_steps = ORACLE_GFL(N, const_cast<double *>(y.data()), const_cast<double *>(w.data()),
                                    ntrails, trails.data(), breakpoints.data(), lam2, &alpha, inflate,
                                    ninner_max, converge, beta_impl.data(), z.data(), u.data());
            beta_full = beta_impl;
        } else {
            vec yt, wt; yt.reserve(N); wt.reserve(N);
            for (int i = 0; i < nrows; ++i) { /* extract_trapeze (:93-106) */
                unsigned b = row_base(i), len = std::min<unsigned>(maxdiag, nrows - i);
                yt.insert(yt.end(), y.begin() + b, y.begin() + b + len);
                wt.insert(wt.end(), w.begin() + b, w.begin() + b + len);
            }
            last_steps = ORACLE_GFL(N, yt.data(), wt.data(), ntrails, trails.data(), breakpoints.data(),
                                    lam2, &alpha, inflate, ninner_max, converge, beta_impl.data(),
                                    z.data(), u.data());
            /* fill_triangle (:108-131) */
            double sum_wy = 0, sum_w = 0;
            for (unsigned i = 0; i < (unsigned)nrows; ++i) {
                unsigned b = row_base(i);
                for (unsigned dd = maxdiag; i + dd < (unsigned)nrows; ++dd) {
                    sum_wy += w[b + dd] * y[b + dd]; sum_w += w[b + dd];
                }
            }
            double avg = sum_wy / sum_w;
            beta_full.assign((size_t)nrows * (nrows + 1) / 2, avg);
            unsigned j = 0;
            for (unsigned i = 0; i < (unsigned)nrows; ++i) {
                unsigned b = row_base(i);
                for (unsigned dd = 0; dd < std::min<unsigned>(maxdiag, nrows - i); ++dd) beta_full[b + dd] = beta_impl[j++];
            }
        }
        counter += last_steps;
    }
    /* FusedLassoGaussianEstimator::get (:49-62) with offset=0, lambda1=0: clamp at +-50 */
    vec get() const {
        vec b = beta_full;
        for (auto &v : b) {
            v = std::min(50., std::max(-50., v));
            v = (v > 0) ? std::max(0., v - 0.) : std::min(0., v + 0.); /* soft_threshold(.,0,0), util.cpp:8-18 */
        }
        return b;
    }
};

/* ------------------------------------------------------------------------
 * The handle: data + three legs (+ restored state)
 * ---------------------------------------------------------------------- */
struct Model {
    Data d;
    Leg expo_leg, bias_leg, decay_leg;
    MeanFit expo, bias;
    DecayGam dec;
    double alpha = 1;
    bool legs_built = false, has_state = false;
    std::string err;
    int (*interrupt_cb)(void *) = nullptr; void *interrupt_arg = nullptr;

    void build_legs(double tol_val, unsigned free_decay) {
        build_exposure_leg(expo_leg, d);
        build_decay_leg(decay_leg, d, FB_DECAY_K, 100);
        build_bias_leg(bias_leg, d, d.nbins);
        expo.estimate.assign(expo_leg.nb, 0); expo.mean = 0;
        bias.estimate.assign(bias_leg.nb, 0); bias.mean = 0;
        dec.setup(decay_leg, tol_val, free_decay);
        legs_built = true;
    }

    /* get_log_expected (fast_expected.ipp:1-19) */
    vec log_expected(const vec &signal) const {
        vec e = expo_leg.spread(expo.get_binned_estimate());
        vec b = bias_leg.spread(bias.get_binned_estimate());
        vec dc = decay_leg.spread(dec.get_binned_estimate());
        vec r(d.N);
        for (unsigned i = 0; i < d.N; ++i) r[i] = b[i] + dc[i] + signal[i] + e[i];
        return r;
    }
    /* NB residuals (fast_residuals.ipp:47-67) */
    void residuals(const vec &signal, vec &res, vec &wts) const {
        vec le = log_expected(signal);
        res.resize(d.N); wts.resize(d.N);
        for (unsigned i = 0; i < d.N; ++i) {
            if (d.nobs[i] > 0) {
                double mu = std::exp(le[i]);
                res[i] = d.observed[i] / (d.nobs[i] * mu) - 1;
                wts[i] = d.nobs[i] / (1 / mu + 1 / alpha);
            } else { res[i] = 0; wts[i] = 0; }
        }
    }
    /* Summarizer::set_poisson_lsq_summary (fast_summarizer.ipp:3-30) */
    static void poisson_lsq(Leg &L, const vec &le, const Data &d, double pseudocount = 0.01) {
        vec obs = to_double(d.observed), nobs = to_double(d.nobs);
        vec sum_obs = L.summarize(obs);
        vec t(d.N);
        for (unsigned i = 0; i < d.N; ++i) t[i] = std::exp(le[i]) * nobs[i];
        vec sum_exp = L.summarize(t);
        L.phihat.resize(L.nb); L.weight.resize(L.nb);
        for (unsigned k = 0; k < L.nb; ++k) {
            double est = std::log(pseudocount + sum_obs[k] / sum_exp[k]);
            if (sum_exp[k] == 0) est = 0;
            L.phihat[k] = est;
            L.weight[k] = std::exp(est) * L.nobs[k];
        }
    }
    /* Summarizer::update_summary (fast_summarizer.ipp:32-51) */
    static void update_summary(Leg &L, const vec &res, const vec &wts, const vec &estimate) {
        vec wsum = L.summarize(wts);
        vec t(res.size());
        for (size_t i = 0; i < res.size(); ++i) t[i] = res[i] * wts[i];
        vec ph = L.summarize(t);
        for (unsigned k = 0; k < L.nb; ++k) {
            ph[k] = ph[k] / wsum[k] + estimate[k];
            if (wsum[k] == 0) ph[k] = 0;
        }
        L.phihat = ph; L.weight = wsum;
    }
};

/* get_precision (fast_binless.cpp:22-33) */
void get_precision(const vec &a, const vec &old, double &abs_, double &rel_) {
    double delta = std::abs(a[0] - old[0]), mx = a[0], mn = a[0];
    for (size_t i = 1; i < a.size(); ++i) {
        delta = std::max(std::abs(a[i] - old[i]), delta);
        mn = std::min(a[i], mn); mx = std::max(a[i], mx);
    }
    abs_ = delta; rel_ = delta / (mx - mn);
}

/* remove_signal_degeneracy (fast_binless.cpp:35-65) */
vec remove_signal_degeneracy(const Data &d) {
    vec ls = d.beta;
    double mxs = *std::max_element(ls.begin(), ls.end());
    std::vector<vec> min_row(d.nbins, vec(d.ndatasets, mxs)), min_diag(d.nbins, vec(d.ndatasets, mxs));
    for (unsigned i = 0; i < d.N; ++i) {
        unsigned nm = d.name[i] - 1, b1 = d.bin1[i] - 1, b2 = d.bin2[i] - 1;
        min_diag[b2 - b1][nm] = std::min(min_diag[b2 - b1][nm], ls[i]);
        min_row[b1][nm] = std::min(min_row[b1][nm], ls[i]);
        if (b1 != b2) min_row[b2][nm] = std::min(min_row[b2][nm], ls[i]);
    }
    for (unsigned i = 0; i < d.N; ++i) {
        unsigned nm = d.name[i] - 1, b1 = d.bin1[i] - 1, b2 = d.bin2[i] - 1;
        double adj = std::max(std::max(min_row[b1][nm], min_row[b2][nm]), min_diag[b2 - b1][nm]);
        ls[i] = std::max(ls[i] - adj, 0.);
    }
    return ls;
}
/* shift_signal (fast_binless.cpp:67-83) */
vec shift_signal(const Data &d) {
    vec ls = d.beta;
    double mxs = *std::max_element(ls.begin(), ls.end());
    vec mn(d.ndatasets, mxs);
    for (unsigned i = 0; i < d.N; ++i) mn[d.name[i] - 1] = std::min(mn[d.name[i] - 1], ls[i]);
    for (unsigned i = 0; i < d.N; ++i) ls[i] = std::max(ls[i] - mn[d.name[i] - 1], 0.);
    return ls;
}

/* get_patch_numbers (graph_helpers.cpp:12-59): components of the graph whose
 * edges join (b1,b2)-(b1,b2-1) and (b1,b2)-(b1+1,b2) when |dvalue| < tol_val;
 * Boost numbers components in order of their lowest vertex. */
void patch_numbers(int nbins, double tol_val, const double *value, int32_t *out) {
    const int n = nbins * (nbins + 1) / 2;
    ivec parent(n);
    std::iota(parent.begin(), parent.end(), 0);
    auto find = [&](int a) { while (parent[a] != a) { parent[a] = parent[parent[a]]; a = parent[a]; } return a; };
    auto unite = [&](int a, int b) { a = find(a); b = find(b); if (a == b) return; if (a < b) parent[b] = a; else parent[a] = b; };
    int v = 1;
    for (int l = nbins - 1; l >= 1; ++v, --l)
        for (int j = 1; j <= l; ++v, ++j) {
            if (std::abs(value[v - 1] - value[v]) < tol_val) unite(v - 1, v);
            if (std::abs(value[v] - value[v + l]) < tol_val) unite(v, v + l);
        }
    ivec label(n, -1); int next = 0;
    for (int i = 0; i < n; ++i) { int r = find(i); if (label[r] < 0) label[r] = next++; out[i] = label[r]; }
}

struct SolveLog {
    int64_t iters = 0, cell_iters = 0, n = 0;
    int32_t *buf = nullptr; int64_t cap = 0;
    void add(int steps, unsigned cells) {
        int it = steps - 1;
        if (buf && n < cap) buf[n] = it;
        ++n; iters += it; cell_iters += (int64_t)it * cells;
    }
};

vec broadcast(const double *p, int np, unsigned n) {
    if (!p || np <= 0) return vec(n, 0.0);
    if ((unsigned)np == n) return vec(p, p + n);
    return vec(n, p[0]);
}

double now_ms() {
    return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

/* step_signal (fast_binless.ipp:2-48). All FusedLassoGaussianEstimator copies in
 * `flos` share ONE GFLLibrary strategy object (std::vector fill-constructor copies
 * the shared_ptr, GFLLibrary.hpp:59 / fast_binless.cpp:117-118), so a single
 * solver carries beta/z/u/alpha from one dataset to the next. */
void step_signal(const Data &d, const vec &res, const vec &wts, GflSolver &flo, double tol,
                 const vec &lam2, unsigned group, vec &phihat, vec &weights, vec &signal,
                 SolveLog &log, double &ms_solver) {
    phihat.resize(d.N); weights.resize(d.N);
    for (unsigned i = 0; i < d.N; ++i) {
        unsigned cvgroup = 1 + ((d.bin2[i] + d.bin1[i]) % 2);
        if (cvgroup == group) { phihat[i] = -100; weights[i] = 0; }
        else { phihat[i] = res[i] + d.beta[i]; weights[i] = wts[i]; }
    }
    signal.clear(); signal.reserve(d.N);
    for (unsigned ds = 0; ds < d.ndatasets; ++ds) {
        vec y(phihat.begin() + (size_t)ds * d.ncells, phihat.begin() + (size_t)(ds + 1) * d.ncells);
        vec wt(weights.begin() + (size_t)ds * d.ncells, weights.begin() + (size_t)(ds + 1) * d.ncells);
        double t0 = now_ms();
        flo.optimize(y, wt, lam2[ds], tol);
        ms_solver += now_ms() - t0;
        log.add(flo.last_steps, flo.N);
        vec beta = flo.get();
        double avg = 0, wsum = 0;
        for (unsigned i = 0; i < d.ncells; ++i) {
            int no = d.nobs[(size_t)ds * d.ncells + i];
            avg += beta[i] * no; wsum += no;
        }
        avg = avg / wsum;
        for (unsigned i = 0; i < d.ncells; ++i) beta[i] = beta[i] - avg;
        signal.insert(signal.end(), beta.begin(), beta.end());
    }
}

void copy_out(double *dst, const vec &src) { if (dst) std::memcpy(dst, src.data(), src.size() * sizeof(double)); }

/* get_as_dataframe, Signal overload (fast_dataframe.cpp:9-110) */
void signal_dataframe(const Model &m, const vec &lam1, double tol_val, bool compute_patchnos,
                      const char *csv_out, double *background, double *biasmat, double *decaymat,
                      double *signal, double *binless, double *phihat, double *weight, int32_t *patchno) {
    const Data &d = m.d;
    vec e = m.expo_leg.spread(m.expo.get_binned_estimate());
    vec b = m.bias_leg.spread(m.bias.get_binned_estimate());
    vec dc = m.decay_leg.spread(m.dec.get_binned_estimate());
    FILE *csv = nullptr;
    if (csv_out && csv_out[0]) {
        csv = std::fopen(csv_out, "w");
        if (csv) std::fprintf(csv, "\"phihat\",\"weight\",\"biasmat\",\"decaymat\",\"signal\",\"binless\",\"background\"\n"); /* csvfile.hpp:86-99 */
    }
    for (unsigned i = 0; i < d.N; ++i) {
        double bipbj = b[i], ldec = dc[i], l1 = lam1[d.name[i] - 1], lsig = d.beta[i];
        lsig = (lsig > 0) ? std::max(0., lsig - l1) : std::min(0., lsig + l1);
        double v_bias = std::exp(bipbj), v_dec = std::exp(ldec), v_sig = std::exp(lsig);
        double v_binless = std::exp(ldec + lsig), v_bg = d.nobs[i] * std::exp(bipbj + ldec + e[i]);
        if (csv) std::fprintf(csv, "%.4f,%.4f,%.4f,%.4f,%.4f,%.4f,%.4f\n", d.betahat[i], d.weights[i],
                              v_bias, v_dec, v_sig, v_binless, v_bg);
        if (biasmat) biasmat[i] = v_bias;
        if (decaymat) decaymat[i] = v_dec;
        if (signal) signal[i] = v_sig;
        if (binless) binless[i] = v_binless;
        if (background) background[i] = v_bg;
    }
    if (csv) std::fclose(csv);
    copy_out(phihat, d.betahat); copy_out(weight, d.weights);
    if (patchno) {
        if (compute_patchnos)
            for (unsigned ds = 0; ds < d.ndatasets; ++ds)
                patch_numbers(d.nbins, tol_val, d.beta.data() + (size_t)ds * d.ncells, patchno + (size_t)ds * d.ncells);
        else std::fill(patchno, patchno + d.N, -1);
    }
}

void export_state(const Model &m, fb_mean_state *biases, fb_gam_state *decay, fb_mean_state *expos) {
    if (biases) {
        if (biases->estimate) std::memcpy(biases->estimate, m.bias.estimate.data(), m.bias.estimate.size() * sizeof(double));
        biases->n = m.bias.estimate.size(); biases->mean = m.bias.mean;
    }
    if (decay) {
        for (unsigned k = 0; k < FB_DECAY_K; ++k) decay->beta[k] = m.dec.beta[k];
        decay->lambda = m.dec.lambda; decay->mean = m.dec.mean;
    }
    if (expos) {
        if (expos->estimate) std::memcpy(expos->estimate, m.expo.estimate.data(), m.expo.estimate.size() * sizeof(double));
        expos->n = m.expo.estimate.size(); expos->mean = m.expo.mean;
    }
}

/* binless() (fast_binless.cpp:85-184) */
void run_signal(Model &m, const fb_params *p, fb_signal_out *out) {
    Data &d = m.d;
    const double t_begin = now_ms();
    double ms_solver = 0;
    if (p->verbose) std::printf("init\n");
    m.alpha = p->alpha;
    const double tol_val = p->tol_val;
    unsigned bg_steps = p->bg_steps;
    d.beta.assign(d.N, 0); d.betahat.assign(d.N, 0); d.weights.assign(d.N, 0);
    m.build_legs(tol_val, p->free_decay);
    /* initial Poisson fits, in this order (fast_binless.cpp:104-115) */
    vec expected = m.log_expected(d.beta);
    Model::poisson_lsq(m.expo_leg, expected, d);
    m.expo.update_params(m.expo_leg.phihat, m.expo_leg.nobs, false, false);
    expected = m.log_expected(d.beta);
    Model::poisson_lsq(m.decay_leg, expected, d);
    m.dec.update_params(m.decay_leg.phihat, m.decay_leg.weight, m.decay_leg.nobs);
    expected = m.log_expected(d.beta);
    Model::poisson_lsq(m.bias_leg, expected, d);
    m.bias.update_params(m.bias_leg.phihat, m.bias_leg.nobs, true, true);

    double current_tol_val = 1.;
    GflSolver flo(d.nbins, p->maxdiag);
    double flo_tol = current_tol_val / 20.;
    const vec lam2 = broadcast(p->lam2, p->n_lam2, d.ndatasets);
    const vec lam1 = broadcast(p->lam1, p->n_lam1, d.ndatasets);
    SolveLog log; log.buf = out->solve_iters; log.cap = out->solve_iters_cap;
    vec res, wts, old_expected;
    out->nsteps = 0; out->converged = 0;
    for (unsigned step = 1; step <= p->nouter; ++step) {
        if (m.interrupt_cb && m.interrupt_cb(m.interrupt_arg)) throw std::runtime_error("interrupted");
        if (p->verbose) std::printf("step %u", step);
        /* biases */
        m.residuals(d.beta, res, wts);
        Model::update_summary(m.bias_leg, res, wts, m.bias.get_binned_estimate());
        m.bias.update_params(m.bias_leg.phihat, m.bias_leg.nobs, true, true);
        /* decay */
        if (step <= bg_steps) old_expected = m.log_expected(d.beta);
        m.residuals(d.beta, res, wts);
        Model::update_summary(m.decay_leg, res, wts, m.dec.get_binned_estimate());
        m.dec.update_params(m.decay_leg.phihat, m.decay_leg.weight, m.decay_leg.nobs);
        if (step <= bg_steps) expected = m.log_expected(d.beta);
        /* signal */
        if (step > bg_steps) {
            old_expected = m.log_expected(d.beta);
            m.residuals(d.beta, res, wts);
            vec phihat, weights, signal;
            step_signal(d, res, wts, flo, flo_tol, lam2, 0, phihat, weights, signal, log, ms_solver);
            d.beta = signal; d.betahat = phihat; d.weights = weights;
            expected = m.log_expected(d.beta);
        }
        double pabs, prel;
        get_precision(expected, old_expected, pabs, prel);
        if (p->verbose) std::printf(" : reached precision abs = %g rel = %g\n", pabs, prel);
        if (step - 1 < FB_MAX_OUTER) { out->prec_abs[step - 1] = pabs; out->prec_rel[step - 1] = prel; }
        bool converged = prel <= tol_val && current_tol_val <= tol_val * 1.01;
        if (converged && step <= bg_steps) { converged = false; bg_steps = step; }
        current_tol_val = std::min(current_tol_val, std::max(tol_val, pabs));
        flo_tol = current_tol_val / 20;
        if (converged || step == p->nouter) d.beta = shift_signal(d);
        else d.beta = remove_signal_degeneracy(d);
        /* exposures */
        m.residuals(d.beta, res, wts);
        Model::update_summary(m.expo_leg, res, wts, m.expo.get_binned_estimate());
        m.expo.update_params(m.expo_leg.phihat, m.expo_leg.nobs, false, false);
        out->nsteps = step;
        if (converged) { if (p->verbose) std::printf("converged\n"); out->converged = 1; break; }
    }
    if (p->verbose) std::printf("done\n");
    signal_dataframe(m, lam1, tol_val, p->compute_patchnos != 0, p->csv_out, out->background, out->biasmat,
                     out->decaymat, out->signal, out->binless, out->phihat, out->weight, out->patchno);
    copy_out(out->log_signal, d.beta);
    export_state(m, &out->biases, &out->decay, &out->exposures);
    out->admm_iters = log.iters; out->admm_cell_iters = log.cell_iters; out->n_solves = log.n;
    out->ms_solver = ms_solver; out->ms_total = now_ms() - t_begin;
    m.has_state = true;
}

/* restore: set_state calls of fast_binless.cpp:244-260 */
void restore(Model &m, const double *log_signal, const fb_mean_state *biases, const fb_gam_state *decay,
             const fb_mean_state *expos, double tol_val) {
    Data &d = m.d;
    if (!m.legs_built) m.build_legs(tol_val, 10000);
    if (log_signal) d.beta.assign(log_signal, log_signal + d.N);
    if (biases) { m.bias.estimate.assign(biases->estimate, biases->estimate + biases->n); m.bias.mean = biases->mean; }
    if (decay) { m.dec.beta.assign(decay->beta, decay->beta + FB_DECAY_K); m.dec.lambda = decay->lambda; m.dec.mean = decay->mean; }
    if (expos) { m.expo.estimate.assign(expos->estimate, expos->estimate + expos->n); m.expo.mean = expos->mean; }
    if (m.bias.estimate.size() != m.bias_leg.nb || m.expo.estimate.size() != m.expo_leg.nb)
        throw ArgError("restored leg state has the wrong size");
    m.has_state = true;
}

/* binless_difference (fast_binless.cpp:235-279) + step_difference (fast_binless.ipp:51-105)
 * + Difference dataframe (fast_dataframe.cpp:113-145) */
void run_difference(Model &m, int ref, const fb_params *p, fb_diff_out *out) {
    Data &d = m.d;
    const double t_begin = now_ms();
    double ms_solver = 0;
    if (!m.has_state) throw std::logic_error("run_difference needs a fitted or restored model");
    if (p->verbose) std::printf("init\n");
    m.alpha = p->alpha;
    if ((int)d.ndatasets < ref || ref < 1) throw ArgError("Ref is not in the allowed range");
    const vec log_signal = d.beta; /* out$log_signal */
    Data dd = d; /* work on a Difference view, leave the signal model untouched */
    dd.ref0 = ref - 1; dd.phi_ref.assign(d.N, 0);
    diff_set_log_signal(dd, log_signal);
    const double converge = p->tol_val / 20.;
    GflSolver flo(d.nbins, p->maxdiag);
    const vec lam2 = broadcast(p->lam2, p->n_lam2, d.ndatasets - 1);
    if (p->verbose) std::printf("compute\n");
    vec res, wts;
    m.residuals(diff_get_log_signal(dd), res, wts);
    SolveLog log; log.buf = out->solve_iters; log.cap = out->solve_iters_cap;
    /* step_difference */
    vec dsignal = diff_get_log_signal(dd);
    vec phihat(d.N), deltahat(d.N), weights(d.N);
    for (unsigned i = 0; i < d.N; ++i) {
        phihat[i] = res[i] + dsignal[i];
        deltahat[i] = phihat[i] - dd.phi_ref[i];
        weights[i] = wts[i];
    }
    vec delta; delta.reserve(d.N);
    bool ref_seen = false;
    for (unsigned ds = 0; ds < d.ndatasets; ++ds) {
        if ((int)ds == ref - 1) { delta.resize(delta.size() + d.ncells, 0); ref_seen = true; continue; }
        vec y(deltahat.begin() + (size_t)ds * d.ncells, deltahat.begin() + (size_t)(ds + 1) * d.ncells);
        vec wt(weights.begin() + (size_t)ds * d.ncells, weights.begin() + (size_t)(ds + 1) * d.ncells);
        unsigned idx = ds - (ref_seen ? 1 : 0);
        double t0 = now_ms();
        flo.optimize(y, wt, lam2[idx], converge);
        ms_solver += now_ms() - t0;
        log.add(flo.last_steps, flo.N);
        vec ld = flo.get();
        delta.insert(delta.end(), ld.begin(), ld.end());
    }
    vec phi_ref(d.N);
    const size_t ro = (size_t)(ref - 1) * d.ncells;
    for (unsigned i = 0; i < d.N; ++i) {
        double pr = phihat[ro + i % d.ncells], wr = weights[ro + i % d.ncells], val;
        if (wr == 0 && weights[i] == 0) val = (pr + phihat[i]) / 2;
        else val = (pr * wr + (phihat[i] - delta[i]) * weights[i]) / (wr + weights[i]);
        phi_ref[i] = std::max(val, 0.);
    }
    if (out->difference) for (unsigned i = 0; i < d.N; ++i) out->difference[i] = std::exp(delta[i]);
    copy_out(out->deltahat, deltahat); copy_out(out->weight, weights); copy_out(out->phi_ref, phi_ref);
    if (out->patchno) {
        if (p->compute_patchnos)
            for (unsigned ds = 0; ds < d.ndatasets; ++ds)
                patch_numbers(d.nbins, p->tol_val, delta.data() + (size_t)ds * d.ncells, out->patchno + (size_t)ds * d.ncells);
        else std::fill(out->patchno, out->patchno + d.N, -1);
    }
    if (p->verbose) std::printf("done\n");
    out->admm_iters = log.iters; out->admm_cell_iters = log.cell_iters; out->n_solves = log.n;
    out->ms_solver = ms_solver; out->ms_total = now_ms() - t_begin;
}

/* binless_eval_cv (fast_binless.cpp:186-233) */
void eval_cv(Model &m, const fb_params *p, int nlam, const double *lam2v, unsigned group, fb_cv_out *outs) {
    Data &d = m.d;
    if (!m.has_state) throw std::logic_error("eval_cv needs a fitted or restored model");
    if (d.ndatasets != 1) throw ArgError("Provide only 1 dataset!");
    m.alpha = p->alpha;
    const vec signal_ori = d.beta;
    const vec lam1 = broadcast(p->lam1, p->n_lam1, 1);
    GflSolver flo(d.nbins, p->maxdiag);
    SolveLog log; double ms = 0;
    for (int i = 0; i < nlam; ++i) {
        d.beta = signal_ori;
        vec res, wts;
        m.residuals(d.beta, res, wts);
        vec phihat, weights, signal;
        step_signal(d, res, wts, flo, p->tol_val / 20., vec(1, lam2v[i]), group, phihat, weights, signal, log, ms);
        d.beta = signal; d.betahat = phihat; d.weights = weights;
        fb_cv_out &o = outs[i];
        signal_dataframe(m, lam1, p->tol_val, true, nullptr, nullptr, nullptr, nullptr, o.signal, o.binless,
                         o.phihat, o.weight, o.patchno);
        copy_out(o.log_signal, d.beta);
    }
    d.beta = signal_ori;
}

/* lambda2 / BIC sweep (fastbinless.h: sweep_lambda): cold-start solve per (lambda2, dataset); chi square of
 * Likelihoods.hpp:25-42 with eCprime = 0, dof of DOFComputer.hpp:16-24, BIC of ScoreAssembler.hpp:41-55 */
void sweep_lambda(Model &m, const fb_params *p, int nlam, const double *lam2v, fb_sweep_out *out) {
    Data &d = m.d;
    if (!m.has_state) throw std::logic_error("sweep_lambda needs a fitted or restored model");
    m.alpha = p->alpha;
    const vec lam1 = broadcast(p->lam1, p->n_lam1, d.ndatasets);
    vec res, wts;
    m.residuals(d.beta, res, wts);
    vec phihat(d.N);
    for (unsigned i = 0; i < d.N; ++i) phihat[i] = res[i] + d.beta[i];
    double ms = 0; int64_t cell_iters = 0;
    std::vector<int32_t> patch(d.ncells);
    for (int k = 0; k < nlam; ++k)
        for (unsigned ds = 0; ds < d.ndatasets; ++ds) {
            const size_t off = (size_t)ds * d.ncells;
            GflSolver flo(d.nbins, p->maxdiag);
            vec y(phihat.begin() + off, phihat.begin() + off + d.ncells), wt(wts.begin() + off, wts.begin() + off + d.ncells);
            double t0 = now_ms();
            flo.optimize(y, wt, lam2v[k], p->tol_val / 20.);
            ms += now_ms() - t0;
            cell_iters += (int64_t)(flo.last_steps - 1) * flo.N;
            vec beta = flo.get();
            double avg = 0, wsum = 0;
            for (unsigned i = 0; i < d.ncells; ++i) { int no = d.nobs[off + i]; avg += beta[i] * no; wsum += no; }
            avg = avg / wsum;
            for (unsigned i = 0; i < d.ncells; ++i) beta[i] = beta[i] - avg;
            patch_numbers(d.nbins, p->tol_val, beta.data(), patch.data());
            double chisq = 0;
            std::vector<char> sel(d.ncells, 0);
            for (unsigned i = 0; i < d.ncells; ++i) {
                const double b = beta[i], l1 = lam1[ds];
                const double soft = (b > 0) ? std::max(0., b - l1) : std::min(0., b + l1);
                chisq += wt[i] * (y[i] - soft) * (y[i] - soft) + l1 * l1;
                if (std::fabs(soft) > p->tol_val / 2) sel[patch[i]] = 1;
            }
            int dof = 0;
            for (char c : sel) dof += c;
            const size_t o = (size_t)k * d.ndatasets + ds;
            if (out->chisq) out->chisq[o] = chisq;
            if (out->dof) out->dof[o] = dof;
            if (out->steps) out->steps[o] = flo.last_steps;
            if (out->bic) out->bic[o] = chisq + std::log(wsum) * dof;
        }
    out->ms_solver = ms; out->admm_cell_iters = cell_iters;
}

/* checkerboard CV score (fastbinless.h: cv_score): ScorePreparator<CVkSD>::compute (ScorePreparator.ipp:20-49) with the
 * fast path's hold-out convention (fast_binless.ipp:14-22) and centring (:33-44), chi square of Likelihoods.hpp:25-42
 * (eCprime = 0), assembly of ScoreAssembler.hpp:24-36 */
void cv_score(Model &m, const fb_params *p, int nlam, const double *lam2v, fb_cv_score_out *out) {
    Data &d = m.d;
    if (!m.has_state) throw std::logic_error("cv_score needs a fitted or restored model");
    m.alpha = p->alpha;
    const vec lam1 = broadcast(p->lam1, p->n_lam1, d.ndatasets);
    vec res, wts;
    m.residuals(d.beta, res, wts);
    vec phihat(d.N);
    for (unsigned i = 0; i < d.N; ++i) phihat[i] = res[i] + d.beta[i];
    double ms = 0; int64_t cell_iters = 0;
    const int ngroups = 2;
    for (int k = 0; k < nlam; ++k)
        for (unsigned ds = 0; ds < d.ndatasets; ++ds) {
            const size_t off = (size_t)ds * d.ncells;
            vec y(phihat.begin() + off, phihat.begin() + off + d.ncells), w(wts.begin() + off, wts.begin() + off + d.ncells);
            ivec cvgroup(d.ncells);
            for (unsigned i = 0; i < d.ncells; ++i) cvgroup[i] = (d.bin2[off + i] + d.bin1[off + i]) % ngroups;
            vec beta_cv(d.ncells, 0.0);
            const size_t o = (size_t)k * d.ndatasets + ds;
            for (int g = 0; g < ngroups; ++g) {
                vec y_cv(y), w_cv(w);
                for (unsigned i = 0; i < d.ncells; ++i) if (cvgroup[i] == g) { y_cv[i] = -100; w_cv[i] = 0; }
                GflSolver flo(d.nbins, p->maxdiag);   /* cold start */
                double t0 = now_ms();
                flo.optimize(y_cv, w_cv, lam2v[k], p->tol_val / 20.);
                ms += now_ms() - t0;
                cell_iters += (int64_t)(flo.last_steps - 1) * flo.N;
                if (out->steps) out->steps[2 * o + g] = flo.last_steps;
                vec values = flo.get();
                double avg = 0, wsum = 0;
                for (unsigned i = 0; i < d.ncells; ++i) { int no = d.nobs[off + i]; avg += values[i] * no; wsum += no; }
                avg = avg / wsum;
                for (unsigned i = 0; i < d.ncells; ++i) if (cvgroup[i] == g) beta_cv[i] = values[i] - avg;
            }
            vec chisq(d.ncells);
            const double l1 = lam1[ds];
            for (unsigned i = 0; i < d.ncells; ++i) {
                const double b = beta_cv[i];
                const double soft = (b > 0) ? std::max(0., b - l1) : std::min(0., b + l1);
                chisq[i] = w[i] * (y[i] - soft) * (y[i] - soft) + l1 * l1;
            }
            vec gw, gcv;
            for (int g = 0; g < ngroups; ++g) {
                double Ni = 0, sum = 0;
                for (unsigned i = 0; i < d.ncells; ++i) if (cvgroup[i] == g) { Ni += 1; sum += chisq[i]; }
                gw.push_back(Ni); gcv.push_back(sum / Ni);
            }
            const double Nt = gw[0] + gw[1];
            const double CV = (gw[0] * gcv[0] + gw[1] * gcv[1]) / Nt;
            const double CV_sd = std::sqrt((gw[0] * (gcv[0] - CV) * (gcv[0] - CV) + gw[1] * (gcv[1] - CV) * (gcv[1] - CV)) / (Nt - 1));
            if (out->cv) out->cv[o] = CV;
            if (out->cv_sd) out->cv_sd[o] = CV_sd;
        }
    out->ms_solver = ms; out->admm_cell_iters = cell_iters;
}

/* chromosome_binless post-processing and matrix triplets (R/chromosome.R:262-330, signal case), statement by statement on
 * the data frame columns of fast_binless (fast_dataframe.cpp:62-88) */
void band_export(Model &m, const fb_band_params *p, fb_band_out *out) {
    Data &d = m.d;
    if (!m.has_state) throw std::logic_error("band_export needs a fitted or restored model");
    const vec lam1 = broadcast(p->lam1, p->n_lam1, d.ndatasets);
    const vec dc = m.decay_leg.spread(m.dec.get_binned_estimate());
    vec signal(d.N), binless(d.N);
    for (unsigned i = 0; i < d.N; ++i) {
        double l1 = lam1[d.name[i] - 1], lsig = d.beta[i];
        lsig = (lsig > 0) ? std::max(0., lsig - l1) : std::min(0., lsig + l1);
        signal[i] = std::exp(lsig); binless[i] = std::exp(dc[i] + lsig);
    }
    vec lam1_sig(d.ndatasets);
    for (unsigned ds = 0; ds < d.ndatasets; ++ds) {     /* mat[distance > far.cutoff, median(log(signal)), by = name] */
        vec v;
        for (unsigned c = 0; c < d.ncells; ++c) { const size_t i = (size_t)ds * d.ncells + c; if (d.distance[i] > p->far_cutoff) v.push_back(std::log(signal[i])); }
        if (v.empty()) { lam1_sig[ds] = p->lam1_final; continue; }
        std::sort(v.begin(), v.end());
        const size_t nn = v.size();
        lam1_sig[ds] = (nn & 1) ? v[nn / 2] : (v[nn / 2 - 1] + v[nn / 2]) / 2;
        if (std::isnan(lam1_sig[ds])) lam1_sig[ds] = p->lam1_final;
    }
    if (out->lam1_sig) std::copy(lam1_sig.begin(), lam1_sig.end(), out->lam1_sig);
    for (unsigned i = 0; i < d.N; ++i) signal[i] = std::max(signal[i] / std::exp(lam1_sig[d.name[i] - 1]), 1.);   /* :277 */
    int64_t nband = 0;
    for (unsigned c = 0; c < d.ncells; ++c) if ((unsigned)(d.bin2[c] - d.bin1[c]) <= p->maxdiag) ++nband;     /* :303 */
    out->nband = nband;
    for (unsigned ds = 0; ds < d.ndatasets; ++ds) {
        int64_t k = 0;
        for (unsigned c = 0; c < d.ncells; ++c) {
            const size_t i = (size_t)ds * d.ncells + c;
            if ((unsigned)(d.bin2[i] - d.bin1[i]) > p->maxdiag) continue;
            if (ds == 0) { if (out->band_bin1) out->band_bin1[k] = d.bin1[i]; if (out->band_bin2) out->band_bin2[k] = d.bin2[i]; }
            if (out->band_observed) out->band_observed[(size_t)ds * nband + k] = d.observed[i];
            if (out->band_binless) out->band_binless[(size_t)ds * nband + k] = binless[i];
            ++k;
        }
    }
    int64_t total = 0;
    for (unsigned ds = 0; ds < d.ndatasets; ++ds) {       /* mat[signal != 1] -> log10(signal) (:318-321) */
        int64_t cnt = 0;
        for (unsigned c = 0; c < d.ncells; ++c) {
            const size_t i = (size_t)ds * d.ncells + c;
            if (signal[i] == 1.) continue;
            if (out->sig_bin1 && out->sig_bin2 && out->sig_log10) {
                if (total + cnt >= out->sig_cap) throw ArgError("sig_cap is too small");
                out->sig_bin1[total + cnt] = d.bin1[i]; out->sig_bin2[total + cnt] = d.bin2[i]; out->sig_log10[total + cnt] = std::log10(signal[i]);
            }
            ++cnt;
        }
        if (out->nsig) out->nsig[ds] = cnt;
        total += cnt;
    }
}

} // namespace

#include "perf_iteration_oracle.hpp"

struct fb_handle { Model m; };

#define ORACLE_TRY(h, body)                                                     \
    try { body; return FB_OK; }                                                 \
    catch (const ArgError &e) { (h)->m.err = e.what(); return FB_ERR_ARG; }     \
    catch (const NumError &e) { (h)->m.err = e.what(); return FB_ERR_NUMERIC; } \
    catch (const std::logic_error &e) { (h)->m.err = e.what(); return FB_ERR_STATE; } \
    catch (const std::exception &e) { (h)->m.err = e.what(); return FB_ERR_INTERRUPT; }

extern "C" {

FB_API const char *fbo_gfl_kind(void) { return kGflKind; }

FB_API int fbo_create(int nbins, const fb_mat *mat, fb_handle **out) {
    if (!out) return FB_ERR_ARG;
    *out = nullptr;
    std::unique_ptr<fb_handle> h(new fb_handle);
    try { build_data(h->m.d, nbins, mat); }
    catch (const std::exception &e) { g_create_error = e.what(); return FB_ERR_ARG; }
    *out = h.release();
    return FB_OK;
}
/* bin_data() for a CSnorm object (R/preprocess.R:114-146) + add_bin_bounds_and_distance (R/common.R:115-132),
 * restated with plain loops. */
FB_API int fbo_create_binned(const fb_reads *r, int *nbins_out, fb_handle **out) {
    if (!out || !r || !nbins_out) return FB_ERR_ARG;
    *out = nullptr;
    try {
        if (r->ndatasets < 1 || r->resolution < 1 || r->ncuts < 1) throw ArgError("empty input");
        int mn = r->cut_pos[0], mx = r->cut_pos[0];
        for (int64_t k = 0; k < r->ncuts; ++k) { mn = std::min(mn, r->cut_pos[k]); mx = std::max(mx, r->cut_pos[k]); }
        const int64_t res = r->resolution, b0 = (int64_t)mn - 1;
        /* seq(min(pos)-1, max(pos)+1, resolution): borders b0 + k res <= mx + 1; one bin between consecutive borders */
        const int64_t nb = ((int64_t)mx + 1 - b0) / res;
        if (nb < 1) throw ArgError("no complete bin at this resolution");
        const int64_t top = b0 + nb * res;
        auto bin_of = [&](int pos) -> int64_t { /* cut(pos, borders, right=F, include.lowest=T): -1 is NA */
            if (pos < b0 || pos > top) return -1;
            if (pos == top) return nb - 1;
            return ((int64_t)pos - b0) / res;
        };
        const int64_t D = r->ndatasets, nc = nb * (nb + 1) / 2, N = D * nc;
        std::vector<int64_t> cuts(D * nb, 0);
        for (int64_t k = 0; k < r->ncuts; ++k) {
            const int64_t b = bin_of(r->cut_pos[k]);
            if (b >= 0) cuts[(r->cut_name[k] - 1) * nb + b] += 1;
        }
        ivec name(N), bin1(N), bin2(N), pos1(N), pos2(N), observed(N, 0), nobs(N, 0);
        std::vector<char> present(N, 0);   /* the cell has a row in poscounts (its counts may sum to zero) */
        vec distance(N);
        auto cell = [&](int64_t d, int64_t i, int64_t j) { return d * nc + i * nb - i * (i - 1) / 2 + (j - i); };
        for (int64_t k = 0; k < r->nreads; ++k) {
            const int64_t i = bin_of(r->rd_pos1[k]), j = bin_of(r->rd_pos2[k]);
            if (i < 0 || j < 0 || i > j) continue;               /* NA bins and pairs below the diagonal match no cell */
            const int64_t g = cell(r->rd_name[k] - 1, i, j);
            observed[g] += r->contact_close[k] + r->contact_far[k] + r->contact_up[k] + r->contact_down[k];
            present[g] = 1;
        }
        for (int64_t d = 0; d < D; ++d)
            for (int64_t i = 0; i < nb; ++i)
                for (int64_t j = i; j < nb; ++j) {
                    const int64_t g = cell(d, i, j);
                    name[g] = (int)d + 1; bin1[g] = (int)i + 1; bin2[g] = (int)j + 1;
                    const int64_t beg1 = b0 + i * res, beg2 = b0 + j * res;
                    pos1[g] = (int)((beg1 + beg1 + res) / 2); pos2[g] = (int)((beg2 + beg2 + res) / 2);
                    distance[g] = (double)(beg2 - beg1) + (double)res / 2.0;
                    int64_t no = 4 * cuts[d * nb + i] * cuts[d * nb + j];
                    if (i == j) no /= 2;
                    nobs[g] = !present[g] ? 0 : (int)std::min<int64_t>(no, INT32_MAX);   /* is.na(observed) -> both 0 */
                }
        for (int64_t d = 0; d < D; ++d) observed[cell(d, 0, nb - 1)] += 1;
        fb_mat m{N, name.data(), bin1.data(), pos1.data(), bin2.data(), pos2.data(), observed.data(), nobs.data(), distance.data()};
        std::unique_ptr<fb_handle> h(new fb_handle);
        build_data(h->m.d, (int)nb, &m);
        *nbins_out = (int)nb;
        *out = h.release();
        return FB_OK;
    } catch (const std::exception &e) { g_create_error = e.what(); return FB_ERR_ARG; }
}
FB_API int64_t fbo_mat_rows(const fb_handle *h) { return h ? (int64_t)h->m.d.N : 0; }
FB_API int fbo_get_mat(fb_handle *h, int32_t *name, int32_t *bin1, int32_t *pos1, int32_t *bin2, int32_t *pos2,
                       double *distance, int32_t *observed, int32_t *nobs) {
    if (!h) return FB_ERR_ARG;
    const Data &d = h->m.d;
    auto cp = [&](int32_t *dst, const ivec &src) { if (dst) std::copy(src.begin(), src.end(), dst); };
    cp(name, d.name); cp(bin1, d.bin1); cp(pos1, d.pos1); cp(bin2, d.bin2); cp(pos2, d.pos2); cp(observed, d.observed); cp(nobs, d.nobs);
    if (distance) std::copy(d.distance.begin(), d.distance.end(), distance);
    return FB_OK;
}
FB_API void fbo_destroy(fb_handle *h) { delete h; }
FB_API const char *fbo_last_error(const fb_handle *h) { return h ? h->m.err.c_str() : g_create_error.c_str(); }
FB_API int fbo_run_signal(fb_handle *h, const fb_params *p, fb_signal_out *out) { ORACLE_TRY(h, run_signal(h->m, p, out)) }
FB_API int fbo_restore(fb_handle *h, const double *log_signal, const fb_mean_state *biases,
                       const fb_gam_state *decay, const fb_mean_state *exposures) {
    ORACLE_TRY(h, restore(h->m, log_signal, biases, decay, exposures, 0.2))
}
FB_API int fbo_run_difference(fb_handle *h, int ref, const fb_params *p, fb_diff_out *out) { ORACLE_TRY(h, run_difference(h->m, ref, p, out)) }
FB_API int fbo_eval_cv(fb_handle *h, const fb_params *p, int nlam, const double *lam2, unsigned group, fb_cv_out *outs) {
    ORACLE_TRY(h, eval_cv(h->m, p, nlam, lam2, group, outs))
}
FB_API int fbo_sweep_lambda(fb_handle *h, const fb_params *p, int nlam, const double *lam2, fb_sweep_out *out) {
    ORACLE_TRY(h, sweep_lambda(h->m, p, nlam, lam2, out))
}
FB_API int fbo_band_export(fb_handle *h, const fb_band_params *p, fb_band_out *out) { ORACLE_TRY(h, band_export(h->m, p, out)) }
FB_API int fbo_cv_score(fb_handle *h, const fb_params *p, int nlam, const double *lam2, fb_cv_score_out *out) {
    ORACLE_TRY(h, cv_score(h->m, p, nlam, lam2, out))
}
/* csvfile.hpp:30-117 row by row, plain text (the oracle does not link zlib: tests gunzip the product's file and compare) */
FB_API int fbo_csv_write(const char *path, const char *header, int64_t nrows, int ncols, const double *const *cols, int) {
    FILE *f = std::fopen(path, "w");
    if (!f) return 1;
    std::fprintf(f, "%s\n", header);
    for (int64_t i = 0; i < nrows; ++i)
        for (int c = 0; c < ncols; ++c) std::fprintf(f, c + 1 == ncols ? "%.4f\n" : "%.4f,", cols[c][i]);
    std::fclose(f);
    return FB_OK;
}
FB_API int fbo_wgfl_signal_bic(int nbins, const fb_cts *cts, const fb_perf_meta *meta, const fb_perf_params *p,
                               const double *beta_init, int warm, fb_gfl_state *gfl, fb_perf_signal_out *out) {
    try { perf_wgfl_signal(nbins, cts, meta, p, beta_init, warm, gfl, out); return FB_OK; }
    catch (const std::exception &e) { g_create_error = e.what(); return FB_ERR_ARG; }
}
FB_API int fbo_wgfl_diff_bic(int nbins, const fb_cts *cts, const fb_cts *ref, const fb_perf_meta *meta, const fb_perf_params *p,
                             const double *phi_ref_init, const double *beta_init, int warm, fb_gfl_state *gfl, fb_perf_diff_out *out) {
    try { perf_wgfl_diff(nbins, cts, ref, meta, p, phi_ref_init, beta_init, warm, gfl, out); return FB_OK; }
    catch (const std::exception &e) { g_create_error = e.what(); return FB_ERR_ARG; }
}
/* test hook: the scatter front end alone (cts_core.c:5-47: the reference's object in libfbo_ref.so, the restatement in the port) */
FB_API int fbo_test_cts_core(int nbins, double dispersion, const fb_cts *c, const double *phi, double *phihat, double *phihat_var,
                             double *phihat_var_nodecay, double *ncounts) {
    PerfCts cts = perf_copy_cts(c);
    const int nb = nbins * (nbins + 1) / 2;
    vec ph(phi, phi + nb);
    ivec b1(nb), b2(nb);
    std::fill(phihat, phihat + nb, 0.0); std::fill(phihat_var, phihat_var + nb, 0.0);
    std::fill(phihat_var_nodecay, phihat_var_nodecay + nb, 0.0); std::fill(ncounts, ncounts + nb, 0.0);
    ORACLE_CTS_CORE((int)cts.bin1.size(), cts.bin1.data(), cts.bin2.data(), cts.count.data(), cts.lmu_nosig.data(), cts.nobs.data(),
                    cts.log_decay.data(), nbins, dispersion, ph.data(), phihat, phihat_var, phihat_var_nodecay, ncounts, b1.data(), b2.data());
    return FB_OK;
}
FB_API int64_t fbo_kernel_launches(const fb_handle *) { return 0; }
FB_API void fbo_release_cached_memory(void) {}

/* The multi-GPU entry points of the ABI. The reference is one thread in one process (gfl_graph_fl.c:53-213), so the
 * oracle is always a single rank; fbo_shard_plan restates the equal-cell split by counting cells line by line (the
 * product uses the closed-form offsets) so that the host-side tests can compare the two. */
FB_API int fbo_comm_unique_id(void *) { g_create_error = "the CPU oracle is a single rank"; return 6; }
FB_API int fbo_comm_init(int rank, int world, const void *) {
    if (rank == 0 && world == 1) return FB_OK;
    g_create_error = "the CPU oracle is a single rank"; return 6;
}
FB_API int fbo_comm_destroy(void) { return FB_OK; }
FB_API double fbo_gfl_last_solve_ms(void) { return 0.0; }
/* fb_far_median_noalloc has no counterpart in the reference (fastbinless.h); the oracle side of the ABI is the same
 * median on HOST pointers, so that tests can check the device result: R's median of log(exp(soft(log_signal, lam1))) over
 * the cells with bin2 - bin1 >= far_from_diag (chromosome.R:270-276), NaN for none. */
FB_API int fbo_far_median_noalloc(int nbins, const double *log_signal, double lam1, int far_from_diag, double *median, void *) {
    if (!log_signal || !median || nbins < 1 || far_from_diag < 0) { g_create_error = "invalid argument"; return FB_ERR_ARG; }
    std::vector<double> v;
    int64_t pos = 0;
    for (int i = 0; i < nbins; ++i)
        for (int j = i; j < nbins; ++j, ++pos)
            if (j - i >= far_from_diag) {
                const double x = log_signal[pos];
                const double s = x > 0 ? std::max(0., x - lam1) : std::min(0., x + lam1);
                v.push_back(std::log(std::exp(s)));
            }
    if (v.empty()) { *median = NAN; return FB_OK; }
    std::sort(v.begin(), v.end());
    const size_t n = v.size();
    *median = (n & 1) ? v[(n - 1) / 2] : (v[(n - 1) / 2] + v[n / 2]) / 2;
    return FB_OK;
}
/* single-process multi-GPU entry points: the reference is one thread on one CPU, so "ngpus" changes nothing here */
FB_API int fbo_create(int nbins, const fb_mat *mat, fb_handle **out);
FB_API int fbo_create_multi(int nbins, const fb_mat *mat, int, fb_handle **out) { return fbo_create(nbins, mat, out); }
FB_API int fbo_handle_ngpus(const fb_handle *h) { return h ? 1 : 0; }
FB_API int fbo_gfl_solve(int nbins, unsigned maxdiag, const double *y, const double *w, double lam2,
                         double converge, int maxsteps, int warm, fb_gfl_state *st, double *beta_full);
FB_API int fbo_gfl_solve_multi(int, int nbins, unsigned maxdiag, const double *y, const double *w, double lam2,
                               double converge, int maxsteps, int warm, fb_gfl_state *st, double *beta_full) {
    return fbo_gfl_solve(nbins, maxdiag, y, w, lam2, converge, maxsteps, warm, st, beta_full);
}
FB_API int fbo_comm_rank(void) { return 0; }
FB_API int fbo_comm_world(void) { return 1; }
FB_API int fbo_shard_plan(int nbins, unsigned maxdiag, int world, int32_t *row_begin, int32_t *col_begin) {
    if (nbins < 1 || world < 1 || !row_begin || !col_begin) return FB_ERR_ARG;
    unsigned md = maxdiag < (unsigned)nbins ? maxdiag : (unsigned)nbins;
    if (md > (unsigned)(nbins + 1) / 2 || md <= 1) md = nbins;
    long long total = 0;
    for (int i = 0; i < nbins; ++i) total += std::min<long long>(md, nbins - i);
    for (int pass = 0; pass < 2; ++pass) {
        int32_t *out = pass ? col_begin : row_begin;
        long long before = 0; int line = 0;
        for (int k = 0; k < world; ++k) {
            const long long target = (total * k + world - 1) / world;
            while (line < nbins && before < target) {
                before += pass ? std::min<long long>(md, line + 1) : std::min<long long>(md, nbins - line);
                ++line;
            }
            out[k] = line;
        }
        out[world] = nbins; out[0] = 0;
    }
    return FB_OK;
}
FB_API int fbo_set_interrupt_cb(fb_handle *h, int (*cb)(void *), void *arg) { h->m.interrupt_cb = cb; h->m.interrupt_arg = arg; return FB_OK; }

FB_API int fbo_gfl_solve(int nbins, unsigned maxdiag, const double *y, const double *w, double lam2,
                         double converge, int maxsteps, int warm, fb_gfl_state *st, double *beta_full) {
    try {
        GflSolver s(nbins, maxdiag);
        const size_t nc = (size_t)nbins * (nbins + 1) / 2;
        s.ninner_max = maxsteps;
        if (warm) {
            s.alpha = st->alpha;
            if (st->beta) s.beta_impl.assign(st->beta, st->beta + s.N);
            if (st->z) s.z.assign(st->z, st->z + s.tsz);
            if (st->u) s.u.assign(st->u, st->u + s.tsz);
        }
        s.optimize(vec(y, y + nc), vec(w, w + nc), lam2, converge);
        st->alpha = s.alpha; st->steps = s.last_steps;
        if (st->beta) std::memcpy(st->beta, s.beta_impl.data(), s.N * sizeof(double));
        if (st->z) std::memcpy(st->z, s.z.data(), s.tsz * sizeof(double));
        if (st->u) std::memcpy(st->u, s.u.data(), s.tsz * sizeof(double));
        if (beta_full) std::memcpy(beta_full, s.beta_full.data(), nc * sizeof(double));
        return FB_OK;
    } catch (const std::exception &e) { g_create_error = e.what(); return FB_ERR_ARG; }
}

/* exposed for unit tests of the restated third-party pieces */
FB_API int fbo_test_qp(int n, int m, const double *G, const double *g0, const double *CI, double *x) {
    try {
        DMat A(n, n), L, C(n, m);
        std::memcpy(A.a.data(), G, sizeof(double) * n * n);
        std::memcpy(C.a.data(), CI, sizeof(double) * n * m);
        if (!cholesky(A, L)) return FB_ERR_NUMERIC;
        vec g(g0, g0 + n), xx(n);
        for (int i = 0; i < n; ++i) xx[i] = g[i];
        solve_lower(L, xx); solve_lower_T(L, xx);
        for (auto &v : xx) v = -v;
        if (m > 0) gi_solve(L, g, C, xx);
        std::memcpy(x, xx.data(), sizeof(double) * n);
        return FB_OK;
    } catch (const std::exception &e) { g_create_error = e.what(); return FB_ERR_NUMERIC; }
}
FB_API int fbo_test_spline(int n, const double *x, double xmin, double xmax, int K, double *out) {
    DMat X = generate_spline_base(vec(x, x + n), xmin, xmax, K);
    std::memcpy(out, X.a.data(), sizeof(double) * n * K);
    return FB_OK;
}
FB_API int fbo_test_trails(int nbins, unsigned maxdiag, unsigned *trails, unsigned *breakpoints, int *ntrails, unsigned *tsz, unsigned *ncells) {
    try {
        GflSolver s(nbins, maxdiag);
        *ntrails = s.ntrails; *tsz = s.tsz; *ncells = s.N;
        if (trails) std::memcpy(trails, s.trails.data(), sizeof(unsigned) * s.tsz);
        if (breakpoints) std::memcpy(breakpoints, s.breakpoints.data(), sizeof(unsigned) * s.ntrails);
        return FB_OK;
    } catch (const std::exception &e) { g_create_error = e.what(); return FB_ERR_ARG; }
}
FB_API int fbo_test_patchnos(int nbins, double tol, const double *value, int32_t *out) { patch_numbers(nbins, tol, value, out); return FB_OK; }

} // extern "C"
