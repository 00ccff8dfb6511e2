// fb_gam.cpp -- host-side decay GAM (see fb_gam.hpp for the reference map).
#include "fb_gam.hpp"
#include "fb_util.cuh"
#include <algorithm>
#include <cmath>
#include <limits>

namespace fb {

bool chol_lower(int n, const std::vector<double> &A, std::vector<double> &L) {
    L.assign((size_t)n * n, 0.0);
    for (int j = 0; j < n; ++j) {
        double s = A[(size_t)j * n + j];
        for (int k = 0; k < j; ++k) s -= L[(size_t)j * n + k] * L[(size_t)j * n + k];
        if (!(s > 0.0)) return false;
        const double ljj = std::sqrt(s);
        L[(size_t)j * n + j] = ljj;
        for (int i = j + 1; i < n; ++i) {
            double t = A[(size_t)i * n + j];
            for (int k = 0; k < j; ++k) t -= L[(size_t)i * n + k] * L[(size_t)j * n + k];
            L[(size_t)i * n + j] = t / ljj;
        }
    }
    return true;
}

// solves (L L') x = b in place
void chol_solve(int n, const std::vector<double> &L, std::vector<double> &b) {
    for (int i = 0; i < n; ++i) {
        double s = b[i];
        for (int k = 0; k < i; ++k) s -= L[(size_t)i * n + k] * b[k];
        b[i] = s / L[(size_t)i * n + i];
    }
    for (int i = n - 1; i >= 0; --i) {
        double s = b[i];
        for (int k = i + 1; k < n; ++k) s -= L[(size_t)k * n + i] * b[k];
        b[i] = s / L[(size_t)i * n + i];
    }
}

namespace {
// slack of constraint c: x[c] - x[c+1]
inline double slack(const std::vector<double> &x, int c) { return x[c] - x[c + 1]; }
} // namespace

void qp_monotone_tail(int n, const std::vector<double> &L, const std::vector<int> &cons, std::vector<double> &x) {
    const int m = (int)cons.size();
    if (m == 0) return;
    const double eps = std::numeric_limits<double>::epsilon();
    const double inf = std::numeric_limits<double>::infinity();
    // scale of the stopping rule (QuadProg++.cpp:203,211-219,328): c1 = |diag L|^2, c2 = trace(L^-T)
    double c1 = 0.0, c2 = 0.0;
    for (int i = 0; i < n; ++i) { c1 += L[(size_t)i * n + i] * L[(size_t)i * n + i]; c2 += 1.0 / L[(size_t)i * n + i]; }
    std::vector<int> act;         // active constraints, in order of activation
    std::vector<double> u;        // their multipliers
    std::vector<char> is_act(m, 0);
    std::vector<double> ginv_np(n), z(n), r, M, Lm, rhs;
    std::vector<std::vector<double>> ginv_n; // G^-1 n_k for every active k

    for (int iter = 0; iter <= 10 * m; ++iter) { // QuadProg++.cpp:289
        // sum of infeasibilities and most violated constraint, both over the inactive set (:314-318, :407-416)
        double psi = 0.0;
        int ip = -1;
        double worst = 0.0;
        for (int k = 0; k < m; ++k) {
            if (is_act[k]) continue;
            const double s = slack(x, cons[k]);
            psi += std::min(0.0, s);
            if (s < worst) { worst = s; ip = k; }
        }
        if (std::fabs(psi) <= (m - (int)act.size()) * eps * c1 * c2 * 100.0 || ip < 0) return;
        const int cp = cons[ip];
        double uplus = 0.0;
        // G^-1 n_p, n_p = e_cp - e_{cp+1}
        std::fill(ginv_np.begin(), ginv_np.end(), 0.0);
        ginv_np[cp] = 1.0; ginv_np[cp + 1] = -1.0;
        chol_solve(n, L, ginv_np);
        for (;;) {
            const int q = (int)act.size();
            // step directions: z = H n_p (primal), r = N* n_p (dual)
            r.assign(q, 0.0);
            z = ginv_np;
            if (q > 0) {
                M.assign((size_t)q * q, 0.0);
                rhs.assign(q, 0.0);
                for (int a = 0; a < q; ++a) {
                    const int ca = cons[act[a]];
                    rhs[a] = ginv_np[ca] - ginv_np[ca + 1];
                    for (int b = 0; b < q; ++b) M[(size_t)a * q + b] = ginv_n[b][ca] - ginv_n[b][ca + 1];
                }
                if (!chol_lower(q, M, Lm)) throw NumError("quadratic program: dependent constraints");
                r = rhs;
                chol_solve(q, Lm, r);
                for (int b = 0; b < q; ++b)
                    for (int k = 0; k < n; ++k) z[k] -= ginv_n[b][k] * r[b];
            }
            double t1 = inf; int l = -1;
            for (int a = 0; a < q; ++a)
                if (r[a] > 0.0 && u[a] / r[a] < t1) { t1 = u[a] / r[a]; l = a; }
            double zz = 0.0;
            for (int k = 0; k < n; ++k) zz += z[k] * z[k];
            double t2 = inf;
            if (zz > eps) t2 = -slack(x, cp) / (z[cp] - z[cp + 1]);
            const double t = std::min(t1, t2);
            if (!(t < inf)) throw NumError("quadratic program is infeasible");
            auto drop = [&](int a) {
                is_act[act[a]] = 0;
                act.erase(act.begin() + a); u.erase(u.begin() + a); ginv_n.erase(ginv_n.begin() + a);
            };
            if (!(t2 < inf)) { // dual step only
                for (int a = 0; a < q; ++a) u[a] -= t * r[a];
                uplus += t;
                drop(l);
                continue;
            }
            for (int k = 0; k < n; ++k) x[k] += t * z[k];
            for (int a = 0; a < q; ++a) u[a] -= t * r[a];
            uplus += t;
            if (t2 <= t1) { // full step: constraint becomes active
                act.push_back(ip); u.push_back(uplus); ginv_n.push_back(ginv_np); is_act[ip] = 1;
                break;
            }
            drop(l); // partial step
        }
    }
    // iteration cap reached: the reference returns the current iterate
}

void DecayGam::setup(const std::vector<double> &support, double smin, double smax, double tol_val, unsigned free_decay) {
    const int K = kDecayK;
    nb_ = (int)support.size();
    tol_val_ = tol_val;
    // uniform cubic B-spline basis on the knot grid of generate_spline_base (spline.cpp:19-26):
    // dx = 1.01 (max-min)/(K-3), first knot = (min - 0.01 dx) - 3 dx. Four cubic pieces are non-zero at x.
    const double dx = 1.01 * (smax - smin) / (K - 3);
    const double knot0 = (smin - dx * 0.01) - 3 * dx;
    xcol_.assign(nb_, 0);
    xval_.assign((size_t)nb_ * 4, 0.0);
    for (int r = 0; r < nb_; ++r) {
        const double p = (support[r] - knot0) / dx;
        int mi = (int)std::floor(p);
        const double t = p - mi;
        const double t2 = t * t, t3 = t2 * t, omt = 1.0 - t;
        const double v[4] = {omt * omt * omt / 6.0, (3 * t3 - 6 * t2 + 4) / 6.0, (-3 * t3 + 3 * t2 + 3 * t + 1) / 6.0, t3 / 6.0};
        xcol_[r] = mi - 3;
        for (int a = 0; a < 4; ++a) {
            const int col = mi - 3 + a;
            xval_[(size_t)r * 4 + a] = (col >= 0 && col < K) ? v[a] : 0.0; // leftCols(K): spline.cpp:16
        }
    }
    // (K*D2)'(K*D2), D2 rows e_i - 2 e_{i+1} + e_{i+2} (gam.cpp:15-21, QuadProgGAMLibrary.cpp:30-31)
    pen_.assign((size_t)K * K, 0.0);
    const double d2[3] = {1.0, -2.0, 1.0};
    for (int i = 0; i + 2 < K; ++i)
        for (int a = 0; a < 3; ++a)
            for (int b = 0; b < 3; ++b) pen_[(size_t)(i + a) * K + (i + b)] += (K * d2[a]) * (K * d2[b]);
    // decreasing_constraint(K, free_first) (gam.cpp:6-13, fast_decay.hpp:85-87)
    const double ff = (double)K * (std::log((double)free_decay) - smin) / (smax - smin);
    const unsigned free_first = (unsigned)(long long)ff;
    cons_.clear();
    if (K > 1 && free_first < (unsigned)K)
        for (int c = (int)free_first; c <= K - 2; ++c) cons_.push_back(c);
    qp_beta_.assign(K, 1.0); qp_lambda_ = 1.0;
    beta.assign(K, 0.0); lambda = -1.0; mean = 0.0;
}

std::vector<double> DecayGam::estimate() const {
    std::vector<double> e(nb_);
    for (int r = 0; r < nb_; ++r) {
        double s = 0.0;
        for (int a = 0; a < 4; ++a) {
            const int col = xcol_[r] + a;
            if (col >= 0 && col < kDecayK) s += xval_[(size_t)r * 4 + a] * beta[col];
        }
        e[r] = s;
    }
    return e;
}

std::vector<double> DecayGam::binned_estimate() const {
    std::vector<double> e = estimate();
    for (auto &v : e) v -= mean;
    return e;
}

void DecayGam::fit(const std::vector<double> &phihat, const std::vector<double> &weight, const std::vector<double> &nobs) {
    const int K = kDecayK;
    // X' S^-2 X and X' S^-2 y with S^-1 = sqrt(weight) (fast_fitter_gam.hpp:77-79, QuadProgGAMLibrary.cpp:27-29)
    std::vector<double> xtx((size_t)K * K, 0.0), xty(K, 0.0);
    for (int r = 0; r < nb_; ++r) {
        const double s = std::sqrt(weight[r]);
        const double sy = s * phihat[r];
        for (int a = 0; a < 4; ++a) {
            const int ca = xcol_[r] + a;
            if (ca < 0 || ca >= K) continue;
            const double ta = s * xval_[(size_t)r * 4 + a];
            xty[ca] += ta * sy;
            for (int b = 0; b < 4; ++b) {
                const int cb = xcol_[r] + b;
                if (cb < 0 || cb >= K) continue;
                xtx[(size_t)ca * K + cb] += ta * (s * xval_[(size_t)r * 4 + b]);
            }
        }
    }
    std::vector<double> beta_old = qp_beta_, A((size_t)K * K), L;
    int iter = 0;
    bool converged = false;
    do {
        for (int a = 0; a < K; ++a)
            for (int b = 0; b < K; ++b)
                A[(size_t)a * K + b] = xtx[(size_t)a * K + b] + qp_lambda_ * qp_lambda_ * pen_[(size_t)a * K + b] + (a == b ? 1e-10 : 0.0);
        if (!chol_lower(K, A, L)) throw NumError("decomposition failed"); // QuadProgGAMLibrary.cpp:21
        qp_beta_ = xty;
        chol_solve(K, L, qp_beta_);
        if (!cons_.empty()) qp_monotone_tail(K, L, cons_, qp_beta_);
        double sq = 0.0;
        for (int r = 0; r + 2 < K; ++r) {
            const double t = K * (qp_beta_[r] - 2 * qp_beta_[r + 1] + qp_beta_[r + 2]);
            sq += t * t;
        }
        qp_lambda_ = std::sqrt((K - 2) / (sq + 1.0 / (kDecaySigma * kDecaySigma)));
        const double mx = *std::max_element(qp_beta_.begin(), qp_beta_.end());
        const double mn = *std::min_element(qp_beta_.begin(), qp_beta_.end());
        double precision = mx - mn;
        if (precision > 0) {
            double dm = 0.0;
            for (int a = 0; a < K; ++a) dm = std::max(dm, std::fabs(qp_beta_[a] - beta_old[a]));
            precision = dm / precision;
        }
        converged = precision < tol_val_;
        beta_old = qp_beta_;
        ++iter;
    } while (iter < kDecayMaxIter && !converged);
    last_iters = iter;
    lambda = qp_lambda_;
    beta = qp_beta_;
    // center_estimate (fast_fitter_gam.hpp:108-111)
    const std::vector<double> e = estimate();
    double num = 0.0, den = 0.0;
    for (int r = 0; r < nb_; ++r) { num += nobs[r] * e[r]; den += nobs[r]; }
    mean = num / den;
}

} // namespace fb
