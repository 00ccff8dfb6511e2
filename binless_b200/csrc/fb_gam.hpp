// fb_gam.hpp -- host-side decay leg: penalised cubic B-spline GAM with a monotone tail.
//
// The K = 50 problem is a few thousand flops per iteration; it stays on the host
// ("not shardable ... replicated on every rank"). Replaces
//   FitterSettingsImpl<Decay,GAM>      fast_decay.hpp:71-89
//   FitterImpl<Leg,GAM>::update_params fast_fitter_gam.hpp:75-111
//   QuadProgGAMLibrary::optimize       QuadProgGAMLibrary.cpp:24-73
//   generate_spline_base               spline.cpp:4-26
//   second_order_difference_matrix, decreasing_constraint   gam.cpp:6-21
//   QuadProgPP::solve_quadprog_with_guess                   QuadProg++.cpp:165-
// (paths relative to /root/reference/binless/src).
#pragma once
#include <vector>

namespace fb {

constexpr int kDecayK = 50;          // Config<Decay,GAM>::K, fast_decay.hpp:25
constexpr int kDecayBinsPerBf = 100; // bins_per_bf, fast_decay.hpp:28
constexpr int kDecayMaxIter = 100;   // max_iter, fast_decay.hpp:26
constexpr double kDecaySigma = 1.0;  // sigma, fast_decay.hpp:27

// Strictly convex QP  min 1/2 x'Gx - d'x  s.t.  x[c] - x[c+1] >= 0 for c in cons
// by the Goldfarb-Idnani dual active-set method, each step solved directly from
// the Cholesky factor of G (n <= 50, so no factor updating is needed).
// L is the lower Cholesky factor of G (row-major n x n). x holds the
// unconstrained minimiser on entry and the constrained one on exit.
void qp_monotone_tail(int n, const std::vector<double> &L, const std::vector<int> &cons, std::vector<double> &x);

// dense helpers (row-major)
bool chol_lower(int n, const std::vector<double> &A, std::vector<double> &L);
void chol_solve(int n, const std::vector<double> &L, std::vector<double> &b);

class DecayGam {
public:
    // support: mean log-distance of each decay bin; [smin, smax]: range of the log-distances
    void setup(const std::vector<double> &support, double smin, double smax, double tol_val, unsigned free_decay);
    // one update_params: phihat / weight are the decay summary; nobs the per-bin observable counts
    void fit(const std::vector<double> &phihat, const std::vector<double> &weight, const std::vector<double> &nobs);
    // X * beta - mean for every bin (get_binned_estimate, fast_fitter.hpp:31-33)
    std::vector<double> binned_estimate() const;
    std::vector<double> estimate() const;

    int nb() const { return nb_; }
    // Params<GAM> (fast_fitter_gam.hpp:17-41)
    std::vector<double> beta = std::vector<double>(kDecayK, 0.0);
    double lambda = -1.0, mean = 0.0;
    int last_iters = 0;

private:
    int nb_ = 0;
    double tol_val_ = 0.2;
    // the cubic B-spline design has 4 non-zeros per row: first column + 4 values
    std::vector<int> xcol_;
    std::vector<double> xval_;
    std::vector<double> pen_;   // K x K, (K*D2)'(K*D2)
    std::vector<int> cons_;     // indices c with beta[c] - beta[c+1] >= 0
    // library state persisting across fits (QuadProgGAMLibrary.hpp:50-52, .cpp:12)
    std::vector<double> qp_beta_ = std::vector<double>(kDecayK, 1.0);
    double qp_lambda_ = 1.0;
};

} // namespace fb
