/* ref_gfl_wrap.cpp -- TEST INFRASTRUCTURE. A plain C face on the reference's own GFLLibrary object
 * (/root/reference/binless/src/GFLLibrary.hpp:15-64 with GFLLibraryTriangleImpl.cpp and GFLLibraryTrapezeImpl.cpp,
 * compiled where they lie by oracle/Makefile against oracle/stub/rcpp/Rcpp.h), so that the restatement in
 * oracle/fast_binless_oracle.cpp (trail decomposition, extract_trapeze / fill_triangle, get_state / set_state) can be
 * checked against the reference itself, bit for bit (tests/test_oracle_refgfl.py). */
#include <cstring>
#include <vector>
#define private public          /* the test reads trails_ / breakpoints_ of the reference objects */
#include "GFLLibrary.hpp"
#undef private

extern "C" {

void *refgfl_new(unsigned nrows, unsigned maxdiag) { try { return new GFLLibrary(nrows, maxdiag); } catch (...) { return nullptr; } }
void refgfl_free(void *h) { delete static_cast<GFLLibrary *>(h); }
void refgfl_reset(void *h) { static_cast<GFLLibrary *>(h)->reset(); }

/* 1 = trapeze, 0 = triangle; N = cells of the lasso problem, tsz = trail slots, ntrails */
int refgfl_geometry(void *h, unsigned *N, unsigned *tsz, int *ntrails) {
    GFLLibrary *g = static_cast<GFLLibrary *>(h);
    if (auto *t = dynamic_cast<GFLLibraryTriangleImpl *>(g->strategy_.get())) { *N = t->N_; *tsz = t->tsz_; *ntrails = t->ntrails_; return 0; }
    auto *t = dynamic_cast<GFLLibraryTrapezeImpl *>(g->strategy_.get());
    *N = t->N_; *tsz = t->tsz_; *ntrails = t->ntrails_; return 1;
}
void refgfl_trails(void *h, unsigned *trails, unsigned *breakpoints) {
    GFLLibrary *g = static_cast<GFLLibrary *>(h);
    const std::vector<unsigned> *tr, *bp;
    if (auto *t = dynamic_cast<GFLLibraryTriangleImpl *>(g->strategy_.get())) { tr = &t->trails_; bp = &t->breakpoints_; }
    else { auto *t2 = dynamic_cast<GFLLibraryTrapezeImpl *>(g->strategy_.get()); tr = &t2->trails_; bp = &t2->breakpoints_; }
    std::memcpy(trails, tr->data(), tr->size() * sizeof(unsigned));
    std::memcpy(breakpoints, bp->data(), bp->size() * sizeof(unsigned));
}
/* y, w: the full packed triangle (ncells = nrows (nrows + 1) / 2); beta_full likewise */
void refgfl_optimize(void *h, unsigned ncells, const double *y, const double *w, double lam2, double converge, double *beta_full) {
    GFLLibrary *g = static_cast<GFLLibrary *>(h);
    g->optimize(std::vector<double>(y, y + ncells), std::vector<double>(w, w + ncells), lam2, converge);
    const std::vector<double> b = g->get_beta();
    std::memcpy(beta_full, b.data(), b.size() * sizeof(double));
}
void refgfl_get_state(void *h, double *z, double *u, double *beta, double *alpha, unsigned *counter, int *maxdiag) {
    const Rcpp::List st = static_cast<GFLLibrary *>(h)->get_state();
    const std::vector<double> vz = Rcpp::as<std::vector<double> >(st["z"]), vu = Rcpp::as<std::vector<double> >(st["u"]),
                              vb = Rcpp::as<std::vector<double> >(st["beta"]);
    std::memcpy(z, vz.data(), vz.size() * sizeof(double)); std::memcpy(u, vu.data(), vu.size() * sizeof(double));
    std::memcpy(beta, vb.data(), vb.size() * sizeof(double));
    *alpha = Rcpp::as<double>(st["alpha"]); *counter = Rcpp::as<unsigned>(st["counter"]); *maxdiag = Rcpp::as<int>(st["maxdiag"]);
}
void refgfl_set_state(void *h, unsigned N, unsigned tsz, const double *z, const double *u, const double *beta, double alpha,
                      unsigned counter, int maxdiag) {
    using namespace Rcpp;
    static_cast<GFLLibrary *>(h)->set_state(List::create(_["z"] = std::vector<double>(z, z + tsz), _["u"] = std::vector<double>(u, u + tsz),
                                                         _["alpha"] = alpha, _["beta"] = std::vector<double>(beta, beta + N),
                                                         _["counter"] = counter, _["maxdiag"] = maxdiag));
}
unsigned refgfl_ninner(void *h) { return static_cast<GFLLibrary *>(h)->get_ninner(); }
double refgfl_alpha(void *h) { return static_cast<GFLLibrary *>(h)->get_alpha(); }

}  /* extern "C" */
