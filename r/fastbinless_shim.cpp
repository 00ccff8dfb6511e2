// r/fastbinless_shim.cpp -- Rcpp shim that keeps the two R entry points of the reference unchanged and routes
// them to libfastbinless_cuda.so through the C ABI (include/fastbinless.h).
//
// It replaces, in the reference package, the two registrations
//   function("fast_binless", &binless::fast::binless, ...)                     binless/src/csnorm_cpp.cpp:18-22
//   function("fast_binless_difference", &binless::fast::binless_difference...) binless/src/csnorm_cpp.cpp:23-26
// with the functions below (same argument names, order and defaults; same returned list / data-frame layout:
// fast_binless.cpp:179-183, fast_dataframe.cpp:90-105 and :133-144).
//
// NOT compiled in this repository's image (no R, Rcpp or RcppEigen here: SURVEY.md F5). Build inside the
// reference package with  PKG_LIBS += -L<dir> -lfastbinless_cuda  and  PKG_CPPFLAGS += -I<repo>/include.
// The Python mirror binless_b200/_abi.py performs exactly the same marshalling and is what the tests exercise.
#include <Rcpp.h>
#include <string>
#include <vector>
#include "fastbinless.h"

using namespace Rcpp;

namespace {

struct HandleGuard {
    fb_handle *h = nullptr;
    ~HandleGuard() { if (h) fb_destroy(h); }
};

void check(int rc, const fb_handle *h) {
    // the callback below saw a pending user interrupt: raise it as the interrupt it is (the reference's
    // Rcpp::checkUserInterrupt() in the outer loop, fast_binless.cpp:124), not as an error
    if (rc == FB_ERR_INTERRUPT) throw Rcpp::internal::InterruptedException();
    if (rc != FB_OK) Rcpp::stop(std::string(fb_last_error(h)));   // same channel as the reference's Rcpp::stop
}

// asks R whether an interrupt is pending WITHOUT unwinding through the library's frames (R_CheckUserInterrupt would longjmp)
void check_interrupt_fn(void *) { R_CheckUserInterrupt(); }
int interrupt_cb(void *) { return R_ToplevelExec(check_interrupt_fn, nullptr) == FALSE ? 1 : 0; }

// options(binless.ngpus = N): the call shares every fused-lasso solve between N devices of this R process (fb_create_multi);
// the R signatures of the reference stay as they are
int option_ngpus() {
    SEXP v = Rf_GetOption1(Rf_install("binless.ngpus"));
    if (v == R_NilValue) return 1;
    const int n = Rf_asInteger(v);
    return n == NA_INTEGER || n < 1 ? 1 : n;
}
int create_handle(unsigned nbins, const fb_mat *m, fb_handle **h) {
    const int ngpus = option_ngpus();
    return ngpus > 1 ? fb_create_multi((int)nbins, m, ngpus, h) : fb_create((int)nbins, m, h);
}

// FastDataCore's view of the data frame (FastData.ipp:5-14): factors arrive as their integer codes
fb_mat as_fb_mat(const DataFrame &obs, IntegerVector &name, IntegerVector &bin1, IntegerVector &pos1,
                 IntegerVector &bin2, IntegerVector &pos2, IntegerVector &observed, IntegerVector &nobs,
                 NumericVector &distance) {
    name = obs["name"]; bin1 = obs["bin1"]; pos1 = obs["pos1"]; bin2 = obs["bin2"]; pos2 = obs["pos2"];
    observed = obs["observed"]; nobs = obs["nobs"]; distance = obs["distance"];
    fb_mat m;
    m.N = observed.size();
    m.name = name.begin(); m.bin1 = bin1.begin(); m.pos1 = pos1.begin(); m.bin2 = bin2.begin(); m.pos2 = pos2.begin();
    m.observed = observed.begin(); m.nobs = nobs.begin(); m.distance = distance.begin();
    return m;
}

fb_params as_fb_params(double alpha, const NumericVector &lam2, const NumericVector &lam1, unsigned nouter, double tol_val,
                       unsigned bg_steps, unsigned free_decay, bool compute_patchnos, const std::string &csv_out,
                       unsigned maxdiag) {
    fb_params p;
    p.alpha = alpha;
    p.lam2 = lam2.begin(); p.n_lam2 = lam2.size();
    p.lam1 = lam1.begin(); p.n_lam1 = lam1.size();
    p.nouter = nouter; p.tol_val = tol_val; p.bg_steps = bg_steps; p.free_decay = free_decay;
    p.compute_patchnos = compute_patchnos; p.csv_out = csv_out.c_str(); p.maxdiag = maxdiag;
    p.verbose = 1;                                                // the reference prints its progress lines
    return p;
}

} // namespace

// binless::fast::binless (fast_binless.hpp:29-31)
List fast_binless_cuda(const DataFrame obs, unsigned nbins, double alpha, const NumericVector lam2,
                       const NumericVector lam1, unsigned nouter, double tol_val, unsigned bg_steps,
                       unsigned free_decay, bool compute_patchnos, const std::string csv_out, unsigned maxdiag) {
    IntegerVector name, bin1, pos1, bin2, pos2, observed, nobs;
    NumericVector distance;
    fb_mat m = as_fb_mat(obs, name, bin1, pos1, bin2, pos2, observed, nobs, distance);
    HandleGuard g;
    check(create_handle(nbins, &m, &g.h), nullptr);
    fb_set_interrupt_cb(g.h, interrupt_cb, nullptr);   // refused (and not needed) on a multi-GPU handle
    const R_xlen_t N = m.N;
    const int D = Rcpp::max(name);
    // csv_out: the reference streams the seven computed columns to the (gzip) file and returns mat = 1
    // (fast_dataframe.cpp:23-57); the library then needs no host columns at all, only log_signal for the list
    const bool csv = !csv_out.empty();
    const R_xlen_t Nc = csv ? 0 : N;
    NumericVector background(Nc), biasmat(Nc), decaymat(Nc), signal(Nc), binless(Nc), phihat(Nc), weight(Nc), log_signal(N);
    IntegerVector patchno(Nc);
    NumericVector bias_est(nbins), expo_est(D);
    fb_signal_out o = fb_signal_out();
    if (!csv) {
        o.background = background.begin(); o.biasmat = biasmat.begin(); o.decaymat = decaymat.begin();
        o.signal = signal.begin(); o.binless = binless.begin(); o.phihat = phihat.begin(); o.weight = weight.begin();
        o.patchno = patchno.begin();
    }
    o.log_signal = log_signal.begin();
    o.biases.estimate = bias_est.begin(); o.exposures.estimate = expo_est.begin();
    fb_params p = as_fb_params(alpha, lam2, lam1, nouter, tol_val, bg_steps, free_decay, compute_patchnos, csv_out, maxdiag);
    check(fb_run_signal(g.h, &p, &o), g.h);
    SEXP mat_sexp;
    if (csv) mat_sexp = Rcpp::wrap(1);
    else {
        // the reference returns name/bin1/bin2 as the ordered factors it was given (FastData.hpp:27-29,50-55): pass them through
        DataFrame mat = DataFrame::create(_["name"] = obs["name"], _["bin1"] = obs["bin1"], _["pos1"] = pos1,
                                          _["bin2"] = obs["bin2"], _["pos2"] = pos2, _["observed"] = observed,
                                          _["nobs"] = nobs, _["distance"] = distance, _["background"] = background,
                                          _["biasmat"] = biasmat, _["decaymat"] = decaymat, _["signal"] = signal,
                                          _["binless"] = binless, _["phihat"] = phihat, _["weight"] = weight,
                                          _["patchno"] = NumericVector(patchno.begin(), patchno.end()));
        mat_sexp = mat;
    }
    NumericVector beta(o.decay.beta, o.decay.beta + FB_DECAY_K);
    if (o.biases.n < (int)nbins) bias_est = NumericVector(bias_est.begin(), bias_est.begin() + o.biases.n);
    return List::create(_["mat"] = mat_sexp,
                        _["biases"] = List::create(_["estimate"] = bias_est, _["mean"] = o.biases.mean),
                        _["decay"] = List::create(_["beta"] = beta, _["lambda"] = o.decay.lambda, _["mean"] = o.decay.mean),
                        _["exposures"] = List::create(_["estimate"] = expo_est, _["mean"] = o.exposures.mean),
                        _["log_signal"] = log_signal, _["nbins"] = nbins);
}

// binless::fast::binless_difference (fast_binless.hpp:36-38)
DataFrame fast_binless_difference_cuda(const List out, int ref, double alpha, const NumericVector lam2,
                                       const NumericVector lam1, double tol_val, bool compute_patchnos, unsigned maxdiag) {
    const unsigned nbins = out["nbins"];
    const DataFrame obs = as<DataFrame>(out["mat"]);
    IntegerVector name, bin1, pos1, bin2, pos2, observed, nobs;
    NumericVector distance;
    fb_mat m = as_fb_mat(obs, name, bin1, pos1, bin2, pos2, observed, nobs, distance);
    HandleGuard g;
    check(create_handle(nbins, &m, &g.h), nullptr);
    // set_state calls of fast_binless.cpp:244-260
    NumericVector log_signal = out["log_signal"];
    List lb = out["biases"], ld = out["decay"], le = out["exposures"];
    NumericVector be = lb["estimate"], ee = le["estimate"], db = ld["beta"];
    fb_mean_state sb = {be.begin(), (int)be.size(), as<double>(lb["mean"])};
    fb_mean_state se = {ee.begin(), (int)ee.size(), as<double>(le["mean"])};
    fb_gam_state sd;
    for (int k = 0; k < FB_DECAY_K; ++k) sd.beta[k] = db[k];
    sd.lambda = as<double>(ld["lambda"]); sd.mean = as<double>(ld["mean"]);
    check(fb_restore(g.h, log_signal.begin(), &sb, &sd, &se), g.h);
    const R_xlen_t N = m.N;
    NumericVector difference(N), deltahat(N), weight(N), phi_ref(N);
    IntegerVector patchno(N);
    fb_diff_out o = fb_diff_out();
    o.difference = difference.begin(); o.deltahat = deltahat.begin(); o.weight = weight.begin();
    o.phi_ref = phi_ref.begin(); o.patchno = patchno.begin();
    fb_params p = as_fb_params(alpha, lam2, lam1, 25, tol_val, 5, 10000, compute_patchnos, "", maxdiag);
    check(fb_run_difference(g.h, ref, &p, &o), g.h);
    return DataFrame::create(_["name"] = obs["name"], _["bin1"] = obs["bin1"], _["pos1"] = pos1, _["bin2"] = obs["bin2"],
                             _["pos2"] = pos2, _["observed"] = observed, _["nobs"] = nobs, _["difference"] = difference,
                             _["deltahat"] = deltahat, _["weight"] = weight, _["phi_ref"] = phi_ref,
                             _["patchno"] = NumericVector(patchno.begin(), patchno.end()));
}

// In binless/src/csnorm_cpp.cpp the two registrations become (everything else in RCPP_MODULE(binless_cpp) stays):
//   function("fast_binless", &fast_binless_cuda, List::create(_["obs"], _["nbins"], _["alpha"], _["lam2"], _["lam1"]=0.,
//            _["nouter"]=25, _["tol_val"]=2e-1, _["bg_steps"]=5, _["free_decay"]=10000, _["compute_patchnos"]=true,
//            _["csv_out"]="", _["maxdiag"]=0), "documentation for fast_binless ");
//   function("fast_binless_difference", &fast_binless_difference_cuda, List::create(_["out"], _["ref"], _["alpha"],
//            _["lam2"], _["lam1"]=0., _["tol_val"]=2e-1, _["compute_patchnos"]=true, _["maxdiag"]=0),
//            "documentation for fast_binless_difference ");

// ---- input side: bin_data(cs, resolution) for a CSnorm object (R/preprocess.R:114-146) on the device -----------------
// biases = cs@biases[, .(name, pos)] and counts = cs@counts[, .(name, pos1, pos2, contact.close, contact.far, contact.up,
// contact.down)] with `name` as integer codes; returns the data.table columns bin_data() builds (begin/end follow from
// pos and resolution in R).
// [[Rcpp::export]]
DataFrame bin_data_cuda(const DataFrame biases, const DataFrame counts, int ndatasets, int resolution) {
    IntegerVector cn = biases["name"], cp = biases["pos"];
    IntegerVector rn = counts["name"], p1 = counts["pos1"], p2 = counts["pos2"];
    IntegerVector cc = counts["contact.close"], cf = counts["contact.far"], cu = counts["contact.up"], cd = counts["contact.down"];
    fb_reads r;
    r.ndatasets = ndatasets; r.resolution = resolution;
    r.ncuts = cn.size(); r.cut_name = cn.begin(); r.cut_pos = cp.begin();
    r.nreads = rn.size(); r.rd_name = rn.begin(); r.rd_pos1 = p1.begin(); r.rd_pos2 = p2.begin();
    r.contact_close = cc.begin(); r.contact_far = cf.begin(); r.contact_up = cu.begin(); r.contact_down = cd.begin();
    HandleGuard g;
    int nbins = 0;
    check(fb_create_binned(&r, &nbins, &g.h), NULL);
    const R_xlen_t N = fb_mat_rows(g.h);
    IntegerVector name(N), bin1(N), pos1(N), bin2(N), pos2(N), observed(N), nobs(N);
    NumericVector distance(N);
    check(fb_get_mat(g.h, name.begin(), bin1.begin(), pos1.begin(), bin2.begin(), pos2.begin(), distance.begin(),
                     observed.begin(), nobs.begin()), g.h);
    return DataFrame::create(_["name"] = name, _["bin1"] = bin1, _["pos1"] = pos1, _["bin2"] = bin2, _["pos2"] = pos2,
                             _["distance"] = distance, _["observed"] = observed, _["nobs"] = nobs);
}

// ---- multi-GPU: one R process per GPU (e.g. under mpirun); the 128-byte id travels by whatever the host has --------
// [[Rcpp::export]]
RawVector fastbinless_comm_id() { RawVector id(128); check(fb_comm_unique_id(id.begin()), NULL); return id; }
// [[Rcpp::export]]
void fastbinless_comm_init(int rank, int world, RawVector id) { check(fb_comm_init(rank, world, world > 1 ? id.begin() : NULL), NULL); }
