// fb_model.cuh -- host driver of the fast-binless path: owns the device-resident matrix, the three
// leg estimators and the outer IRLS loop. One Model = one fb_handle.
//
// Follows (paths relative to /root/reference/binless/src):
//   binless()             fast_binless.cpp:85-184
//   binless_eval_cv()     fast_binless.cpp:186-233
//   binless_difference()  fast_binless.cpp:235-279
//   step_signal / step_difference   fast_binless.ipp:2-105
//   FastDataCore          FastData.ipp:3-26
//   leg estimators        fast_exposure.hpp, fast_bias_mean.hpp, fast_decay.hpp, fast_fitter_mean.hpp
#pragma once
#include "../../include/fastbinless.h"
#include "fb_gam.hpp"
#include "fb_gfl.cuh"
#include "fb_irls.cuh"
#include <memory>
#include <string>
#include <vector>

namespace fb {

// pinned host buffer
template <typename T>
struct PinBuf {
    T *p = nullptr;
    size_t n = 0;
    PinBuf() {}
    PinBuf(const PinBuf &) = delete;
    PinBuf &operator=(const PinBuf &) = delete;
    ~PinBuf() { release(); }
    void release() { if (p) BlockCache::host().put(p, n * sizeof(T)); p = nullptr; n = 0; }
    void alloc(size_t n_) { release(); n = n_; if (n) p = static_cast<T *>(BlockCache::host().get(n * sizeof(T))); }
};

class Model {
public:
    Model(int nbins, const fb_mat *mat);
    Model(const fb_reads *reads, int *nbins_out);     // bin_data() on the device (fb_bin.cu)
    void get_mat(int32_t *name, int32_t *bin1, int32_t *pos1, int32_t *bin2, int32_t *pos2, double *distance,
                 int32_t *observed, int32_t *nobs);
    int64_t rows() const { return (int64_t)t_.D * t_.ncells; }
    ~Model();
    void run_signal(const fb_params *p, fb_signal_out *out);
    void restore(const double *log_signal, const fb_mean_state *biases, const fb_gam_state *decay,
                 const fb_mean_state *exposures);
    void run_difference(int ref, const fb_params *p, fb_diff_out *out);
    void eval_cv(const fb_params *p, int nlam, const double *lam2, unsigned group, fb_cv_out *outs);
    void sweep_lambda(const fb_params *p, int nlam, const double *lam2, fb_sweep_out *out);
    void cv_score(const fb_params *p, int nlam, const double *lam2, fb_cv_score_out *out);
    void band_export(const fb_band_params *p, fb_band_out *out);     // fb_band.cu

    std::string err;
    int (*interrupt_cb)(void *) = nullptr;
    void *interrupt_arg = nullptr;
    int64_t kernel_launches = 0;

private:
    Tri t_;
    cudaStream_t st_ = nullptr;
    int dev_ = 0;                     // the device the handle lives on (current at construction)
    cudaEvent_t ev0_ = nullptr, ev1_ = nullptr;
    // early download of the output columns that are final before the last fused-lasso solves (run_signal): a second stream
    cudaStream_t st_dl_ = nullptr;
    cudaEvent_t ev_in_ = nullptr, ev_sk_ = nullptr, ev_dl_ = nullptr;
    // device-resident matrix
    DevBuf<int32_t> obs_, nobs_;
    DevBuf<double> phi_, phi_new_, phihat_, wgt_;
    // structure tables found at ingest
    std::vector<double> pos_;   // [n] position of bin r
    std::vector<double> dist_;  // [n] distance of diagonal k
    // legs (host side; all O(n))
    bool legs_built_ = false, has_state_ = false, struct_built_ = false;
    double alpha_ = 1.0;
    int nb_bias_ = 0, nb_dec_ = 0;
    std::vector<int> rowbin_, dbin_;                  // row -> bias bin, diagonal -> decay bin
    std::vector<double> expo_nobs_, bias_nobs_, dec_nobs_;
    std::vector<double> expo_est_, bias_est_;         // Params<Mean>::estimate (mean = 0)
    DecayGam dec_;
    // device leg tables: [brow n | ddiag n | expo D], current and the copy taken before the decay update
    DevBuf<double> tab_, tab_old_, tab_sig_;          // tab_sig_: the tables of the last signal step (shared band solves)
    PinBuf<double> tab_h_, items_h_, prec_h_;
    DevBuf<double> part_, items_, items_g_, scratch_;   // items_g_: the per-group sums of a leg pass (fb_irls.cuh)
    DevBuf<long long> keys_;

    // per-cell fallback (util.cpp:56-111 bins every cell): set by ingest when pos / distance are not
    // functions of the bin / diagonal. Per cell the int32 bins of pos1, pos2 and log(distance); per leg bin the list of
    // its cells (CSR; bit 31 of an entry = the cell counts twice); the tables are then indexed by leg bin.
    bool generic_ = false;
    DevBuf<int32_t> cb1_, cb2_, cdb_;
    DevBuf<int64_t> bias_off_, dec_off_;
    DevBuf<uint32_t> bias_list_, dec_list_;
    DevBuf<double> cell_a_, cell_b_;
    std::vector<double> dec_support_;
    double dec_smin_ = 0, dec_smax_ = 0;
    int tab_b_ = 0, tab_d_ = 0;                       // entries of the bias / decay part of the device tables

    LegTables tables(const DevBuf<double> &b) const {
        LegTables L{b.p, b.p + tab_b_, b.p + (size_t)tab_b_ + tab_d_};
        if (generic_) { L.cb1 = cb1_.p; L.cb2 = cb2_.p; L.cdb = cdb_.p; }
        return L;
    }
    std::unique_ptr<GflSolver> flo_;                  // kept across calls (structure only; every call starts from reset())
    unsigned flo_maxdiag_ = 0; int flo_comm_ = 1;
    GflSolver &solver(unsigned maxdiag);
    Share share() const;
    void retire_solver();
    void open_stream();
    void close_stream();
    void ingest_cells(const fb_mat *mat);
    void alloc_tables();
    void ingest(int nbins, const fb_mat *mat);
    void ingest_reads(const fb_reads *reads, int *nbins_out);
    void alloc_work();
    void build_legs(double tol_val, unsigned free_decay);
    void upload_tables();
    // runs one leg pass and returns the per-bin sums (a, b) on the host
    void leg_sums(int leg, int mode, std::vector<double> &a, std::vector<double> &b);
    void fit_expo(int mode);
    void fit_bias(int mode);
    void fit_decay(int mode);
    void check_interrupt();
    void patchnos(const double *d_value, double tol, bool compute, int32_t *h_out);
    void download(double *h_dst, const double *d_src, size_t n);
    void download_bytes(void *h_dst, const void *d_src, size_t bytes, cudaStream_t st = nullptr);   // nullptr: the handle's stream
    void upload_bytes(void *d_dst, const void *h_src, size_t bytes);
};

extern thread_local std::string g_last_error;
void release_idle_solvers();      // the pool of idle solver objects of destroyed handles (fb_model.cu)

} // namespace fb
