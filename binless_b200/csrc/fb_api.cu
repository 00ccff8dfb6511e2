// fb_api.cu -- the C ABI of libfastbinless_cuda.so (include/fastbinless.h). Thin: argument checks,
// exception -> status translation, nothing else. The Rcpp shim (r/fastbinless_shim.cpp) and the ctypes
// mirror (binless_b200/_abi.py) bind exactly these symbols.
#include "fb_model.cuh"
#include "fb_comm.cuh"
#include "fb_csv.hpp"
#include "fb_perf.cuh"
#include <atomic>
#include <exception>
#include <memory>
#include <string>
#include <thread>
#include <vector>

// One handle = one Model on one device, or (fb_create_multi) one Model per device of this process, every one driven by
// its own host thread with its own NCCL communicator for the duration of a call: the ranks make the same calls on the
// same inputs and share every fused-lasso solve (trail blocks), exactly like the one-process-per-GPU mode.
struct fb_handle {
    struct Rank { int dev = 0; fb::Comm comm; std::unique_ptr<fb::Model> m; };
    std::vector<std::unique_ptr<Rank>> ranks;
    fb::Model *lead = nullptr;                    // rank 0: the one whose outputs the caller sees
    fb::Model &m() { return *lead; }
    const fb::Model &m() const { return *lead; }
    bool multi() const { return ranks.size() > 1; }

    // f(rank index, model) on every rank, each on its own thread / device / communicator; the first exception wins
    template <typename F>
    void on_all(F &&f) {
        if (!multi()) { f(0, *lead); return; }
        std::vector<std::exception_ptr> ex(ranks.size());
        std::vector<std::thread> th;
        for (size_t r = 0; r < ranks.size(); ++r)
            th.emplace_back([&, r] {
                try {
                    FB_CUDA(cudaSetDevice(ranks[r]->dev));
                    fb::Comm::set_current(&ranks[r]->comm);
                    f((int)r, *ranks[r]->m);
                } catch (...) { ex[r] = std::current_exception(); }
                fb::Comm::set_current(nullptr);
            });
        for (auto &t : th) t.join();
        for (auto &e : ex) if (e) std::rethrow_exception(e);
    }

    fb_handle(int nbins, const fb_mat *mat) { single(new fb::Model(nbins, mat)); }
    fb_handle(const fb_reads *reads, int *nbins) { single(new fb::Model(reads, nbins)); }
    fb_handle(int nbins, const fb_mat *mat, int ngpus) {
        int ndev = 0;
        if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) throw fb::CudaError("no CUDA device available (there is no CPU fallback)");
        if (ngpus < 1 || ngpus > ndev) throw fb::ArgError("ngpus must be between 1 and the number of visible devices");
        if (ngpus == 1) { single(new fb::Model(nbins, mat)); return; }
        char id[128];
        { fb::Comm tmp; tmp.unique_id(id); }
        for (int r = 0; r < ngpus; ++r) { ranks.emplace_back(new Rank()); ranks.back()->dev = r; }
        std::vector<std::exception_ptr> ex(ngpus);
        std::vector<std::thread> th;
        for (int r = 0; r < ngpus; ++r)
            th.emplace_back([&, r] {
                try {
                    FB_CUDA(cudaSetDevice(r));
                    ranks[r]->comm.init(r, ngpus, id);
                    fb::Comm::set_current(&ranks[r]->comm);
                    ranks[r]->m.reset(new fb::Model(nbins, mat));
                } catch (...) { ex[r] = std::current_exception(); }
                fb::Comm::set_current(nullptr);
            });
        for (auto &t : th) t.join();
        for (auto &e : ex) if (e) { destroy_ranks(); std::rethrow_exception(e); }
        lead = ranks[0]->m.get();
    }
    ~fb_handle() { destroy_ranks(); }

private:
    void single(fb::Model *p) { ranks.emplace_back(new Rank()); ranks[0]->m.reset(p); lead = p; cudaGetDevice(&ranks[0]->dev); }
    void destroy_ranks() {
        if (!multi()) { ranks.clear(); return; }
        std::vector<std::thread> th;
        for (size_t r = 0; r < ranks.size(); ++r)
            th.emplace_back([&, r] {     // device buffers go back to the cache of THEIR device
                cudaSetDevice(ranks[r]->dev);
                fb::Comm::set_current(&ranks[r]->comm);   // the solvers leave the exchange collectively
                ranks[r]->m.reset();
                cudaDeviceSynchronize();
                ranks[r]->comm.destroy();
                fb::Comm::set_current(nullptr);
            });
        for (auto &t : th) t.join();
        ranks.clear();
    }
};

namespace {
template <typename F>
int guarded(std::string &err, F &&f) {
    try { f(); return FB_OK; }
    catch (const fb::ArgError &e) { err = e.what(); return FB_ERR_ARG; }
    catch (const fb::CudaError &e) { err = e.what(); return FB_ERR_CUDA; }
    catch (const fb::StateError &e) { err = e.what(); return FB_ERR_STATE; }
    catch (const fb::NumError &e) { err = e.what(); return FB_ERR_NUMERIC; }
    catch (const fb::Interrupted &e) { err = e.what(); return FB_ERR_INTERRUPT; }
    catch (const fb::CommError &e) { err = e.what(); return FB_ERR_COMM; }
    catch (const std::bad_alloc &) { err = "out of host memory"; return FB_ERR_ARG; }
    catch (const std::exception &e) { err = e.what(); return FB_ERR_NUMERIC; }
}
} // namespace

extern "C" {

FB_API int fb_create(int nbins, const fb_mat *mat, fb_handle **out) {
    if (!out) { fb::g_last_error = "null handle pointer"; return FB_ERR_ARG; }
    *out = nullptr;
    return guarded(fb::g_last_error, [&] { *out = new fb_handle(nbins, mat); });
}

FB_API int fb_create_multi(int nbins, const fb_mat *mat, int ngpus, fb_handle **out) {
    if (!out) { fb::g_last_error = "null handle pointer"; return FB_ERR_ARG; }
    *out = nullptr;
    return guarded(fb::g_last_error, [&] { *out = new fb_handle(nbins, mat, ngpus); });
}
FB_API int fb_handle_ngpus(const fb_handle *h) { return h ? (int)h->ranks.size() : 0; }

FB_API int fb_create_binned(const fb_reads *reads, int *nbins, fb_handle **out) {
    if (!out) { fb::g_last_error = "null handle pointer"; return FB_ERR_ARG; }
    *out = nullptr;
    return guarded(fb::g_last_error, [&] { *out = new fb_handle(reads, nbins); });
}
FB_API int fb_get_mat(fb_handle *h, int32_t *name, int32_t *bin1, int32_t *pos1, int32_t *bin2, int32_t *pos2,
                      double *distance, int32_t *observed, int32_t *nobs) {
    if (!h) { fb::g_last_error = "null handle"; return FB_ERR_ARG; }
    return guarded(h->m().err, [&] { h->m().get_mat(name, bin1, pos1, bin2, pos2, distance, observed, nobs); });
}
FB_API int64_t fb_mat_rows(const fb_handle *h) { return h ? h->m().rows() : 0; }

FB_API void fb_destroy(fb_handle *h) { delete h; }

FB_API const char *fb_last_error(const fb_handle *h) { return h ? h->m().err.c_str() : fb::g_last_error.c_str(); }

FB_API int fb_run_signal(fb_handle *h, const fb_params *p, fb_signal_out *out) {
    if (!h) { fb::g_last_error = "null handle"; return FB_ERR_ARG; }
    return guarded(h->m().err, [&] {
        h->on_all([&](int r, fb::Model &m) {
            if (r == 0) { m.run_signal(p, out); return; }
            fb_signal_out o = fb_signal_out();      // the other ranks compute the same thing and keep it to themselves
            fb_params q = *p; q.csv_out = nullptr; q.verbose = 0;
            m.run_signal(&q, &o);
        });
    });
}

FB_API int fb_restore(fb_handle *h, const double *log_signal, const fb_mean_state *biases, const fb_gam_state *decay,
                      const fb_mean_state *exposures) {
    if (!h) { fb::g_last_error = "null handle"; return FB_ERR_ARG; }
    return guarded(h->m().err, [&] { h->on_all([&](int, fb::Model &m) { m.restore(log_signal, biases, decay, exposures); }); });
}

FB_API int fb_run_difference(fb_handle *h, int ref, const fb_params *p, fb_diff_out *out) {
    if (!h) { fb::g_last_error = "null handle"; return FB_ERR_ARG; }
    return guarded(h->m().err, [&] {
        h->on_all([&](int r, fb::Model &m) {
            if (r == 0) { m.run_difference(ref, p, out); return; }
            fb_diff_out o = fb_diff_out();
            fb_params q = *p; q.verbose = 0;
            m.run_difference(ref, &q, &o);
        });
    });
}

FB_API int fb_eval_cv(fb_handle *h, const fb_params *p, int nlam, const double *lam2, unsigned group, fb_cv_out *outs) {
    if (!h) { fb::g_last_error = "null handle"; return FB_ERR_ARG; }
    return guarded(h->m().err, [&] {
        h->on_all([&](int r, fb::Model &m) {
            if (r == 0) { m.eval_cv(p, nlam, lam2, group, outs); return; }
            std::vector<fb_cv_out> o((size_t)std::max(0, nlam));
            m.eval_cv(p, nlam, lam2, group, o.data());
        });
    });
}

FB_API int fb_sweep_lambda(fb_handle *h, const fb_params *p, int nlam, const double *lam2, fb_sweep_out *out) {
    if (!h) { fb::g_last_error = "null handle"; return FB_ERR_ARG; }
    return guarded(h->m().err, [&] {
        h->on_all([&](int r, fb::Model &m) {
            if (r == 0) { m.sweep_lambda(p, nlam, lam2, out); return; }
            fb_sweep_out o = fb_sweep_out();
            m.sweep_lambda(p, nlam, lam2, &o);
        });
    });
}

FB_API int fb_cv_score(fb_handle *h, const fb_params *p, int nlam, const double *lam2, fb_cv_score_out *out) {
    if (!h) { fb::g_last_error = "null handle"; return FB_ERR_ARG; }
    return guarded(h->m().err, [&] {
        h->on_all([&](int r, fb::Model &m) {
            if (r == 0) { m.cv_score(p, nlam, lam2, out); return; }
            fb_cv_score_out o = fb_cv_score_out();
            m.cv_score(p, nlam, lam2, &o);
        });
    });
}

FB_API int fb_band_export(fb_handle *h, const fb_band_params *p, fb_band_out *out) {
    if (!h) { fb::g_last_error = "null handle"; return FB_ERR_ARG; }
    return guarded(h->m().err, [&] { h->m().band_export(p, out); });   // no solve inside: rank 0 alone on a multi-GPU handle
}

FB_API int fb_csv_write(const char *path, const char *header, int64_t nrows, int ncols, const double *const *cols, int gzip) {
    if (!path || !header || nrows < 0 || ncols <= 0 || !cols) { fb::g_last_error = "bad argument"; return FB_ERR_ARG; }
    return guarded(fb::g_last_error, [&] {
        if (!fb::write_csv_gz(path, header, nrows, ncols, cols, gzip != 0)) throw fb::ArgError(std::string("cannot write ") + path);
    });
}

FB_API int fb_wgfl_signal_bic(int nbins, const fb_cts *cts, const fb_perf_meta *meta, const fb_perf_params *p, const double *beta_init,
                              int warm, fb_gfl_state *gfl, fb_perf_signal_out *out) {
    return guarded(fb::g_last_error, [&] { fb::wgfl_signal_bic(nbins, cts, meta, p, beta_init, warm, gfl, out); });
}
FB_API int fb_wgfl_diff_bic(int nbins, const fb_cts *cts, const fb_cts *ref, const fb_perf_meta *meta, const fb_perf_params *p,
                            const double *phi_ref_init, const double *beta_init, int warm, fb_gfl_state *gfl, fb_perf_diff_out *out) {
    return guarded(fb::g_last_error, [&] { fb::wgfl_diff_bic(nbins, cts, ref, meta, p, phi_ref_init, beta_init, warm, gfl, out); });
}

FB_API int fb_set_interrupt_cb(fb_handle *h, int (*cb)(void *), void *arg) {
    if (!h) return FB_ERR_ARG;
    if (h->multi()) return FB_ERR_ARG;   // the ranks could not agree on an interrupt seen by one host thread
    h->m().interrupt_cb = cb; h->m().interrupt_arg = arg;
    return FB_OK;
}

FB_API void fb_release_cached_memory(void) {
    cudaDeviceSynchronize();
    fb::release_idle_solvers();
    fb::BlockCache::device().trim();
    fb::BlockCache::host().trim();
}

FB_API int64_t fb_kernel_launches(const fb_handle *h) { return h ? h->m().kernel_launches : 0; }

namespace { thread_local double tl_last_solve_ms = 0.0; std::atomic<double> g_multi_solve_ms{0.0}; }
namespace {
// FusedLassoGaussianEstimator<GFLLibrary>::optimize on the current device (and communicator)
void gfl_solve_body(int nbins, unsigned maxdiag, const double *y, const double *w, double lam2, double converge, int maxsteps,
                    int warm, fb_gfl_state *st, double *beta_full) {
    const size_t nc = (size_t)nbins * (nbins + 1) / 2;
    struct Stream { // a real stream (ordered with the default stream): the ADMM loop is captured into a graph
        cudaStream_t st = nullptr;
        Stream() { FB_CUDA(cudaStreamCreate(&st)); }
        ~Stream() { if (st) cudaStreamDestroy(st); }
    } stream;
    fb::GflSolver s(nbins, maxdiag, stream.st);
    s.ninner_max = maxsteps;
    fb::DevBuf<double> dy(nc), dw(nc), dout(nc);
    dy.upload(y, nc); dw.upload(w, nc);
    if (warm) { s.alpha = st->alpha; s.import_state(st->beta, st->z, st->u); }
    s.optimize(dy.p, dw.p, lam2, converge);
    tl_last_solve_ms = s.ms_last;
    if (fb::Comm::get().rank() == 0) g_multi_solve_ms.store(s.ms_last);
    st->alpha = s.alpha; st->steps = s.last_steps;
    s.export_state(st->beta, st->z, st->u);
    if (beta_full) {
        s.get_beta_full(dout.p);
        FB_CUDA(cudaDeviceSynchronize());
        dout.download(beta_full, nc);
        FB_CUDA(cudaDeviceSynchronize());
    }
}
} // namespace

FB_API int fb_gfl_solve(int nbins, unsigned maxdiag, const double *y, const double *w, double lam2, double converge,
                        int maxsteps, int warm, fb_gfl_state *st, double *beta_full) {
    return guarded(fb::g_last_error, [&] {
        int ndev = 0;
        if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0)
            throw fb::CudaError("no CUDA device available (there is no CPU fallback)");
        if (!y || !w || !st) throw fb::ArgError("null argument");
        gfl_solve_body(nbins, maxdiag, y, w, lam2, converge, maxsteps, warm, st, beta_full);
    });
}

// the same solve shared by `ngpus` devices of THIS process: one host thread and one NCCL communicator per device
FB_API int fb_gfl_solve_multi(int ngpus, int nbins, unsigned maxdiag, const double *y, const double *w, double lam2, double converge,
                              int maxsteps, int warm, fb_gfl_state *st, double *beta_full) {
    return guarded(fb::g_last_error, [&] {
        int ndev = 0;
        if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0)
            throw fb::CudaError("no CUDA device available (there is no CPU fallback)");
        if (!y || !w || !st) throw fb::ArgError("null argument");
        if (ngpus < 1 || ngpus > ndev) throw fb::ArgError("ngpus must be between 1 and the number of visible devices");
        if (ngpus == 1) { gfl_solve_body(nbins, maxdiag, y, w, lam2, converge, maxsteps, warm, st, beta_full); return; }
        char id[128];
        { fb::Comm tmp; tmp.unique_id(id); }
        const size_t nc = (size_t)nbins * (nbins + 1) / 2;
        std::vector<std::exception_ptr> ex(ngpus);
        std::vector<std::thread> th;
        for (int r = 0; r < ngpus; ++r)
            th.emplace_back([&, r] {
                fb::Comm comm;
                try {
                    FB_CUDA(cudaSetDevice(r));
                    comm.init(r, ngpus, id);
                    fb::Comm::set_current(&comm);
                    if (r == 0) gfl_solve_body(nbins, maxdiag, y, w, lam2, converge, maxsteps, warm, st, beta_full);
                    else {   // the same call with private output arrays (state export is collective)
                        std::vector<double> b(st->beta ? nc : 0), z(st->z ? 2 * nc : 0), u(st->u ? 2 * nc : 0);
                        if (warm) {
                            if (st->beta) std::copy(st->beta, st->beta + nc, b.begin());   // full-triangle solves only carry nc / 2nc-2 entries
                            if (st->z) std::copy(st->z, st->z + (2 * nc - 2), z.begin());
                            if (st->u) std::copy(st->u, st->u + (2 * nc - 2), u.begin());
                        }
                        fb_gfl_state s2 = *st;
                        s2.beta = st->beta ? b.data() : nullptr; s2.z = st->z ? z.data() : nullptr; s2.u = st->u ? u.data() : nullptr;
                        gfl_solve_body(nbins, maxdiag, y, w, lam2, converge, maxsteps, warm, &s2, nullptr);
                    }
                } catch (...) { ex[r] = std::current_exception(); }
                fb::Comm::set_current(nullptr);
                cudaDeviceSynchronize();
                comm.destroy();
            });
        for (auto &t : th) t.join();
        for (auto &e : ex) if (e) std::rethrow_exception(e);
        tl_last_solve_ms = g_multi_solve_ms.load();      // rank 0's thread measured it
    });
}

FB_API double fb_gfl_last_solve_ms(void) { return tl_last_solve_ms; }

// ---- one process per GPU: trail-block sharding of the fused-lasso solve over the ranks of one NCCL communicator ----
FB_API int fb_comm_unique_id(void *id128) {
    if (!id128) { fb::g_last_error = "null argument"; return FB_ERR_ARG; }
    return guarded(fb::g_last_error, [&] { fb::Comm::get().unique_id(id128); });
}
FB_API int fb_comm_init(int rank, int world, const void *id128) {
    if (world > 1 && !id128) { fb::g_last_error = "null argument"; return FB_ERR_ARG; }
    return guarded(fb::g_last_error, [&] {
        int ndev = 0;
        if (world > 1 && (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0))
            throw fb::CudaError("no CUDA device available (there is no CPU fallback)");
        fb::Comm::get().init(rank, world, id128);
    });
}
FB_API int fb_comm_destroy(void) { return guarded(fb::g_last_error, [&] { fb::Comm::get().destroy(); }); }
FB_API int fb_comm_rank(void) { return fb::Comm::get().rank(); }
FB_API int fb_comm_world(void) { return fb::Comm::get().world(); }
FB_API int fb_shard_plan(int nbins, unsigned maxdiag, int world, int32_t *row_begin, int32_t *col_begin) {
    if (!row_begin || !col_begin) { fb::g_last_error = "null argument"; return FB_ERR_ARG; }
    return guarded(fb::g_last_error, [&] {
        unsigned md = maxdiag < (unsigned)nbins ? maxdiag : (unsigned)nbins;        // FusedLassoGaussianEstimator.hpp:33
        if (md > (unsigned)(nbins + 1) / 2 || md <= 1) md = nbins;                  // GFLLibrary.hpp:21-27
        fb::shard_plan(nbins, (int)md, world, row_begin, col_begin);
    });
}

} // extern "C"
