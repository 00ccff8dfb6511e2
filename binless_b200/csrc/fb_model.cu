// fb_model.cu -- host driver of the fast-binless path (see fb_model.cuh).
#include "fb_model.cuh"
#include "fb_csv.hpp"
#include "fb_comm.cuh"
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <atomic>
#include <chrono>
#include <cstdlib>
#include <future>
#include <map>
#include <thread>

namespace fb {

thread_local std::string g_last_error;   // per host thread: handles may be created from several threads

namespace {

// development aid: FB_TIMING=1 prints host wall-clock per phase of run_signal
struct PhaseTimer {
    bool on = std::getenv("FB_TIMING") != nullptr;
    double acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    std::chrono::steady_clock::time_point t;
    void start() { if (on) t = std::chrono::steady_clock::now(); }
    void stop(int k) { if (on) acc[k] += std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t).count(); }
};
PhaseTimer g_pt;

// pos1[k] == pi, pos2[k] == pj[k], distance[k] == dj[k] for every k < len
__attribute__((optimize("O3"))) bool row_is_layout(const int32_t *__restrict__ p1, const int32_t *__restrict__ p2,
                                                   const double *__restrict__ dd, const int32_t pi, const int32_t *__restrict__ pj,
                                                   const double *__restrict__ dj, const int64_t len) {
    int bad = 0;
    for (int64_t k = 0; k < len; ++k) bad |= (p1[k] != pi) | (p2[k] != pj[k]) | (dd[k] != dj[k]);
    return bad == 0;
}

// name[k] == d1, bin1[k] == i1, bin2[k] == i1 + k for every k < len (columns that are absent are not looked at)
__attribute__((optimize("O3"))) bool row_is_ordered(const int32_t *__restrict__ nm, const int32_t *__restrict__ b1,
                                                    const int32_t *__restrict__ b2, const int32_t d1, const int32_t i1, const int64_t len) {
    int bad = 0;
    if (nm) for (int64_t k = 0; k < len; ++k) bad |= (nm[k] != d1);
    if (b1) for (int64_t k = 0; k < len; ++k) bad |= (b1[k] != i1);
    if (b2) for (int64_t k = 0; k < len; ++k) bad |= (b2[k] != i1 + (int32_t)k);
    return bad == 0;
}

const char *kLayoutMsg = "pos1/pos2 do not depend on the bin only or distance not on bin2-bin1 only";   // -> per-cell fallback
const char *kSizeMsg = "Input size should be N(N+1)/2 dense matrix of observed counts for each dataset";

// make_even_bins (util.cpp:56-68) as a sorted edge list: (edge value, index). The reference keeps the edges in a
// std::map, so equal edge values (a degenerate range) keep the LAST index assigned; that case goes through a map here
// too, the usual strictly increasing edges do not need one.
struct EvenBins {
    std::vector<double> key; std::vector<unsigned> val; unsigned Nbins;
    EvenBins(double lower, double upper, unsigned Nb) : Nbins(Nb) {
        const double size = upper - lower;
        key.resize(Nb + 1); val.resize(Nb + 1);
        bool increasing = true;
        for (unsigned i = 0; i <= Nb; ++i) {
            key[i] = lower + i * size / double(Nb); val[i] = i;
            if (i && !(key[i] > key[i - 1])) increasing = false;
        }
        if (!increasing) {
            std::map<double, unsigned> edges;
            for (unsigned i = 0; i <= Nb; ++i) edges[lower + i * size / double(Nb)] = i;
            key.clear(); val.clear();
            for (auto &kv : edges) { key.push_back(kv.first); val.push_back(kv.second); }
        }
    }
    int bin(double v) const {   // bin_data (util.cpp:70-81): upper_bound, past the end -> last bin
        const size_t k = std::upper_bound(key.begin(), key.end(), v) - key.begin();
        return (int)((k == key.size() ? val.back() : val[k]) - 1);
    }
};

// make_even_bins + bin_data with drop=true (util.cpp:56-111): bin index of each value, unused bins removed
std::vector<int> even_bins_dropped(const std::vector<double> &v, double lower, double upper, unsigned Nbins, int &nkept) {
    const EvenBins eb(lower, upper, Nbins);
    std::vector<int> raw(v.size());
    std::vector<char> used(Nbins, 0);
    for (size_t k = 0; k < v.size(); ++k) {
        const int j = eb.bin(v[k]);
        raw[k] = j;
        used[j] = 1;
    }
    std::vector<int> remap(Nbins, -1);
    nkept = 0;
    for (unsigned b = 0; b < Nbins; ++b) if (used[b]) remap[b] = nkept++;
    for (auto &r : raw) r = remap[r];
    return raw;
}

std::vector<double> broadcast(const double *p, int np, int n) {
    if (!p || np <= 0) return std::vector<double>(n, 0.0);
    if (np == n) return std::vector<double>(p, p + n);
    return std::vector<double>(n, p[0]);
}

// csv_out (fast_dataframe.cpp:23-57): gzip like the reference as built (USE_BOOST = 1, Makevars:1-2; csvfile.hpp:46-48);
// FB_CSV_PLAIN=1 writes the plain text of a USE_BOOST = 0 build
void write_signal_csv(const char *path, int64_t N, const double *const cols[7]) {
    const char *e = std::getenv("FB_CSV_PLAIN");
    write_csv_gz(path, "\"phihat\",\"weight\",\"biasmat\",\"decaymat\",\"signal\",\"binless\",\"background\"", N, 7, cols,
                 !(e && std::atoi(e) != 0));
}

// has everything recorded before `e` finished? (a "not ready" answer must not stay behind as the thread's last CUDA error)
bool event_done(cudaEvent_t e) {
    const cudaError_t r = cudaEventQuery(e);
    if (r == cudaSuccess) return true;
    cudaGetLastError();
    if (r != cudaErrorNotReady) throw CudaError(std::string("cudaEventQuery: ") + cudaGetErrorString(r));
    return false;
}

int host_threads();
// host threads that move one staging chunk (FB_STAGE_THREADS; default: up to 8 of the process's host threads)
int stage_threads() {
    if (const char *e = std::getenv("FB_STAGE_THREADS")) return std::max(1, std::min(64, std::atoi(e)));
    return std::max(1, std::min(8, host_threads()));
}

// true for host memory the driver can DMA into directly (cudaHostAlloc / cudaHostRegister / managed)
bool host_pinned(const void *p) {
    cudaPointerAttributes at{};
    if (cudaPointerGetAttributes(&at, p) != cudaSuccess) { cudaGetLastError(); return false; }
    return at.type == cudaMemoryTypeHost || at.type == cudaMemoryTypeManaged;
}

// copies of at least this many bytes to / from pageable host memory are pipelined through pinned staging
// (FB_STAGE_MIN_MB lowers the limit so that the tests exercise the path at small sizes)
size_t stage_min_bytes() {
    static const size_t v = [] {
        const char *e = std::getenv("FB_STAGE_MIN_MB");
        return e ? (size_t)std::max(1, std::atoi(e)) << 20 : size_t(64) << 20;
    }();
    return v;
}

// Host threads of this process for validation / staging copies: FB_HOST_THREADS, else the cores divided by the ranks that
// share the machine (LOCAL_WORLD_SIZE, set by torchrun / mpirun wrappers): eight ranks with sixteen threads each on a
// sixteen-core host was what made the end-to-end time collapse at 8 GPUs.
int host_threads() {
    static const int v = [] {
        const unsigned h = std::max(1u, std::thread::hardware_concurrency());
        if (const char *e = std::getenv("FB_HOST_THREADS")) return std::max(1, std::min(64, std::atoi(e)));
        int local = 1;
        if (const char *e = std::getenv("LOCAL_WORLD_SIZE")) local = std::max(1, std::atoi(e));
        return (int)std::max(1u, std::min(16u, h / (unsigned)local));
    }();
    return v;
}

} // namespace

// ---------------------------------------------------------------------------------------------
// ingest: FastDataCore (FastData.ipp:3-26) + verification of the bin_data layout
// ---------------------------------------------------------------------------------------------
void Model::open_stream() {
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0)
        throw CudaError("no CUDA device available (there is no CPU fallback)");
    FB_CUDA(cudaGetDevice(&dev_));
    FB_CUDA(cudaStreamCreateWithFlags(&st_, cudaStreamNonBlocking));
    FB_CUDA(cudaEventCreate(&ev0_));
    FB_CUDA(cudaEventCreate(&ev1_));
    FB_CUDA(cudaStreamCreateWithFlags(&st_dl_, cudaStreamNonBlocking));
    FB_CUDA(cudaEventCreateWithFlags(&ev_in_, cudaEventDisableTiming));
    FB_CUDA(cudaEventCreateWithFlags(&ev_sk_, cudaEventDisableTiming));
    FB_CUDA(cudaEventCreateWithFlags(&ev_dl_, cudaEventDisableTiming));
}
void Model::close_stream() {
    retire_solver();
    if (st_) cudaStreamSynchronize(st_); // buffers go back to the block cache only when nothing uses them any more
    if (st_dl_) cudaStreamSynchronize(st_dl_);
    if (ev0_) cudaEventDestroy(ev0_);
    if (ev1_) cudaEventDestroy(ev1_);
    for (cudaEvent_t e : {ev_in_, ev_sk_, ev_dl_}) if (e) cudaEventDestroy(e);
    if (st_) cudaStreamDestroy(st_);
    if (st_dl_) cudaStreamDestroy(st_dl_);
    st_ = st_dl_ = nullptr; ev0_ = ev1_ = ev_in_ = ev_sk_ = ev_dl_ = nullptr;
}

// a constructor that throws (rejected input: the normal validation path) does not run the destructor
Model::Model(int nbins, const fb_mat *mat) {
    open_stream();
    try { ingest(nbins, mat); } catch (...) { close_stream(); throw; }
}

Model::Model(const fb_reads *reads, int *nbins_out) {
    open_stream();
    try { ingest_reads(reads, nbins_out); } catch (...) { close_stream(); throw; }
}

// the per-handle work buffers that do not depend on how the matrix arrived
void Model::alloc_work() {
    const int64_t n = t_.n, D = t_.D, N = D * t_.ncells;
    phi_.alloc(N); phi_new_.alloc(N); phihat_.alloc(N); wgt_.alloc(N);
    phi_.zero(st_); phihat_.zero(st_); wgt_.zero(st_);
    alloc_tables();
    prec_h_.alloc(4);
    part_.alloc(leg_part_doubles(t_));
    scratch_.alloc(4096);
    keys_.alloc((size_t)(2 * n + 1) * D + 8);
    FB_CUDA(cudaStreamSynchronize(st_));
}

// device / pinned leg tables [bias part | decay part | exposures] and the per-item sums of one leg pass
void Model::alloc_tables() {
    const int64_t n = t_.n, D = t_.D;
    if (!generic_) { tab_b_ = (int)n; tab_d_ = (int)n; }
    const size_t tn = (size_t)tab_b_ + tab_d_ + D;
    tab_.alloc(tn); tab_old_.alloc(tn); tab_sig_.alloc(tn); tab_h_.alloc(tn);
    const size_t items = 2 * (size_t)std::max<int64_t>(std::max<int64_t>(n, D), std::max(tab_b_, tab_d_));
    items_.alloc(items); items_h_.alloc(items);
    items_g_.alloc((size_t)LEG_GROUPS * items);
}

Model::~Model() { close_stream(); }

void Model::ingest(int nbins, const fb_mat *m) {
    if (!m || m->N <= 0 || nbins <= 0) throw ArgError("empty input");
    if (!m->pos1 || !m->pos2 || !m->observed || !m->nobs || !m->distance) throw ArgError("missing input column");
    if (nbins > 65535) throw ArgError("at most 65535 bins are supported");
    const int64_t N = m->N, n = nbins, ncells = n * (n + 1) / 2;
    const int nthr = host_threads();
    auto par_rows = [&](int64_t D, auto &&fn) { // fn(d, i) -> error string or nullptr, over all rows of all datasets
        std::vector<std::future<const char *>> fut;
        for (int tid = 0; tid < nthr; ++tid)
            fut.push_back(std::async(std::launch::async, [&, tid]() -> const char * {
                for (int64_t q = tid; q < D * n; q += nthr)
                    if (const char *e = fn(q / n, q % n)) return e;
                return nullptr;
            }));
        const char *err = nullptr;
        for (auto &f : fut) { const char *e = f.get(); if (e && !err) err = e; }
        return err;
    };
    int64_t D;
    if (m->name) {
        std::vector<std::future<std::pair<int, int>>> fut;
        for (int tid = 0; tid < nthr; ++tid)
            fut.push_back(std::async(std::launch::async, [&, tid]() {
                int mn = INT32_MAX, mx = INT32_MIN;
                const int64_t lo = N * tid / nthr, hi = N * (tid + 1) / nthr;
                for (int64_t k = lo; k < hi; ++k) { mn = std::min(mn, m->name[k]); mx = std::max(mx, m->name[k]); }
                return std::make_pair(mn, mx);
            }));
        int mn = INT32_MAX, mx = INT32_MIN;
        for (auto &f : fut) { auto r = f.get(); mn = std::min(mn, r.first); mx = std::max(mx, r.second); }
        D = mx;
        if (D <= 0 || D * ncells != N) throw ArgError(kSizeMsg);
        if (mn != 1) throw ArgError("Name column values should start at 1");
    } else {
        if (N % ncells != 0) throw ArgError(kSizeMsg);
        D = N / ncells;
    }
    t_.n = nbins; t_.D = (int)D; t_.ncells = ncells;

    // start the uploads, validate on the host meanwhile
    const auto ti0 = std::chrono::steady_clock::now();
    auto since = [&] { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - ti0).count(); };
    obs_.alloc(N); nobs_.alloc(N);
    auto checks = std::async(std::launch::async, [&]() -> const char * {
        if (m->bin1 || m->bin2 || m->name) {
            // FastData.ipp:22-25 asks for bins in [1, nbins]; this library additionally requires the
            // (name, bin1, bin2) sort order that the reference silently relies on (fast_binless.ipp:28-30)
            std::atomic<bool> range_bad{false};
            const char *e = par_rows(D, [&](int64_t d, int64_t i) -> const char * {
                const int64_t base = d * ncells + rm_start(i, n);
                // the usual case, a row at a time without branches; a row that fails is walked again for the message
                if (row_is_ordered(m->name ? m->name + base : nullptr, m->bin1 ? m->bin1 + base : nullptr,
                                   m->bin2 ? m->bin2 + base : nullptr, (int32_t)(d + 1), (int32_t)(i + 1), n - i))
                    return nullptr;
                for (int64_t j = i; j < n; ++j) {
                    const int64_t g = base + (j - i);
                    if (m->bin1 && (m->bin1[g] < 1 || m->bin1[g] > n)) { range_bad = true; return "range"; }
                    if (m->bin2 && (m->bin2[g] < 1 || m->bin2[g] > n)) { range_bad = true; return "range"; }
                    if ((m->name && m->name[g] != d + 1) || (m->bin1 && m->bin1[g] != i + 1) || (m->bin2 && m->bin2[g] != j + 1))
                        return "order";
                }
                return nullptr;
            });
            if (e) return range_bad.load() ? "bins should range from 1 to the number of bins"
                                    : "mat must be sorted by (name, bin1, bin2) and hold every cell of the upper triangle";
        }
        pos_.resize(n); dist_.resize(n);
        std::vector<int32_t> ipos(n);
        for (int64_t r = 0; r < n; ++r) { ipos[r] = m->pos1[rm_start(r, n)]; pos_[r] = ipos[r]; dist_[r] = m->distance[r]; }
        return par_rows(D, [&](int64_t d, int64_t i) -> const char * {
            // a row at a time, without branches (the compiler vectorises it): pos1 is the row's position, pos2 the
            // positions of the bins from i on, distance the per-diagonal table (NaN compares unequal: not a layout)
            return row_is_layout(m->pos1 + d * ncells + rm_start(i, n), m->pos2 + d * ncells + rm_start(i, n),
                                 m->distance + d * ncells + rm_start(i, n), ipos[i], ipos.data() + i, dist_.data(), n - i)
                       ? nullptr : kLayoutMsg;
        });
    });
    upload_bytes(obs_.p, m->observed, N * sizeof(int32_t));
    upload_bytes(nobs_.p, m->nobs, N * sizeof(int32_t));
    { // same buffers as alloc_work(), without its synchronisation: the validation below overlaps the uploads
        phi_.alloc(N); phi_new_.alloc(N); phihat_.alloc(N); wgt_.alloc(N);
        phi_.zero(st_); phihat_.zero(st_); wgt_.zero(st_);
        prec_h_.alloc(4);
        part_.alloc(leg_part_doubles(t_));
        scratch_.alloc(4096);
        keys_.alloc((size_t)(2 * n + 1) * D + 8);
    }
    const double t_enq = since();
    const char *e = checks.get();
    const double t_chk = since();
    FB_CUDA(cudaStreamSynchronize(st_));
    if (g_pt.on) std::fprintf(stderr, "[fb timing] ingest: uploads enqueued %.2f ms, host validation done %.2f ms, uploads done %.2f ms\n", t_enq, t_chk, since());
    if (e == kLayoutMsg || std::getenv("FB_FORCE_CELL_BINS")) {
        // not a bin_data() layout (or the test aid asks for it): bin every cell like the reference does (util.cpp:56-111)
        if (e && e != kLayoutMsg) throw ArgError(e);
        ingest_cells(m);
    } else {
        if (e) throw ArgError(e);
        for (int64_t k = 0; k < n; ++k)
            if (!(dist_[k] > 0)) throw ArgError("distance must be positive");
    }
    alloc_tables();
}

namespace {
template <class F>
void par_for(int64_t N, F &&fn) {   // fn(lo, hi) on host_threads() contiguous ranges
    const int nthr = host_threads();
    std::vector<std::thread> th;
    for (int t = 1; t < nthr; ++t) th.emplace_back([&, t] { fn(N * t / nthr, N * (t + 1) / nthr); });
    fn(0, N / nthr);
    for (auto &x : th) x.join();
}
// drop = true (util.cpp:83-98): bins nobody uses are removed; returns the number kept
int drop_unused(std::vector<int32_t> &raw, unsigned Nbins) {
    std::vector<char> used(Nbins, 0);
    for (int32_t r : raw) used[r] = 1;
    std::vector<int32_t> remap(Nbins, -1);
    int kept = 0;
    for (unsigned b = 0; b < Nbins; ++b) if (used[b]) remap[b] = kept++;
    par_for((int64_t)raw.size(), [&](int64_t lo, int64_t hi) { for (int64_t k = lo; k < hi; ++k) raw[k] = remap[raw[k]]; });
    return kept;
}
} // namespace

// Per-cell binners of the bias and decay legs (fast_bias_mean.hpp:37-68, fast_decay.hpp:48-68 over util.cpp:56-111),
// evaluated for every cell with the reference's double expressions; the cell lists per leg bin feed leg_pass_cells.
void Model::ingest_cells(const fb_mat *m) {
    const int64_t N = m->N;
    if (N >= (int64_t(1) << 31)) throw ArgError("the per-cell binner fallback supports fewer than 2^31 cells");
    generic_ = true;
    const int n = t_.n;
    // bias: even bins between the smallest and the largest position of either end
    double mn = INFINITY, mx = -INFINITY;
    for (int64_t k = 0; k < N; ++k) {
        mn = std::min(mn, std::min((double)m->pos1[k], (double)m->pos2[k]));
        mx = std::max(mx, std::max((double)m->pos1[k], (double)m->pos2[k]));
    }
    std::vector<int32_t> b1(N), b2(N), db(N);
    {
        const EvenBins eb(mn, mx, (unsigned)n);
        par_for(N, [&](int64_t lo, int64_t hi) {
            for (int64_t k = lo; k < hi; ++k) { b1[k] = eb.bin((double)m->pos1[k]); b2[k] = eb.bin((double)m->pos2[k]); }
        });
        const int n1 = drop_unused(b1, (unsigned)n), n2 = drop_unused(b2, (unsigned)n);
        if (n1 != n2) throw ArgError("bias binner: pos1 and pos2 do not populate the same bins");
        nb_bias_ = n1;
    }
    // decay: K * bins_per_bf even bins of log(distance); support = mean log-distance per bin (ascending cell order)
    std::vector<double> ld(N);
    for (int64_t k = 0; k < N; ++k) {
        if (!(m->distance[k] > 0)) throw ArgError("distance must be positive");
        ld[k] = std::log(m->distance[k]);
    }
    dec_smin_ = *std::min_element(ld.begin(), ld.end()); dec_smax_ = *std::max_element(ld.begin(), ld.end());
    {
        const EvenBins eb(dec_smin_, dec_smax_, (unsigned)(kDecayK * kDecayBinsPerBf));
        par_for(N, [&](int64_t lo, int64_t hi) { for (int64_t k = lo; k < hi; ++k) db[k] = eb.bin(ld[k]); });
        nb_dec_ = drop_unused(db, (unsigned)(kDecayK * kDecayBinsPerBf));
    }
    std::vector<double> num(nb_dec_, 0.0), cnt(nb_dec_, 0.0);
    for (int64_t k = 0; k < N; ++k) { num[db[k]] += ld[k]; cnt[db[k]] += 1.0; }
    dec_support_.resize(nb_dec_);
    for (int b = 0; b < nb_dec_; ++b) dec_support_[b] = num[b] / cnt[b];
    // cell lists per bin
    std::vector<int64_t> boff(nb_bias_ + 1, 0), doff(nb_dec_ + 1, 0);
    for (int64_t k = 0; k < N; ++k) { boff[b1[k] + 1] += 1; if (b2[k] != b1[k]) boff[b2[k] + 1] += 1; doff[db[k] + 1] += 1; }
    for (int b = 0; b < nb_bias_; ++b) boff[b + 1] += boff[b];
    for (int b = 0; b < nb_dec_; ++b) doff[b + 1] += doff[b];
    std::vector<uint32_t> blist(boff[nb_bias_]), dlist(doff[nb_dec_]);
    {
        std::vector<int64_t> bp(boff.begin(), boff.end() - 1), dp(doff.begin(), doff.end() - 1);
        for (int64_t k = 0; k < N; ++k) {
            if (b1[k] == b2[k]) blist[bp[b1[k]]++] = (uint32_t)k | 0x80000000u;
            else { blist[bp[b1[k]]++] = (uint32_t)k; blist[bp[b2[k]]++] = (uint32_t)k; }
            dlist[dp[db[k]]++] = (uint32_t)k;
        }
    }
    cb1_.alloc(N); cb2_.alloc(N); cdb_.alloc(N);
    bias_off_.alloc(boff.size()); dec_off_.alloc(doff.size());
    bias_list_.alloc(std::max<size_t>(1, blist.size())); dec_list_.alloc(std::max<size_t>(1, dlist.size()));
    cell_a_.alloc(N); cell_b_.alloc(N);
    FB_CUDA(cudaMemcpy(cb1_.p, b1.data(), N * sizeof(int32_t), cudaMemcpyHostToDevice));
    FB_CUDA(cudaMemcpy(cb2_.p, b2.data(), N * sizeof(int32_t), cudaMemcpyHostToDevice));
    FB_CUDA(cudaMemcpy(cdb_.p, db.data(), N * sizeof(int32_t), cudaMemcpyHostToDevice));
    FB_CUDA(cudaMemcpy(bias_off_.p, boff.data(), boff.size() * sizeof(int64_t), cudaMemcpyHostToDevice));
    FB_CUDA(cudaMemcpy(dec_off_.p, doff.data(), doff.size() * sizeof(int64_t), cudaMemcpyHostToDevice));
    FB_CUDA(cudaMemcpy(bias_list_.p, blist.data(), blist.size() * sizeof(uint32_t), cudaMemcpyHostToDevice));
    FB_CUDA(cudaMemcpy(dec_list_.p, dlist.data(), dlist.size() * sizeof(uint32_t), cudaMemcpyHostToDevice));
    tab_b_ = nb_bias_; tab_d_ = nb_dec_;
}

// ---------------------------------------------------------------------------------------------
// legs
// ---------------------------------------------------------------------------------------------
// The binners and the observables per leg bin are functions of the matrix alone (SummarizerSettingsImpl constructors,
// fast_exposure.hpp:35-51, fast_bias_mean.hpp:37-68, fast_decay.hpp:48-68): built on the first call of a handle and kept
// with it, like the uploaded matrix. What depends on the call (the GAM settings, the zeroed estimates) is redone.
void Model::build_legs(double tol_val, unsigned free_decay) {
    const int n = t_.n, D = t_.D;
    if (!struct_built_) {
        if (!generic_) {
            // bias binner (fast_bias_mean.hpp:37-68): even bins between min and max position
            const double mn = *std::min_element(pos_.begin(), pos_.end()), mx = *std::max_element(pos_.begin(), pos_.end());
            rowbin_ = even_bins_dropped(pos_, mn, mx, (unsigned)n, nb_bias_);
            // decay binner (fast_decay.hpp:48-68): K * bins_per_bf even bins of log(distance), empty ones dropped
            std::vector<double> ld(n);
            for (int k = 0; k < n; ++k) ld[k] = std::log(dist_[k]);
            dec_smin_ = *std::min_element(ld.begin(), ld.end()); dec_smax_ = *std::max_element(ld.begin(), ld.end());
            dbin_ = even_bins_dropped(ld, dec_smin_, dec_smax_, (unsigned)(kDecayK * kDecayBinsPerBf), nb_dec_);
            std::vector<double> sup_num(nb_dec_, 0.0), sup_den(nb_dec_, 0.0);
            for (int k = 0; k < n; ++k) { // mean log-distance per bin over all cells: D*(n-k) cells on diagonal k
                const double cnt = (double)D * (double)(n - k);
                sup_num[dbin_[k]] += cnt * ld[k]; sup_den[dbin_[k]] += cnt;
            }
            dec_support_.resize(nb_dec_);
            for (int b = 0; b < nb_dec_; ++b) dec_support_[b] = sup_num[b] / sup_den[b];
        }   // else: the binners were evaluated cell by cell at ingest (ingest_cells)
        // observables per leg bin (integers: exact in any order)
        std::vector<double> dummy;
        leg_sums(LEG_EXPO, MODE_NOBS, expo_nobs_, dummy);
        leg_sums(LEG_DECAY, MODE_NOBS, dec_nobs_, dummy);
        leg_sums(LEG_BIAS, MODE_NOBS, bias_nobs_, dummy);
        for (auto &v : bias_nobs_) v = v / 2.; // fast_bias_mean.hpp:58
        struct_built_ = true;
    }
    dec_.setup(dec_support_, dec_smin_, dec_smax_, tol_val, free_decay);
    std::fill(tab_h_.p, tab_h_.p + tab_h_.n, 0.0);
    FB_CUDA(cudaMemcpyAsync(tab_.p, tab_h_.p, tab_h_.n * sizeof(double), cudaMemcpyHostToDevice, st_));
    expo_est_.assign(D, 0.0);
    bias_est_.assign(nb_bias_, 0.0);
    legs_built_ = true;
}

void Model::upload_tables() {
    const int n = t_.n, D = t_.D;
    const std::vector<double> dc = dec_.binned_estimate();
    double *h = tab_h_.p;
    if (generic_) {
        for (int b = 0; b < nb_bias_; ++b) h[b] = bias_est_[b];
        for (int k = 0; k < nb_dec_; ++k) h[tab_b_ + k] = dc[k];
    } else {
        for (int r = 0; r < n; ++r) h[r] = bias_est_[rowbin_[r]];       // Params<Mean>::mean stays 0
        for (int k = 0; k < n; ++k) h[n + k] = dc[dbin_[k]];
    }
    for (int d = 0; d < D; ++d) h[(size_t)tab_b_ + tab_d_ + d] = expo_est_[d];
    FB_CUDA(cudaMemcpyAsync(tab_.p, h, tab_h_.n * sizeof(double), cudaMemcpyHostToDevice, st_));
    FB_CUDA(cudaStreamSynchronize(st_)); // tab_h_ is reused by the next fit
}

void Model::leg_sums(int leg, int mode, std::vector<double> &a, std::vector<double> &b) {
    if (generic_) {
        const int nb = leg == LEG_EXPO ? t_.D : (leg == LEG_BIAS ? nb_bias_ : nb_dec_);
        leg_pass_cells(mode, t_, tables(tab_), alpha_, obs_.p, nobs_.p, phi_.p, nb,
                       leg == LEG_EXPO ? nullptr : (leg == LEG_BIAS ? bias_off_.p : dec_off_.p),
                       leg == LEG_EXPO ? nullptr : (leg == LEG_BIAS ? bias_list_.p : dec_list_.p), cell_a_.p, cell_b_.p, items_.p, st_);
        kernel_launches += 2;
        FB_CUDA(cudaMemcpyAsync(items_h_.p, items_.p, 2 * (size_t)nb * sizeof(double), cudaMemcpyDeviceToHost, st_));
        FB_CUDA(cudaStreamSynchronize(st_));
        a.resize(nb); b.resize(nb);
        for (int k = 0; k < nb; ++k) { a[k] = items_h_.p[2 * k]; b[k] = items_h_.p[2 * k + 1]; }
        return;
    }
    // IRLS legs shared between the GPUs of the communicator: every rank sums the tile rows of its groups, the group sums
    // are gathered (one all-gather: a rank's groups lie side by side), and everybody adds the groups in the same order: the
    // result has the bits of the single-GPU sum (fb_irls.cuh). Needs a world size that divides LEG_GROUPS.
    Comm &cm = Comm::get();
    const bool shared = cm.active() && LEG_GROUPS % cm.world() == 0 && !std::getenv("FB_IRLS_REPLICATED");
    const int nit = leg == LEG_EXPO ? t_.D : t_.n;
    leg_pass_groups(leg, mode, t_, tables(tab_), alpha_, obs_.p, nobs_.p, phi_.p, part_.p, items_g_.p, shared ? cm.rank() : 0,
                    shared ? cm.world() : 1, st_);
    if (shared) cm.allgather(items_g_.p, (size_t)(LEG_GROUPS / cm.world()) * 2 * nit, st_);   // a rank's groups lie side by side
    leg_pass_combine(nit, items_g_.p, items_.p, shared ? cm.world() : 1, st_);
    kernel_launches += 1;
    const int nitems = leg == LEG_EXPO ? t_.D : t_.n;
    FB_CUDA(cudaMemcpyAsync(items_h_.p, items_.p, 2 * (size_t)nitems * sizeof(double), cudaMemcpyDeviceToHost, st_));
    FB_CUDA(cudaStreamSynchronize(st_));
    const int nb = leg == LEG_EXPO ? t_.D : (leg == LEG_BIAS ? nb_bias_ : nb_dec_);
    a.assign(nb, 0.0); b.assign(nb, 0.0);
    for (int k = 0; k < nitems; ++k) {
        const int bin = leg == LEG_EXPO ? k : (leg == LEG_BIAS ? rowbin_[k] : dbin_[k]);
        a[bin] += items_h_.p[2 * k]; b[bin] += items_h_.p[2 * k + 1];
    }
}

namespace {
// Summarizer::update_summary (fast_summarizer.ipp:32-51) / set_poisson_lsq_summary (:3-30), given the per-bin sums
void summary_from_sums(int mode, const std::vector<double> &a, const std::vector<double> &b, const std::vector<double> &est,
                       const std::vector<double> &nobs, std::vector<double> &phihat, std::vector<double> &weight) {
    const size_t nb = a.size();
    phihat.resize(nb); weight.resize(nb);
    for (size_t k = 0; k < nb; ++k) {
        if (mode == MODE_RESID) { // a = sum w, b = sum z*w
            double ph = b[k] / a[k] + est[k];
            if (a[k] == 0) ph = 0;
            phihat[k] = ph; weight[k] = a[k];
        } else {                  // a = sum observed, b = sum nobs*exp(log expected); pseudocount 0.01
            double e = std::log(0.01 + a[k] / b[k]);
            if (b[k] == 0) e = 0;
            phihat[k] = e; weight[k] = std::exp(e) * nobs[k];
        }
    }
}
} // namespace

void Model::fit_expo(int mode) {
    std::vector<double> a, b, ph, wt;
    g_pt.start();
    leg_sums(LEG_EXPO, mode, a, b);
    g_pt.stop(0);
    summary_from_sums(mode, a, b, expo_est_, expo_nobs_, ph, wt);
    expo_est_ = ph; // no centring, no cap (fast_exposure.hpp:63-70)
    upload_tables();
}

void Model::fit_bias(int mode) {
    std::vector<double> a, b, ph, wt;
    g_pt.start();
    leg_sums(LEG_BIAS, mode, a, b);
    g_pt.stop(0);
    summary_from_sums(mode, a, b, bias_est_, bias_nobs_, ph, wt);
    // FitterImpl<Leg,Mean>::update_params (fast_fitter_mean.hpp:55-79): centre, then cap at 3 sd
    double num = 0, den = 0;
    for (int k = 0; k < nb_bias_; ++k) { num += bias_nobs_[k] * ph[k]; den += bias_nobs_[k]; }
    const double avg = num / den;
    for (auto &v : ph) v = v - avg;
    double sq = 0;
    for (double v : ph) sq += v * v;
    const double sd = std::sqrt(sq / ph.size());
    for (auto &v : ph) v = std::max(std::min(v, 3 * sd), -3 * sd);
    bias_est_ = ph;
    upload_tables();
}

void Model::fit_decay(int mode) {
    std::vector<double> a, b, ph, wt;
    g_pt.start();
    leg_sums(LEG_DECAY, mode, a, b);
    g_pt.stop(0);
    g_pt.start();
    summary_from_sums(mode, a, b, dec_.binned_estimate(), dec_nobs_, ph, wt);
    dec_.fit(ph, wt, dec_nobs_);
    g_pt.stop(1);
    upload_tables();
}

// Idle single-GPU solver objects of destroyed handles, kept for the next handle of the same shape on the same device:
// fast_binless is called once per locus on same-shaped matrices, and building the object (chain batches of the active
// set, tables, the captured ADMM graph) costs 1.7 ms at 2400 bins against 25 ms for the whole call. The key holds the
// values of the environment switches the constructor reads, so test variants never meet each other's objects.
// fb_release_cached_memory() empties the pool. Never destroyed at exit (the CUDA context may be gone by then).
namespace {
struct SolverPool {
    struct Item { int dev, n; unsigned maxdiag; std::string env; std::unique_ptr<GflSolver> s; };
    std::mutex mu;
    std::vector<Item> items;
    static constexpr size_t kMax = 2;
    static std::string env_sig() {
        static const char *names[] = {"FB_ASET_FROM", "FB_CTL_FOLD", "FB_NO_GRAPH", "FB_PROX_ASET", "FB_PROX_GS", "FB_PROX_RTAB",
                                      "FB_PROX_TRACE", "FB_ASET_CAP", "FB_ASET_MAXPASS", "FB_ASET_STATS", "FB_SHARD_TIMING"};
        std::string sig;
        for (const char *nm : names) { const char *v = std::getenv(nm); sig += v ? v : "~"; sig += '|'; }
        return sig;
    }
    std::unique_ptr<GflSolver> take(int dev, int n, unsigned maxdiag) {
        const std::string sig = env_sig();
        std::lock_guard<std::mutex> lock(mu);
        for (size_t k = 0; k < items.size(); ++k)
            if (items[k].dev == dev && items[k].n == n && items[k].maxdiag == maxdiag && items[k].env == sig) {
                std::unique_ptr<GflSolver> s = std::move(items[k].s);
                items.erase(items.begin() + (long)k);
                return s;
            }
        return nullptr;
    }
    void give(int dev, int n, unsigned maxdiag, std::unique_ptr<GflSolver> s) {
        std::unique_ptr<GflSolver> drop;        // destroyed outside the lock
        {
            std::lock_guard<std::mutex> lock(mu);
            if (items.size() >= kMax) { drop = std::move(items.front().s); items.erase(items.begin()); }
            items.push_back(Item{dev, n, maxdiag, env_sig(), std::move(s)});
        }
    }
    void clear() {
        std::vector<Item> all;
        { std::lock_guard<std::mutex> lock(mu); all.swap(items); }
        for (auto &it : all) {       // a solver frees its buffers into the cache of ITS device
            int cur = 0; cudaGetDevice(&cur);
            cudaSetDevice(it.dev); it.s.reset(); cudaSetDevice(cur);
        }
    }
};
SolverPool &solver_pool() {
    static SolverPool *p = [] { BlockCache::pressure_hook().store(&release_idle_solvers); return new SolverPool(); }();
    return *p;
}
} // namespace

void release_idle_solvers() { solver_pool().clear(); }

// the handle's solver goes to the pool (idle, its stream drained) or, when it is shared between GPUs, away
void Model::retire_solver() {
    if (!flo_) return;
    if (!flo_->sharded() && st_) {
        if (cudaStreamSynchronize(st_) == cudaSuccess) {
            flo_->rebind(nullptr);
            solver_pool().give(dev_, t_.n, flo_maxdiag_, std::move(flo_));
        }
    }
    flo_.reset();
}

// The fused-lasso object of a call: a FRESH FusedLassoGaussianEstimator<GFLLibrary> in the reference (fast_binless.cpp:122).
// Its structure (chain batches, the captured ADMM graph, the device arrays) depends on (nbins, maxdiag) only, so the handle
// keeps one (or takes an idle one of that shape from the pool) and hands it out reset to the cold state
// (beta = z = u = 0, alpha = 5, counters 0).
GflSolver &Model::solver(unsigned maxdiag) {
    if (!flo_ || flo_maxdiag_ != maxdiag || flo_comm_ != (Comm::get().active() ? Comm::get().world() : 1)) {
        retire_solver();
        if (!Comm::get().active()) flo_ = solver_pool().take(dev_, t_.n, maxdiag);
        if (flo_) { flo_->rebind(st_); flo_->reset(); }
        else flo_.reset(new GflSolver(t_.n, maxdiag, st_));
        flo_maxdiag_ = maxdiag; flo_comm_ = Comm::get().active() ? Comm::get().world() : 1;
    } else flo_->reset();
    flo_->kernel_launches = 0;
    return *flo_;
}

// the GPUs of the communicator share the reduction passes (fb_irls.cuh: Share); FB_IRLS_REPLICATED=1 keeps them replicated
Share Model::share() const {
    Share s;
    Comm &c = Comm::get();
    if (c.active() && !std::getenv("FB_IRLS_REPLICATED")) { s.rank = c.rank(); s.world = c.world(); s.cm = &c; }
    return s;
}

void Model::check_interrupt() {
    if (interrupt_cb && interrupt_cb(interrupt_arg)) throw Interrupted("interrupted");
}

// Device -> caller's host buffer. Pinned destinations (bench.py, the Python harness) take one DMA. R hands freshly
// allocated PAGEABLE vectors: there the driver's own staging runs at a few GB/s (10 GB of log_signal at 50 000 bins:
// 2.5 s, most of it first-touch page faults on one thread), so large pageable copies are pipelined through pinned
// staging buffers and scattered to the destination by several host threads while the next chunk is in flight.
void Model::download_bytes(void *h_dst, const void *d_src, size_t bytes, cudaStream_t st) {
    if (!h_dst || !bytes) return;
    if (!st) st = st_;
    constexpr size_t CHUNK = size_t(32) << 20;
    cudaPointerAttributes at{};
    const bool known = cudaPointerGetAttributes(&at, h_dst) == cudaSuccess;
    if (!known) cudaGetLastError();
    const bool pageable = !known || at.type == cudaMemoryTypeUnregistered;
    if (!pageable || bytes < stage_min_bytes()) {
        FB_CUDA(cudaMemcpyAsync(h_dst, d_src, bytes, cudaMemcpyDeviceToHost, st));
        return;
    }
    constexpr int NB = 3;
    void *stage[NB]; cudaEvent_t ev[NB]; std::future<void> busy[NB];
    for (int b = 0; b < NB; ++b) { stage[b] = BlockCache::host().get(CHUNK); FB_CUDA(cudaEventCreateWithFlags(&ev[b], cudaEventDisableTiming)); }
    const int nthr = stage_threads();
    size_t off = 0;
    for (int i = 0; off < bytes; ++i, off += CHUNK) {
        const int b = i % NB;
        const size_t len = std::min(CHUNK, bytes - off);
        if (busy[b].valid()) busy[b].get();                       // the staging buffer has been emptied
        FB_CUDA(cudaMemcpyAsync(stage[b], static_cast<const char *>(d_src) + off, len, cudaMemcpyDeviceToHost, st));
        FB_CUDA(cudaEventRecord(ev[b], st));
        char *dst = static_cast<char *>(h_dst) + off; const char *src = static_cast<const char *>(stage[b]); cudaEvent_t e = ev[b];
        busy[b] = std::async(std::launch::async, [=] {
            cudaEventSynchronize(e);
            std::vector<std::thread> th;
            const size_t part = (len + nthr - 1) / nthr;
            for (int t = 1; t < nthr; ++t)
                if ((size_t)t * part < len) th.emplace_back([=] { std::memcpy(dst + t * part, src + t * part, std::min(part, len - t * part)); });
            std::memcpy(dst, src, std::min(part, len));
            for (auto &x : th) x.join();
        });
    }
    for (int b = 0; b < NB; ++b) { if (busy[b].valid()) busy[b].get(); cudaEventDestroy(ev[b]); BlockCache::host().put(stage[b], CHUNK); }
}

// Host -> device, the mirror image: large pageable sources are copied into pinned staging by several host threads
// and sent chunk by chunk, so the DMA of one chunk overlaps the host copy of the next.
void Model::upload_bytes(void *d_dst, const void *h_src, size_t bytes) {
    if (!bytes) return;
    constexpr size_t CHUNK = size_t(32) << 20;
    cudaPointerAttributes at{};
    const bool known = cudaPointerGetAttributes(&at, h_src) == cudaSuccess;
    if (!known) cudaGetLastError();
    const bool pageable = !known || at.type == cudaMemoryTypeUnregistered;
    if (!pageable || bytes < stage_min_bytes()) {
        FB_CUDA(cudaMemcpyAsync(d_dst, h_src, bytes, cudaMemcpyHostToDevice, st_));
        return;
    }
    constexpr int NB = 3;
    void *stage[NB]; cudaEvent_t ev[NB]; bool used[NB] = {false, false, false};
    for (int b = 0; b < NB; ++b) { stage[b] = BlockCache::host().get(CHUNK); FB_CUDA(cudaEventCreateWithFlags(&ev[b], cudaEventDisableTiming)); }
    const int nthr = stage_threads();
    size_t off = 0;
    for (int i = 0; off < bytes; ++i, off += CHUNK) {
        const int b = i % NB;
        const size_t len = std::min(CHUNK, bytes - off);
        if (used[b]) FB_CUDA(cudaEventSynchronize(ev[b]));       // the DMA out of this staging buffer has finished
        char *dst = static_cast<char *>(stage[b]); const char *src = static_cast<const char *>(h_src) + off;
        std::vector<std::thread> th;
        const size_t part = (len + nthr - 1) / nthr;
        for (int t = 1; t < nthr; ++t)
            if ((size_t)t * part < len) th.emplace_back([=] { std::memcpy(dst + t * part, src + t * part, std::min(part, len - t * part)); });
        std::memcpy(dst, src, std::min(part, len));
        for (auto &x : th) x.join();
        FB_CUDA(cudaMemcpyAsync(static_cast<char *>(d_dst) + off, stage[b], len, cudaMemcpyHostToDevice, st_));
        FB_CUDA(cudaEventRecord(ev[b], st_));
        used[b] = true;
    }
    FB_CUDA(cudaStreamSynchronize(st_));
    for (int b = 0; b < NB; ++b) { cudaEventDestroy(ev[b]); BlockCache::host().put(stage[b], CHUNK); }
}

void Model::download(double *h_dst, const double *d_src, size_t cnt) { download_bytes(h_dst, d_src, cnt * sizeof(double)); }

void Model::patchnos(const double *d_value, double tol, bool compute, int32_t *h_out) {
    if (!h_out) return;
    const int64_t N = t_.N(), nc = t_.ncells;
    DevBuf<int32_t> labels(N);
    if (!compute) { fill_i32(labels.p, N, -1, st_); kernel_launches += 1; }
    else {
        DevBuf<int32_t> parent(nc), scan(nc + nc / 4096 + 8);
        for (int d = 0; d < t_.D; ++d) {
            patch_labels(t_.n, tol, d_value + (size_t)d * nc, parent.p, scan.p, labels.p + (size_t)d * nc, st_);
            kernel_launches += 6;
        }
        FB_CUDA(cudaStreamSynchronize(st_));
    }
    download_bytes(h_out, labels.p, N * sizeof(int32_t));
    FB_CUDA(cudaStreamSynchronize(st_));
}

// ---------------------------------------------------------------------------------------------
// binless() (fast_binless.cpp:85-184)
// ---------------------------------------------------------------------------------------------
void Model::run_signal(const fb_params *p, fb_signal_out *out) {
    if (!p || !out) throw ArgError("null argument");
    if (!p->lam2 || p->n_lam2 <= 0) throw ArgError("lam2 is required");
    const int n = t_.n, D = t_.D;
    const int64_t nc = t_.ncells, N = t_.N();
    FB_CUDA(cudaEventRecord(ev0_, st_));
    double ms_solver = 0;
    kernel_launches = 0;
    if (p->verbose) std::printf("init\n");
    alpha_ = p->alpha;
    const double tol_val = p->tol_val;
    unsigned bg_steps = p->bg_steps;
    phi_.zero(st_); phihat_.zero(st_); wgt_.zero(st_);
    g_pt.start();
    build_legs(tol_val, p->free_decay);
    g_pt.stop(3);
    // initial Poisson fits, in this order (fast_binless.cpp:104-115)
    fit_expo(MODE_POISSON);
    fit_decay(MODE_POISSON);
    fit_bias(MODE_POISSON);

    double current_tol_val = 1.;
    const auto tsolver0 = std::chrono::steady_clock::now();
    GflSolver &flo = solver(p->maxdiag); // shared by all datasets: the flos copies share one strategy (GFLLibrary.hpp:59)
    if (g_pt.on)
        std::fprintf(stderr, "[fb timing] fused-lasso solver object (plans, tables) %.2f ms\n",
                     std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - tsolver0).count());
    double flo_tol = current_tol_val / 20.;
    const std::vector<double> lam2 = broadcast(p->lam2, p->n_lam2, D);
    const std::vector<double> lam1 = broadcast(p->lam1, p->n_lam1, D);
    int64_t n_solves = 0, iters = 0, cell_iters = 0;
    bool sig_partial = false;
    out->nsteps = 0; out->converged = 0;
    // get_as_dataframe columns (fast_dataframe.cpp:9-110). Four of them are final BEFORE the fused-lasso solves of the step
    // that turns out to be the last one: phihat and weight (made by signal_yw for the solves) and biasmat / decaymat (the
    // bias and decay legs are not touched after it; fit_expo changes the exposures only). A signal step that CAN end the
    // loop (the tolerance test below needs current_tol_val <= 1.01 tol_val before the step, or it is step nouter) sends
    // them to the caller's buffers on a second stream while the solves run (PCIe is idle then); the copies of the step
    // that ends the loop stand, those of an earlier step are overwritten (by a later early copy on the same stream, or by
    // the downloads at the end, which wait for that stream). A step that finds the previous early copy still in flight
    // does not add another. Only for pinned host buffers: a copy to pageable memory would hold this thread.
    // FB_NO_EARLY_DOWNLOAD=1 keeps everything for the end.
    DevBuf<double> lam1_d(D), cols;
    DevBuf<int32_t> pl_labels, pl_parent, pl_scan;     // patch numbers made on the second stream (below)
    lam1_d.upload(lam1.data(), D, st_);
    const bool csv = p->csv_out && p->csv_out[0];
    double *want[5] = {out->background, out->biasmat, out->decaymat, out->signal, out->binless};
    double *dcol[5] = {nullptr, nullptr, nullptr, nullptr, nullptr};
    {
        int nwant = 0;
        for (double *w : want) nwant += (w != nullptr || csv);
        if (nwant) {
            cols.alloc((size_t)nwant * N);
            int k = 0;
            for (int c = 0; c < 5; ++c) if (want[c] || csv) dcol[c] = cols.p + (size_t)(k++) * N;
        }
    }
    double *early_h[4] = {out->phihat, out->weight, out->biasmat, out->decaymat};
    bool early = !std::getenv("FB_NO_EARLY_DOWNLOAD") && (early_h[0] || early_h[1] || early_h[2] || early_h[3]);
    // pinned buffers take plain asynchronous copies; pageable ones (R vectors) the staged copy of download_bytes, run by a
    // second host thread while this one drives the solves; a mix of the two gets no early copies
    int n_pinned = 0, n_pageable = 0;
    for (double *h : early_h) if (h) { if (host_pinned(h)) ++n_pinned; else ++n_pageable; }
    if (n_pinned && n_pageable) early = false;
    const bool early_staged = early && n_pageable > 0;
    if (early_staged && (size_t)N * sizeof(double) < stage_min_bytes()) early = false;
    unsigned early_step = 0;                 // the signal step whose early copies are in the caller's buffers (0: none)
    struct DlGuard { cudaStream_t s; ~DlGuard() { if (s) cudaStreamSynchronize(s); } } dl_guard{early ? st_dl_ : nullptr};
    std::future<void> early_task;            // early_staged: joins in its destructor, before the buffers above go
    for (unsigned step = 1; step <= p->nouter; ++step) {
        check_interrupt();
        if (p->verbose) std::printf("step %u", step);
        fit_bias(MODE_RESID);
        const bool bg = step <= bg_steps;
        if (bg) FB_CUDA(cudaMemcpyAsync(tab_old_.p, tab_.p, tab_.n * sizeof(double), cudaMemcpyDeviceToDevice, st_));
        fit_decay(MODE_RESID);
        if (bg) {
            precision(t_, tables(tab_old_), tables(tab_), phi_.p, phi_.p, scratch_.p, scratch_.p + 8, st_, share());
            kernel_launches += 3;
        } else {
            // band solve shared between GPUs: (phihat, w) beyond the band only on this rank's rows; the full columns are
            // rebuilt after the loop from the inputs of the last signal step (phi_new_ after the swap, the tables saved here)
            const Share sh = share();
            const bool partial = sh.on() && flo.trapeze();
            signal_yw(t_, tables(tab_), alpha_, obs_.p, nobs_.p, phi_.p, 0, phihat_.p, wgt_.p, st_, partial ? sh : Share(), flo.grid().md);
            if (partial) FB_CUDA(cudaMemcpyAsync(tab_sig_.p, tab_.p, tab_.n * sizeof(double), cudaMemcpyDeviceToDevice, st_));
            sig_partial = partial;
            kernel_launches += 1;
            const bool early_idle = early_staged ? (!early_task.valid() || early_task.wait_for(std::chrono::seconds(0)) == std::future_status::ready)
                                                 : event_done(ev_dl_);
            if (early && (current_tol_val <= tol_val * 1.01 || step == p->nouter) && early_idle) {
                if (early_task.valid()) early_task.get();
                FB_CUDA(cudaEventRecord(ev_in_, st_));
                FB_CUDA(cudaStreamWaitEvent(st_dl_, ev_in_, 0));
                if (dcol[1] || dcol[2]) {
                    output_signal(t_, tables(tab_), nobs_.p, phi_.p, lam1_d.p, nullptr, dcol[1], dcol[2], nullptr, nullptr, st_dl_);
                    kernel_launches += 1;
                }
                FB_CUDA(cudaEventRecord(ev_sk_, st_dl_));
                const double *src[4] = {partial ? nullptr : phihat_.p, partial ? nullptr : wgt_.p, dcol[1], dcol[2]};
                if (!early_staged) {
                    for (int c = 0; c < 4; ++c)
                        if (early_h[c] && src[c])
                            FB_CUDA(cudaMemcpyAsync(early_h[c], src[c], (size_t)N * sizeof(double), cudaMemcpyDeviceToHost, st_dl_));
                    FB_CUDA(cudaEventRecord(ev_dl_, st_dl_));
                } else {
                    int dev = 0;
                    FB_CUDA(cudaGetDevice(&dev));
                    double *d0 = early_h[0], *d1 = early_h[1], *d2 = early_h[2], *d3 = early_h[3];
                    const double *s0 = src[0], *s1 = src[1], *s2 = src[2], *s3 = src[3];
                    early_task = std::async(std::launch::async, [this, dev, N, d0, d1, d2, d3, s0, s1, s2, s3] {
                        FB_CUDA(cudaSetDevice(dev));
                        double *dst[4] = {d0, d1, d2, d3};
                        const double *s4[4] = {s0, s1, s2, s3};
                        for (int c = 0; c < 4; ++c)
                            if (dst[c] && s4[c]) download_bytes(dst[c], s4[c], (size_t)N * sizeof(double), st_dl_);
                        FB_CUDA(cudaStreamSynchronize(st_dl_));
                    });
                }
                early_step = step;
            }
            for (int d = 0; d < D; ++d) {
                g_pt.start();
                const int steps = flo.optimize(phihat_.p + (size_t)d * nc, wgt_.p + (size_t)d * nc, lam2[d], flo_tol);
                g_pt.stop(2);
                ms_solver += flo.ms_last;
                const int it = steps - 1;
                if (out->solve_iters && n_solves < out->solve_iters_cap) out->solve_iters[n_solves] = it;
                ++n_solves; iters += it; cell_iters += (int64_t)it * flo.grid().ncells;
                if (flo.trapeze()) {   // band solve: scatter and centring in one pass (the far region is one value)
                    flo.get_beta_full_centered(phi_new_.p + (size_t)d * nc, nobs_.p + (size_t)d * nc, expo_nobs_[d]);
                } else {
                    flo.get_beta_full(phi_new_.p + (size_t)d * nc);
                    center_dataset(nc, nobs_.p + (size_t)d * nc, phi_new_.p + (size_t)d * nc, scratch_.p + 16, st_, share());
                    kernel_launches += 3;
                }
            }
            precision(t_, tables(tab_), tables(tab_), phi_.p, phi_new_.p, scratch_.p, scratch_.p + 8, st_, share());
            kernel_launches += 3;
            std::swap(phi_.p, phi_new_.p);
            if (early_step == step) FB_CUDA(cudaStreamWaitEvent(st_, ev_sk_, 0));   // fit_expo re-uploads the tables the early kernel read
        }
        FB_CUDA(cudaMemcpyAsync(prec_h_.p, scratch_.p + 8, 3 * sizeof(double), cudaMemcpyDeviceToHost, st_));
        FB_CUDA(cudaStreamSynchronize(st_));
        const double pabs = prec_h_.p[0], prel = pabs / (prec_h_.p[1] - prec_h_.p[2]);
        if (p->verbose) std::printf(" : reached precision abs = %g rel = %g\n", pabs, prel);
        if (step - 1 < FB_MAX_OUTER) { out->prec_abs[step - 1] = pabs; out->prec_rel[step - 1] = prel; }
        bool converged = prel <= tol_val && current_tol_val <= tol_val * 1.01;
        if (converged && step <= bg_steps) { converged = false; bg_steps = step; }
        current_tol_val = std::min(current_tol_val, std::max(tol_val, pabs));
        flo_tol = current_tol_val / 20;
        if (converged || step == p->nouter) { shift_signal(t_, phi_.p, keys_.p, st_, share()); kernel_launches += 3; }
        else { remove_degeneracy(t_, phi_.p, keys_.p, st_, share()); kernel_launches += 3; }
        fit_expo(MODE_RESID);
        out->nsteps = (int)step;
        if (converged) { if (p->verbose) std::printf("converged\n"); out->converged = 1; break; }
    }
    if (sig_partial) {   // the (phihat, weight) columns of the last signal step in full
        signal_yw(t_, tables(tab_sig_), alpha_, obs_.p, nobs_.p, phi_new_.p, 0, phihat_.p, wgt_.p, st_);
        kernel_launches += 1;
    }
    kernel_launches += flo.kernel_launches;
    if (p->verbose) std::printf("done\n");
    FB_CUDA(cudaEventRecord(ev1_, st_));
    if (g_pt.on) {
        std::fprintf(stderr, "[fb timing] leg passes %.2f ms, decay GAM %.2f ms, solves (host wall) %.2f ms (device %.2f), build_legs %.2f ms\n",
                     g_pt.acc[0], g_pt.acc[1], g_pt.acc[2], ms_solver, g_pt.acc[3]);
        for (double &a : g_pt.acc) a = 0;
    }

    // get_as_dataframe, Signal overload (fast_dataframe.cpp:9-110)
    {
        // the early copies stand when the loop ended on the signal step that made them
        const bool have_early = early_step != 0 && early_step == (unsigned)out->nsteps;
        const auto tail0 = std::chrono::steady_clock::now();
        const bool early_landed = early_step != 0 && (early_staged ? early_task.wait_for(std::chrono::seconds(0)) == std::future_status::ready
                                                                    : event_done(ev_dl_));
        const bool have_yw = have_early && !sig_partial;
        if (early_task.valid()) early_task.get();                          // staged early copies: all in the caller's buffers
        else if (early_step) FB_CUDA(cudaStreamWaitEvent(st_, ev_dl_, 0)); // biasmat / decaymat are written again below
        // the patch numbers (six kernels per dataset) on the second stream, under the column downloads of this one
        if (early && out->patchno && host_pinned(out->patchno)) {
            FB_CUDA(cudaEventRecord(ev_in_, st_));
            FB_CUDA(cudaStreamWaitEvent(st_dl_, ev_in_, 0));
            pl_labels.alloc(N);
            if (!p->compute_patchnos) { fill_i32(pl_labels.p, N, -1, st_dl_); kernel_launches += 1; }
            else {
                pl_parent.alloc(nc); pl_scan.alloc(nc + nc / 4096 + 8);
                for (int d = 0; d < D; ++d) {
                    patch_labels(n, tol_val, phi_.p + (size_t)d * nc, pl_parent.p, pl_scan.p, pl_labels.p + (size_t)d * nc, st_dl_);
                    kernel_launches += 6;
                }
            }
            FB_CUDA(cudaMemcpyAsync(out->patchno, pl_labels.p, (size_t)N * sizeof(int32_t), cudaMemcpyDeviceToHost, st_dl_));
        }
        if (cols.p) {
            output_signal(t_, tables(tab_), nobs_.p, phi_.p, lam1_d.p, dcol[0], dcol[1], dcol[2], dcol[3], dcol[4], st_);
            kernel_launches += 1;
            for (int c = 0; c < 5; ++c)
                if (!(have_early && (c == 1 || c == 2))) download(want[c], dcol[c], N);
        }
        if (!have_yw) {
            download(out->phihat, phihat_.p, N);
            download(out->weight, wgt_.p, N);
        }
        download(out->log_signal, phi_.p, N);
        FB_CUDA(cudaStreamSynchronize(st_));
        if (g_pt.on)
            std::fprintf(stderr, "[fb timing] columns to the host after the loop %.2f ms (early copies of step %u: %s, %s)\n",
                         std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - tail0).count(), early_step,
                         have_early ? "kept" : "not used", !early_step ? "none made" : early_landed ? "landed before the loop ended" : "still in flight");
        if (csv) { // fast_dataframe.cpp:23-57: phihat, weight, biasmat, decaymat, signal, binless, background
            std::vector<double> hbuf((size_t)7 * N);
            const double *src[7] = {phihat_.p, wgt_.p, dcol[1], dcol[2], dcol[3], dcol[4], dcol[0]};
            const double *hc[7];
            for (int c = 0; c < 7; ++c) {
                download(hbuf.data() + (size_t)c * N, src[c], N);     // pageable: staged and pipelined when large
                hc[c] = hbuf.data() + (size_t)c * N;
            }
            FB_CUDA(cudaStreamSynchronize(st_));
            write_signal_csv(p->csv_out, N, hc);
        }
    }
    {
        const auto tp0 = std::chrono::steady_clock::now();
        if (pl_labels.p) FB_CUDA(cudaStreamSynchronize(st_dl_));
        else patchnos(phi_.p, tol_val, p->compute_patchnos != 0, out->patchno);
        if (g_pt.on)
            std::fprintf(stderr, "[fb timing] patch numbers (kernels + download) %.2f ms\n",
                         std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - tp0).count());
    }
    // leg states (fast_binless.cpp:179-183)
    if (out->biases.estimate) std::memcpy(out->biases.estimate, bias_est_.data(), bias_est_.size() * sizeof(double));
    out->biases.n = nb_bias_; out->biases.mean = 0;
    for (int k = 0; k < kDecayK; ++k) out->decay.beta[k] = dec_.beta[k];
    out->decay.lambda = dec_.lambda; out->decay.mean = dec_.mean;
    if (out->exposures.estimate) std::memcpy(out->exposures.estimate, expo_est_.data(), expo_est_.size() * sizeof(double));
    out->exposures.n = D; out->exposures.mean = 0;
    out->admm_iters = iters; out->admm_cell_iters = cell_iters; out->n_solves = n_solves;
    float ms = 0;
    FB_CUDA(cudaEventElapsedTime(&ms, ev0_, ev1_));
    out->ms_total = ms; out->ms_solver = ms_solver;
    has_state_ = true;
}

// set_state calls of fast_binless.cpp:244-260
void Model::restore(const double *log_signal, const fb_mean_state *biases, const fb_gam_state *decay,
                    const fb_mean_state *exposures) {
    if (!legs_built_) build_legs(0.2, 10000);
    if (biases) {
        if (biases->n != nb_bias_ || !biases->estimate) throw ArgError("restored leg state has the wrong size");
        bias_est_.assign(biases->estimate, biases->estimate + biases->n);
        if (biases->mean != 0) for (auto &v : bias_est_) v -= biases->mean;
    }
    if (exposures) {
        if (exposures->n != t_.D || !exposures->estimate) throw ArgError("restored leg state has the wrong size");
        expo_est_.assign(exposures->estimate, exposures->estimate + exposures->n);
        if (exposures->mean != 0) for (auto &v : expo_est_) v -= exposures->mean;
    }
    if (decay) {
        dec_.beta.assign(decay->beta, decay->beta + kDecayK);
        dec_.lambda = decay->lambda; dec_.mean = decay->mean;
    }
    if (log_signal) FB_CUDA(cudaMemcpyAsync(phi_.p, log_signal, t_.N() * sizeof(double), cudaMemcpyHostToDevice, st_));
    upload_tables();
    has_state_ = true;
}

// ---------------------------------------------------------------------------------------------
// binless_difference (fast_binless.cpp:235-279) + step_difference (fast_binless.ipp:51-105)
// ---------------------------------------------------------------------------------------------
void Model::run_difference(int ref, const fb_params *p, fb_diff_out *out) {
    if (!p || !out) throw ArgError("null argument");
    if (!has_state_) throw StateError("run_difference needs a fitted or restored model");
    if (!p->lam2 || p->n_lam2 <= 0) throw ArgError("lam2 is required");
    const int n = t_.n, D = t_.D;
    const int64_t nc = t_.ncells, N = t_.N();
    if (p->verbose) std::printf("init\n");
    if (D < ref || ref < 1) throw ArgError("Ref is not in the allowed range"); // FastData.hpp:90
    alpha_ = p->alpha;
    kernel_launches = 0;
    FB_CUDA(cudaEventRecord(ev0_, st_));
    const double converge = p->tol_val / 20.;
    GflSolver &flo = solver(p->maxdiag);
    const std::vector<double> lam2 = broadcast(p->lam2, p->n_lam2, D - 1);
    if (p->verbose) std::printf("compute\n");
    DevBuf<double> phihat(N), deltahat(N), w(N), delta(N), colbuf(N);
    diff_prepare(t_, tables(tab_), alpha_, obs_.p, nobs_.p, phi_.p, ref - 1, phihat.p, deltahat.p, w.p, st_);
    kernel_launches += 1;
    delta.zero(st_);
    int64_t n_solves = 0, iters = 0, cell_iters = 0;
    double ms_solver = 0;
    for (int d = 0, idx = 0; d < D; ++d) {
        if (d == ref - 1) continue;
        const int steps = flo.optimize(deltahat.p + (size_t)d * nc, w.p + (size_t)d * nc, lam2[idx++], converge);
        ms_solver += flo.ms_last;
        const int it = steps - 1;
        if (out->solve_iters && n_solves < out->solve_iters_cap) out->solve_iters[n_solves] = it;
        ++n_solves; iters += it; cell_iters += (int64_t)it * flo.grid().ncells;
        flo.get_beta_full(delta.p + (size_t)d * nc);
    }
    kernel_launches += flo.kernel_launches;
    FB_CUDA(cudaEventRecord(ev1_, st_));
    // Difference data frame (fast_dataframe.cpp:113-145); lam1 is not applied there
    DevBuf<double> phi_ref;
    if (out->phi_ref) phi_ref.alloc(N);
    diff_finish(t_, ref - 1, phihat.p, w.p, delta.p, phi_ref.p, out->difference ? colbuf.p : nullptr, st_);
    kernel_launches += 1;
    download(out->difference, colbuf.p, N);
    download(out->deltahat, deltahat.p, N);
    download(out->weight, w.p, N);
    download(out->phi_ref, phi_ref.p, N);
    FB_CUDA(cudaStreamSynchronize(st_));
    patchnos(delta.p, p->tol_val, p->compute_patchnos != 0, out->patchno);
    if (p->verbose) std::printf("done\n");
    out->admm_iters = iters; out->admm_cell_iters = cell_iters; out->n_solves = n_solves;
    float ms = 0;
    FB_CUDA(cudaEventElapsedTime(&ms, ev0_, ev1_));
    out->ms_total = ms; out->ms_solver = ms_solver;
}

// ---------------------------------------------------------------------------------------------
// binless_eval_cv (fast_binless.cpp:186-233)
// ---------------------------------------------------------------------------------------------
void Model::eval_cv(const fb_params *p, int nlam, const double *lam2v, unsigned group, fb_cv_out *outs) {
    if (!p || !outs || !lam2v) throw ArgError("null argument");
    if (!has_state_) throw StateError("eval_cv needs a fitted or restored model");
    if (t_.D != 1) throw ArgError("Provide only 1 dataset!");
    const int n = t_.n;
    const int64_t N = t_.N();
    alpha_ = p->alpha;
    kernel_launches = 0;
    const std::vector<double> lam1 = broadcast(p->lam1, p->n_lam1, 1);
    DevBuf<double> lam1_d(1), phi_cv(N), y(N), w(N), sig(N), binl(N);
    lam1_d.upload(lam1.data(), 1, st_);
    GflSolver &flo = solver(p->maxdiag); // not reset between lambdas: warm-start chain of the reference
    for (int k = 0; k < nlam; ++k) {
        check_interrupt();
        signal_yw(t_, tables(tab_), alpha_, obs_.p, nobs_.p, phi_.p, group, y.p, w.p, st_);
        flo.optimize(y.p, w.p, lam2v[k], p->tol_val / 20.);
        flo.get_beta_full(phi_cv.p);
        center_dataset(N, nobs_.p, phi_cv.p, scratch_.p + 16, st_, share());
        fb_cv_out &o = outs[k];
        output_signal(t_, tables(tab_), nobs_.p, phi_cv.p, lam1_d.p, nullptr, nullptr, nullptr,
                      o.signal ? sig.p : nullptr, o.binless ? binl.p : nullptr, st_);
        kernel_launches += 5;
        download(o.signal, sig.p, N); download(o.binless, binl.p, N);
        download(o.phihat, y.p, N); download(o.weight, w.p, N); download(o.log_signal, phi_cv.p, N);
        FB_CUDA(cudaStreamSynchronize(st_));
        patchnos(phi_cv.p, p->tol_val, true, o.patchno);
    }
    kernel_launches += flo.kernel_launches;
}

// ---------------------------------------------------------------------------------------------
// lambda2 / BIC sweep (fastbinless.h: fb_sweep_lambda)
// ---------------------------------------------------------------------------------------------
void Model::sweep_lambda(const fb_params *p, int nlam, const double *lam2v, fb_sweep_out *out) {
    if (!p || !out || !lam2v || nlam < 0) throw ArgError("null argument");
    if (!has_state_) throw StateError("sweep_lambda needs a fitted or restored model");
    const int n = t_.n, D = t_.D;
    const int64_t nc = t_.ncells, N = t_.N();
    alpha_ = p->alpha;
    kernel_launches = 0;
    const std::vector<double> lam1 = broadcast(p->lam1, p->n_lam1, D);
    DevBuf<double> y(N), w(N), beta(nc), partial(SWEEP_BLOCKS), res(2);
    DevBuf<int32_t> parent(nc), scan(nc + nc / 4096 + 8), labels(nc), flags(nc);
    PinBuf<double> res_h; res_h.alloc(2);
    signal_yw(t_, tables(tab_), alpha_, obs_.p, nobs_.p, phi_.p, 0, y.p, w.p, st_);
    kernel_launches += 1;
    GflSolver &flo = solver(p->maxdiag);
    double ms_solver = 0;
    int64_t cell_iters = 0;
    for (int k = 0; k < nlam; ++k) {
        check_interrupt();
        for (int d = 0; d < D; ++d) {
            flo.reset();                                   // cold start: the points are independent of each other
            const int steps = flo.optimize(y.p + (size_t)d * nc, w.p + (size_t)d * nc, lam2v[k], p->tol_val / 20.);
            ms_solver += flo.ms_last;
            cell_iters += (int64_t)(steps - 1) * flo.grid().ncells;
            flo.get_beta_full(beta.p);
            center_dataset(nc, nobs_.p + (size_t)d * nc, beta.p, scratch_.p + 16, st_, share());
            patch_labels(n, p->tol_val, beta.p, parent.p, scan.p, labels.p, st_);
            sweep_score(nc, y.p + (size_t)d * nc, w.p + (size_t)d * nc, beta.p, labels.p, lam1[d], p->tol_val, flags.p,
                        partial.p, res.p, st_);
            kernel_launches += 3 + 6 + 2;
            FB_CUDA(cudaMemcpyAsync(res_h.p, res.p, 2 * sizeof(double), cudaMemcpyDeviceToHost, st_));
            FB_CUDA(cudaStreamSynchronize(st_));
            const size_t o = (size_t)k * D + d;
            const double chisq = res_h.p[0];
            const int dof = (int)res_h.p[1];
            if (out->chisq) out->chisq[o] = chisq;
            if (out->dof) out->dof[o] = dof;
            if (out->steps) out->steps[o] = steps;
            if (out->bic) out->bic[o] = chisq + std::log(expo_nobs_[d]) * dof; // ScoreAssembler<BIC>: lsnc = log(sum nobs)
        }
    }
    kernel_launches += flo.kernel_launches;
    out->ms_solver = ms_solver; out->admm_cell_iters = cell_iters;
}

// ---------------------------------------------------------------------------------------------
// checkerboard cross-validation score (fastbinless.h: fb_cv_score)
// ---------------------------------------------------------------------------------------------
void Model::cv_score(const fb_params *p, int nlam, const double *lam2v, fb_cv_score_out *out) {
    if (!p || !out || !lam2v || nlam < 0) throw ArgError("null argument");
    if (!has_state_) throw StateError("cv_score needs a fitted or restored model");
    const int n = t_.n, D = t_.D;
    const int64_t nc = t_.ncells, N = t_.N();
    alpha_ = p->alpha;
    kernel_launches = 0;
    const std::vector<double> lam1 = broadcast(p->lam1, p->n_lam1, D);
    DevBuf<double> y(N), w(N), ycv(N), wcv(N), beta(nc), partial(2 * CV_BLOCKS), res(4);
    PinBuf<double> res_h; res_h.alloc(4);
    signal_yw(t_, tables(tab_), alpha_, obs_.p, nobs_.p, phi_.p, 0, y.p, w.p, st_);
    kernel_launches += 1;
    GflSolver &flo = solver(p->maxdiag);
    double ms_solver = 0;
    int64_t cell_iters = 0;
    std::vector<int> steps((size_t)std::max(1, nlam) * D * 2, 0);
    std::vector<double> sums((size_t)std::max(1, nlam) * D * 4, 0.0);
    for (int g = 0; g < 2; ++g) {      // group-major: the held-out (y, w) of a group is built once
        signal_yw(t_, tables(tab_), alpha_, obs_.p, nobs_.p, phi_.p, (unsigned)(g + 1), ycv.p, wcv.p, st_);
        kernel_launches += 1;
        for (int k = 0; k < nlam; ++k) {
            check_interrupt();
            for (int d = 0; d < D; ++d) {
                flo.reset();               // cold start (ScorePreparator.ipp:36-37)
                const int st = flo.optimize(ycv.p + (size_t)d * nc, wcv.p + (size_t)d * nc, lam2v[k], p->tol_val / 20.);
                ms_solver += flo.ms_last;
                cell_iters += (int64_t)(st - 1) * flo.grid().ncells;
                flo.get_beta_full(beta.p);
                center_dataset(nc, nobs_.p + (size_t)d * nc, beta.p, scratch_.p + 16, st_, share());
                cv_group_sum(n, g, y.p + (size_t)d * nc, w.p + (size_t)d * nc, beta.p, lam1[d], partial.p, res.p, st_);
                kernel_launches += 1 + 3 + 2;
                FB_CUDA(cudaMemcpyAsync(res_h.p, res.p, 2 * sizeof(double), cudaMemcpyDeviceToHost, st_));
                FB_CUDA(cudaStreamSynchronize(st_));
                const size_t o = (size_t)k * D + d;
                steps[2 * o + g] = st;
                sums[4 * o + 2 * g] = res_h.p[0]; sums[4 * o + 2 * g + 1] = res_h.p[1];
            }
        }
    }
    for (size_t o = 0; o < (size_t)nlam * D; ++o) {   // ScoreAssembler<CVkSD>::assemble (ScoreAssembler.hpp:24-36)
        const double N0 = sums[4 * o + 1], N1 = sums[4 * o + 3];
        const double cv0 = sums[4 * o] / N0, cv1 = sums[4 * o + 2] / N1;
        const double Nt = N0 + N1;
        const double cv = (N0 * cv0 + N1 * cv1) / Nt;
        const double sd = std::sqrt((N0 * (cv0 - cv) * (cv0 - cv) + N1 * (cv1 - cv) * (cv1 - cv)) / (Nt - 1));
        if (out->cv) out->cv[o] = cv;
        if (out->cv_sd) out->cv_sd[o] = sd;
        if (out->steps) { out->steps[2 * o] = steps[2 * o]; out->steps[2 * o + 1] = steps[2 * o + 1]; }
    }
    kernel_launches += flo.kernel_launches;
    out->ms_solver = ms_solver; out->admm_cell_iters = cell_iters;
}

} // namespace fb
