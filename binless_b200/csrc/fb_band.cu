// fb_band.cu -- the output side of chromosome_binless (/root/reference/binless/R/chromosome.R:262-330,
// signal case): from the fitted matrix in HBM
//   lam1.sig  = median(log(signal)) over the cells with distance > far.cutoff, per dataset (:270-276; lam1.final when
//               there is no such cell), signal := pmax(signal / exp(lam1.sig), 1)
//   raw band / binless band : the cells with bin2 - bin1 <= maxdiag as (bin1, bin2, observed) and (bin1, bin2, binless)
//               triplets, what sparseMatrix(i, j, x, symmetric = T) takes (:303-316)
//   log10 signal : the cells with signal != 1 as (bin1, bin2, log10(signal)) triplets (:318-323)
// so that only the band and the sparse signal cross PCIe, not the N-row data frame. The median is exact: a radix select
// over order-preserving 64-bit keys (four 16-bit histogram passes per order statistic), no sort.
#include "fb_model.cuh"
#include <algorithm>
#include <cmath>
#include <cstring>
#include <vector>

namespace fb {

namespace {

constexpr int BT = 256;

__device__ __forceinline__ unsigned long long ukey(double x) {      // order-preserving double -> uint64
    const unsigned long long b = (unsigned long long)__double_as_longlong(x);
    return (b >> 63) ? ~b : (b | 0x8000000000000000ull);
}
inline double ukey_inv(unsigned long long k) {
    const unsigned long long b = (k >> 63) ? (k & 0x7fffffffffffffffull) : ~k;
    double x; std::memcpy(&x, &b, sizeof x); return x;
}
__device__ __forceinline__ double soft1(double x, double l1) { return x > 0 ? fmax(0., x - l1) : fmin(0., x + l1); }
// log(signal) as R computes it from the data frame column: signal = exp(soft(log_signal, lam1)) (fast_dataframe.cpp:71-75)
__device__ __forceinline__ double log_signal_col(double phi, double l1) { return log(exp(soft1(phi, l1))); }

// grid (ceil(n / BT), n): thread -> cell (i, i + off) of dataset d
#define BAND_CELL(n_)                                                        \
    const int i = blockIdx.y, off = blockIdx.x * BT + threadIdx.x, j = i + off; \
    const bool in = off < (n_) - i;                                          \
    const int64_t cell = rm_start(i, (n_)) + off;

// one 16-bit digit of the radix select: histogram of digit `shift` over the far cells whose higher digits equal `prefix`
__global__ void __launch_bounds__(BT)
k_sel_hist(int n, const double *__restrict__ phi_d, double l1, const unsigned char *__restrict__ far_diag, int shift,
           unsigned long long prefix, unsigned int *__restrict__ hist) {
    BAND_CELL(n)
    if (!in || !far_diag[j - i]) return;
    const unsigned long long k = ukey(log_signal_col(phi_d[cell], l1));
    if (shift < 48 && (k >> (shift + 16)) != prefix) return;
    atomicAdd(&hist[(k >> shift) & 0xffffu], 1u);
}

// band triplets at their closed-form slot (row-major over the band of width md1 = maxdiag + 1)
__global__ void __launch_bounds__(BT)
k_band_fill(int n, Grid band, const int32_t *__restrict__ obs_d, const double *__restrict__ phi_d, const double *__restrict__ ddiag,
            double l1, int32_t *__restrict__ b1, int32_t *__restrict__ b2, double *__restrict__ out_obs, double *__restrict__ out_binless) {
    BAND_CELL(n)
    if (!in || off >= band.md) return;
    const int64_t slot = band.rm(i, j);
    if (b1) { b1[slot] = i + 1; b2[slot] = j + 1; }
    if (out_obs) out_obs[slot] = (double)obs_d[cell];
    if (out_binless) out_binless[slot] = exp(ddiag[off] + soft1(phi_d[cell], l1));     // fast_dataframe.cpp:76
}

__device__ __forceinline__ double thresholded_signal(double phi, double l1, double lam1_sig) {
    return fmax(exp(soft1(phi, l1)) / exp(lam1_sig), 1.0);                            // chromosome.R:277
}
// cells with signal != 1 per block, then the triplets in row-major order (block offsets from the host-side scan)
__global__ void __launch_bounds__(BT)
k_sig_count(int n, const double *__restrict__ phi_d, double l1, double lam1_sig, unsigned int *__restrict__ counts) {
    BAND_CELL(n)
    const bool f = in && thresholded_signal(phi_d[cell], l1, lam1_sig) != 1.0;
    const unsigned m = __ballot_sync(0xffffffffu, f);
    if ((threadIdx.x & 31) == 0 && m) atomicAdd(&counts[(int64_t)blockIdx.y * gridDim.x + blockIdx.x], (unsigned)__popc(m));
}
__global__ void __launch_bounds__(BT)
k_sig_fill(int n, const double *__restrict__ phi_d, double l1, double lam1_sig, const int64_t *__restrict__ offs,
           int32_t *__restrict__ b1, int32_t *__restrict__ b2, double *__restrict__ val) {
    __shared__ int s_w[BT / 32];
    BAND_CELL(n)
    const double s = in ? thresholded_signal(phi_d[cell], l1, lam1_sig) : 1.0;
    const bool f = s != 1.0;
    const unsigned m = __ballot_sync(0xffffffffu, f);
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    if (lane == 0) s_w[wid] = __popc(m);
    __syncthreads();
    int base = 0;
    for (int k = 0; k < wid; ++k) base += s_w[k];
    if (f) {
        const int64_t o = offs[(int64_t)blockIdx.y * gridDim.x + blockIdx.x] + base + __popc(m & ((1u << lane) - 1u));
        b1[o] = i + 1; b2[o] = j + 1; val[o] = log10(s);
    }
}

} // namespace

void Model::band_export(const fb_band_params *p, fb_band_out *out) {
    if (!p || !out) throw ArgError("null argument");
    if (!has_state_) throw StateError("band_export needs a fitted or restored model");
    if (generic_) throw ArgError("band_export needs a bin_data() layout (distance a function of bin2 - bin1)");
    const int n = t_.n, D = t_.D;
    const int64_t nc = t_.ncells;
    kernel_launches = 0;
    const std::vector<double> lam1 = [&] {
        std::vector<double> v(D, 0.0);
        if (p->lam1 && p->n_lam1 > 0) for (int d = 0; d < D; ++d) v[d] = p->n_lam1 == D ? p->lam1[d] : p->lam1[0];
        return v;
    }();
    const dim3 grid((unsigned)((n + BT - 1) / BT), (unsigned)n);
    // distance > far.cutoff is a property of the diagonal in this layout
    std::vector<unsigned char> far(n);
    int64_t nfar = 0;
    for (int k = 0; k < n; ++k) { far[k] = dist_[k] > p->far_cutoff; if (far[k]) nfar += n - k; }
    DevBuf<unsigned char> far_d(n);
    FB_CUDA(cudaMemcpyAsync(far_d.p, far.data(), n, cudaMemcpyHostToDevice, st_));
    DevBuf<unsigned int> hist(65536);
    std::vector<unsigned int> hh(65536);
    std::vector<double> lam1_sig(D);
    auto kth = [&](int d, int64_t k) -> double {    // k-th smallest (0-based) of log(signal) over the far cells of dataset d
        unsigned long long prefix = 0;
        for (int shift = 48; shift >= 0; shift -= 16) {
            hist.zero(st_);
            k_sel_hist<<<grid, BT, 0, st_>>>(n, phi_.p + (size_t)d * nc, lam1[d], far_d.p, shift, prefix, hist.p);
            kernel_launches += 1;
            FB_CUDA(cudaMemcpyAsync(hh.data(), hist.p, 65536 * sizeof(unsigned int), cudaMemcpyDeviceToHost, st_));
            FB_CUDA(cudaStreamSynchronize(st_));
            unsigned digit = 0;
            for (; digit < 65536; ++digit) { if (k < (int64_t)hh[digit]) break; k -= hh[digit]; }
            if (digit == 65536) throw NumError("radix select ran past the histogram");
            prefix = (prefix << 16) | digit;
        }
        return ukey_inv(prefix);
    };
    for (int d = 0; d < D; ++d) {
        if (nfar == 0) { lam1_sig[d] = p->lam1_final; continue; }           // median of nothing is NA (chromosome.R:271-275)
        const double a = kth(d, (nfar - 1) / 2);
        lam1_sig[d] = (nfar & 1) ? a : (a + kth(d, nfar / 2)) / 2;          // R's median
        if (std::isnan(lam1_sig[d])) lam1_sig[d] = p->lam1_final;
    }
    if (out->lam1_sig) std::memcpy(out->lam1_sig, lam1_sig.data(), D * sizeof(double));
    // band
    const int md1 = (int)std::min<int64_t>((int64_t)p->maxdiag + 1, n);
    const Grid band = Grid::make(n, md1);
    out->nband = band.ncells;
    if (out->band_bin1 || out->band_bin2 || out->band_observed || out->band_binless) {
        DevBuf<int32_t> b1(band.ncells), b2(band.ncells);
        DevBuf<double> bo(band.ncells), bb(band.ncells);
        for (int d = 0; d < D; ++d) {
            k_band_fill<<<grid, BT, 0, st_>>>(n, band, obs_.p + (size_t)d * nc, phi_.p + (size_t)d * nc, tab_.p + tab_b_, lam1[d],
                                              d == 0 ? b1.p : nullptr, b2.p, out->band_observed ? bo.p : nullptr,
                                              out->band_binless ? bb.p : nullptr);
            kernel_launches += 1;
            if (out->band_observed) download_bytes(out->band_observed + (size_t)d * band.ncells, bo.p, band.ncells * sizeof(double));
            if (out->band_binless) download_bytes(out->band_binless + (size_t)d * band.ncells, bb.p, band.ncells * sizeof(double));
            if (d == 0) {
                if (out->band_bin1) download_bytes(out->band_bin1, b1.p, band.ncells * sizeof(int32_t));
                if (out->band_bin2) download_bytes(out->band_bin2, b2.p, band.ncells * sizeof(int32_t));
            }
            FB_CUDA(cudaStreamSynchronize(st_));
        }
    }
    // sparse log10 signal
    const int64_t nblk = (int64_t)grid.x * grid.y;
    DevBuf<unsigned int> counts(nblk);
    DevBuf<int64_t> offs(nblk);
    std::vector<unsigned int> hc(nblk);
    std::vector<int64_t> ho(nblk);
    int64_t total = 0;
    for (int d = 0; d < D; ++d) {
        counts.zero(st_);
        k_sig_count<<<grid, BT, 0, st_>>>(n, phi_.p + (size_t)d * nc, lam1[d], lam1_sig[d], counts.p);
        kernel_launches += 1;
        FB_CUDA(cudaMemcpyAsync(hc.data(), counts.p, nblk * sizeof(unsigned int), cudaMemcpyDeviceToHost, st_));
        FB_CUDA(cudaStreamSynchronize(st_));
        int64_t acc = 0;
        for (int64_t b = 0; b < nblk; ++b) { ho[b] = acc; acc += hc[b]; }
        if (out->nsig) out->nsig[d] = acc;
        if (acc && out->sig_bin1 && out->sig_bin2 && out->sig_log10) {
            if (total + acc > out->sig_cap) throw ArgError("sig_cap is too small: call once without the sparse arrays to get nsig");
            FB_CUDA(cudaMemcpyAsync(offs.p, ho.data(), nblk * sizeof(int64_t), cudaMemcpyHostToDevice, st_));
            DevBuf<int32_t> s1(acc), s2(acc);
            DevBuf<double> sv(acc);
            k_sig_fill<<<grid, BT, 0, st_>>>(n, phi_.p + (size_t)d * nc, lam1[d], lam1_sig[d], offs.p, s1.p, s2.p, sv.p);
            kernel_launches += 1;
            download_bytes(out->sig_bin1 + total, s1.p, acc * sizeof(int32_t));
            download_bytes(out->sig_bin2 + total, s2.p, acc * sizeof(int32_t));
            download_bytes(out->sig_log10 + total, sv.p, acc * sizeof(double));
            FB_CUDA(cudaStreamSynchronize(st_));
        }
        total += acc;
    }
    FB_LAUNCH_CHECK();
}

} // namespace fb
