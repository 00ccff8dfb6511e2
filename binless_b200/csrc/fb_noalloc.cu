// fb_noalloc.cu -- fb_far_median_noalloc: the far-cell median of fb_band_export (fb_band.cu, k_sel_hist; the lam1.sig of
// R/chromosome.R:270-276) for callers that must not allocate: code under CUDA graph capture, tight serving loops,
// harnesses that forbid cudaMalloc. It has NO counterpart in the reference;
// no other code path of the library uses anything in this file.
//
// Same algorithm as the band export: an exact radix select over order-preserving 64-bit keys, four 16-bit digits per
// order statistic. What differs is where the scratch lives and who reads it:
//   * the 65536-bin digit histogram and the select state are a file-scope `static __device__` array of
//     FB_NOALLOC_SCRATCH_BYTES (256 KiB + 256 B), which every GPU that loads the library holds until the process exits;
//   * the digit is picked ON THE DEVICE (k_na_pick) instead of on the host, so the call enqueues its kernels on the
//     caller's stream and returns: no cudaMalloc / cudaFree / cudaMemcpy / synchronisation, inputs and the result stay
//     in the caller's device buffers.
// The array's name is a host-side stand-in in host code, not a device address: its address is taken once per device
// with cudaGetSymbolAddress (scratch_of_current_device) and handed to the kernels as an ordinary pointer.
//
// All calls on one device SHARE the array and it keeps the last call's contents: k_na_reset clears what must start at
// zero at the head of every order statistic and k_na_pick leaves the histogram zeroed behind it. Two calls whose kernels
// overlap on one GPU (different streams, or several host threads) race on it: the caller serialises them.
#include "fb_model.cuh"
#include "../../include/fastbinless.h"
#include <cmath>
#include <mutex>

namespace fb {

namespace {

constexpr int NA_BT = 256;                 // threads per block of the histogram pass
constexpr int NA_DIGITS = 65536;           // one 16-bit digit per pass
constexpr int NA_PICK_T = 1024;            // k_na_pick: one block, 64 histogram bins per thread
constexpr int NA_MAX_DEVICES = 64;

struct NaState {                           // lives behind the histogram, 256-byte aligned
    unsigned long long prefix;             // digits chosen so far (high to low)
    long long k;                           // rank still to find among the cells that share `prefix`
    double val[2];                         // the one or two middle order statistics
};
static_assert(sizeof(NaState) <= 256, "state slot");
static_assert(FB_NOALLOC_SCRATCH_BYTES == NA_DIGITS * sizeof(unsigned int) + 256, "header and kernel disagree on the scratch size");

static __device__ __align__(256) unsigned char g_na_scratch[FB_NOALLOC_SCRATCH_BYTES];

__device__ __forceinline__ unsigned int *na_hist(unsigned char *scratch) { return reinterpret_cast<unsigned int *>(scratch); }
__device__ __forceinline__ NaState *na_state(unsigned char *scratch) {
    return reinterpret_cast<NaState *>(scratch + NA_DIGITS * sizeof(unsigned int));
}

__device__ __forceinline__ unsigned long long na_ukey(double x) {           // order-preserving double -> uint64
    const unsigned long long b = (unsigned long long)__double_as_longlong(x);
    return (b >> 63) ? ~b : (b | 0x8000000000000000ull);
}
__device__ __forceinline__ double na_ukey_inv(unsigned long long k) {
    const unsigned long long b = (k >> 63) ? (k & 0x7fffffffffffffffull) : ~k;
    return __longlong_as_double((long long)b);
}
__device__ __forceinline__ double na_soft1(double x, double l1) { return x > 0 ? fmax(0., x - l1) : fmin(0., x + l1); }
// log(signal) as R computes it from the data frame column (fast_dataframe.cpp:71-75); the expression of fb_band.cu
__device__ __forceinline__ double na_log_signal_col(double phi, double l1) { return log(exp(na_soft1(phi, l1))); }

// head of one order statistic: the histogram and the select state start from zero whatever the last call left
__global__ void __launch_bounds__(NA_PICK_T)
k_na_reset(unsigned char *scratch, long long k0) {
    unsigned int *hist = na_hist(scratch);
    for (int b = blockIdx.x * NA_PICK_T + threadIdx.x; b < NA_DIGITS; b += gridDim.x * NA_PICK_T) hist[b] = 0u;
    if (blockIdx.x == 0 && threadIdx.x == 0) { NaState *s = na_state(scratch); s->prefix = 0ull; s->k = k0; }
}

// one 16-bit digit: histogram of digit `shift` over the far cells (bin2 - bin1 >= far_from) whose higher digits equal
// the prefix in the scratch. grid (ceil((n - far_from) / NA_BT), n - far_from): only the far corner of the triangle.
__global__ void __launch_bounds__(NA_BT)
k_na_hist(int n, int far_from, const double *__restrict__ phi, double l1, int shift, unsigned char *scratch) {
    const int i = blockIdx.y, off = far_from + blockIdx.x * NA_BT + threadIdx.x;
    if (off >= n - i) return;
    const unsigned long long prefix = na_state(scratch)->prefix;
    const unsigned long long k = na_ukey(na_log_signal_col(phi[rm_start(i, n) + off], l1));
    if (shift < 48 && (k >> (shift + 16)) != prefix) return;
    atomicAdd(&na_hist(scratch)[(k >> shift) & 0xffffu], 1u);
}

// pick the digit that holds rank k (the host loop of Model::band_export, on the device), push it onto the prefix and
// leave the histogram zeroed for the next pass. After the last digit the key is the order statistic `which`; with
// `finish` the median goes to *out: the statistic itself (odd count) or R's mean of the two middle ones.
__global__ void __launch_bounds__(NA_PICK_T)
k_na_pick(unsigned char *scratch, int shift, int which, int finish, int odd, double *__restrict__ out) {
    __shared__ unsigned long long s_warp[NA_PICK_T / 32];
    unsigned int *hist = na_hist(scratch);
    NaState *s = na_state(scratch);
    constexpr int PER = NA_DIGITS / NA_PICK_T;
    const int t = threadIdx.x, lane = t & 31, wid = t >> 5;
    const long long k = s->k;
    const unsigned long long prefix = s->prefix;
    unsigned int h[PER];
    unsigned long long mine = 0;
#pragma unroll
    for (int b = 0; b < PER; ++b) { h[b] = hist[t * PER + b]; mine += h[b]; hist[t * PER + b] = 0u; }
    unsigned long long incl = mine;                         // inclusive scan over the block
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) { const unsigned long long v = __shfl_up_sync(0xffffffffu, incl, d); if (lane >= d) incl += v; }
    if (lane == 31) s_warp[wid] = incl;
    __syncthreads();                                        // also orders every read of s->k / s->prefix before the write below
    unsigned long long before = 0;
    for (int w = 0; w < wid; ++w) before += s_warp[w];
    const unsigned long long excl = before + incl - mine;
    if ((unsigned long long)k >= excl && (unsigned long long)k < excl + mine) {      // exactly one thread
        long long rem = k - (long long)excl;
        int b = 0;
        for (; b < PER - 1; ++b) { if (rem < (long long)h[b]) break; rem -= h[b]; }
        const unsigned long long np = (prefix << 16) | (unsigned long long)(t * PER + b);
        s->prefix = np;
        s->k = rem;
        if (shift == 0) {
            const double v = na_ukey_inv(np);
            s->val[which] = v;
            if (finish) *out = odd ? v : (s->val[0] + v) / 2;
        }
    }
}

__global__ void k_na_store(double *out, double v) { *out = v; }

// device address of the scratch array on the CURRENT device: cudaGetSymbolAddress once per device, with it current
unsigned char *scratch_of_current_device() {
    static std::mutex mu;
    static unsigned char *ptr[NA_MAX_DEVICES] = {};
    int dev = -1;
    const cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) throw CudaError(std::string("no CUDA device available (there is no CPU fallback): ") + cudaGetErrorString(e));
    if (dev < 0 || dev >= NA_MAX_DEVICES) throw CudaError("device index beyond the no-allocation scratch table");
    std::lock_guard<std::mutex> lock(mu);
    if (!ptr[dev]) {
        unsigned char *p = nullptr;
        const cudaError_t g = cudaGetSymbolAddress((void **)&p, g_na_scratch);
        if (g != cudaSuccess || !p) throw CudaError(std::string("cudaGetSymbolAddress(no-allocation scratch): ") + cudaGetErrorString(g));
        ptr[dev] = p;
    }
    return ptr[dev];
}

void far_median_noalloc(int nbins, const double *log_signal_dev, double lam1, int far_from_diag, double *median_dev, cudaStream_t st) {
    if (!log_signal_dev || !median_dev) throw ArgError("null device pointer");
    if (nbins < 1) throw ArgError("nbins must be positive");
    if (nbins > FB_NOALLOC_MAX_NBINS)
        throw ArgError("nbins is beyond FB_NOALLOC_MAX_NBINS (65535), the largest matrix the static scratch of "
                       "fb_far_median_noalloc serves; use fb_band_export");
    if (far_from_diag < 0) throw ArgError("far_from_diag must not be negative");
    unsigned char *scratch = scratch_of_current_device();
    const int64_t m = far_from_diag < nbins ? nbins - far_from_diag : 0;     // rows (and longest row) of the far corner
    const int64_t nfar = m * (m + 1) / 2;
    if (nfar == 0) {                                                         // median of nothing: NA in R (chromosome.R:271-275)
        k_na_store<<<1, 1, 0, st>>>(median_dev, NAN);
        FB_LAUNCH_CHECK();
        return;
    }
    const dim3 grid((unsigned)((m + NA_BT - 1) / NA_BT), (unsigned)m);
    const int odd = (int)(nfar & 1);
    const int nstat = odd ? 1 : 2;
    for (int which = 0; which < nstat; ++which) {
        k_na_reset<<<NA_DIGITS / NA_PICK_T, NA_PICK_T, 0, st>>>(scratch, which == 0 ? (nfar - 1) / 2 : nfar / 2);
        for (int shift = 48; shift >= 0; shift -= 16) {
            k_na_hist<<<grid, NA_BT, 0, st>>>(nbins, far_from_diag, log_signal_dev, lam1, shift, scratch);
            k_na_pick<<<1, NA_PICK_T, 0, st>>>(scratch, shift, which, which == nstat - 1, odd, median_dev);
        }
    }
    FB_LAUNCH_CHECK();
}

} // namespace

} // namespace fb

extern "C" FB_API int fb_far_median_noalloc(int nbins, const double *log_signal_dev, double lam1, int far_from_diag,
                                            double *median_dev, void *stream) {
    try { fb::far_median_noalloc(nbins, log_signal_dev, lam1, far_from_diag, median_dev, (cudaStream_t)stream); return FB_OK; }
    catch (const fb::ArgError &e) { fb::g_last_error = e.what(); return FB_ERR_ARG; }
    catch (const fb::CudaError &e) { fb::g_last_error = e.what(); return FB_ERR_CUDA; }
    catch (const std::exception &e) { fb::g_last_error = e.what(); return FB_ERR_NUMERIC; }
}
