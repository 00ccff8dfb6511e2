// fb_util.cuh -- shared helpers for the fast-binless CUDA library (sm_100a).
#pragma once
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <exception>
#include <map>
#include <atomic>
#include <mutex>
#include <stdexcept>
#include <string>

namespace fb {

struct CudaError : std::runtime_error { using std::runtime_error::runtime_error; };
struct ArgError : std::runtime_error { using std::runtime_error::runtime_error; };
struct NumError : std::runtime_error { using std::runtime_error::runtime_error; };
struct StateError : std::runtime_error { using std::runtime_error::runtime_error; };
struct Interrupted : std::runtime_error { using std::runtime_error::runtime_error; };

#define FB_CUDA(expr)                                                                          \
    do {                                                                                       \
        cudaError_t _e = (expr);                                                               \
        if (_e != cudaSuccess)                                                                 \
            throw fb::CudaError(std::string(#expr) + " failed: " + cudaGetErrorString(_e) +    \
                                " (" __FILE__ ":" + std::to_string(__LINE__) + ")");           \
    } while (0)

#define FB_LAUNCH_CHECK() FB_CUDA(cudaGetLastError())

// Size-keyed cache of device / pinned allocations: fast_binless is called repeatedly with same-shaped inputs (one
// call per locus, then fast_binless_difference on the result), and cudaMalloc / cudaFree / cudaMallocHost cost
// milliseconds each. Blocks are returned to the cache only by owners that have synchronised their stream
// (every API call of this library ends synchronised), so a cached block is never still in use.
class BlockCache {
public:
    explicit BlockCache(bool pinned) : pinned_(pinned) {}
    void *get(size_t bytes) {
        bytes = (bytes + 255) & ~size_t(255);
        {
            std::lock_guard<std::mutex> g(mu_);
            auto it = free_.find(bytes);
            if (it != free_.end()) { void *p = it->second; free_.erase(it); cached_ -= bytes; return p; }
        }
        void *p = nullptr;
        cudaError_t e = pinned_ ? cudaMallocHost(&p, bytes) : cudaMalloc(&p, bytes);
        if (e != cudaSuccess) { // give the cached blocks (and the idle solver objects that hold some) back and retry once
            cudaGetLastError();
            if (void (*hook)() = pressure_hook().load()) hook();
            trim();
            e = pinned_ ? cudaMallocHost(&p, bytes) : cudaMalloc(&p, bytes);
        }
        if (e != cudaSuccess)
            throw CudaError(std::string(pinned_ ? "cudaMallocHost" : "cudaMalloc") + " of " + std::to_string(bytes) + " bytes failed: " + cudaGetErrorString(e));
        return p;
    }
    void put(void *p, size_t bytes) {
        if (!p) return;
        bytes = (bytes + 255) & ~size_t(255);
        std::lock_guard<std::mutex> g(mu_);
        if (cached_ + bytes > limit_) { if (pinned_) cudaFreeHost(p); else cudaFree(p); return; }
        free_.emplace(bytes, p); cached_ += bytes;
    }
    void trim() {
        std::lock_guard<std::mutex> g(mu_);
        for (auto &kv : free_) { if (pinned_) cudaFreeHost(kv.second); else cudaFree(kv.second); }
        free_.clear(); cached_ = 0;
    }
    // one cache per device: a single-process multi-GPU handle runs one host thread per device (fb_create_multi)
    static BlockCache &device() {
        static BlockCache c[16] = {BlockCache(false), BlockCache(false), BlockCache(false), BlockCache(false), BlockCache(false),
                                   BlockCache(false), BlockCache(false), BlockCache(false), BlockCache(false), BlockCache(false),
                                   BlockCache(false), BlockCache(false), BlockCache(false), BlockCache(false), BlockCache(false),
                                   BlockCache(false)};
        int dev = 0;
        if (cudaGetDevice(&dev) != cudaSuccess) { cudaGetLastError(); dev = 0; }
        return c[dev & 15];
    }
    static BlockCache &host() { static BlockCache c(true); return c; }
    // called before the cache is emptied because an allocation failed: owners of idle objects let them go (fb_model.cu)
    static std::atomic<void (*)()> &pressure_hook() { static std::atomic<void (*)()> h{nullptr}; return h; }
private:
    bool pinned_;
    std::mutex mu_;
    std::multimap<size_t, void *> free_;
    size_t cached_ = 0;
    size_t limit_ = pinned_ ? (size_t(1) << 30) : (size_t(48) << 30);
};

// RAII device buffer
template <typename T>
struct DevBuf {
    T *p = nullptr;
    size_t n = 0;
    DevBuf() {}
    explicit DevBuf(size_t n_) { alloc(n_); }
    DevBuf(const DevBuf &) = delete;
    DevBuf &operator=(const DevBuf &) = delete;
    DevBuf(DevBuf &&o) noexcept : p(o.p), n(o.n) { o.p = nullptr; o.n = 0; }
    DevBuf &operator=(DevBuf &&o) noexcept { if (this != &o) { release(); p = o.p; n = o.n; o.p = nullptr; o.n = 0; } return *this; }
    ~DevBuf() { release(); }
    void alloc(size_t n_) {
        release();
        n = n_;
        if (n) p = static_cast<T *>(BlockCache::device().get(n * sizeof(T)));
    }
    void release() {
        if (p) {
            if (std::uncaught_exceptions()) cudaDeviceSynchronize(); // error unwinding: work may still be queued
            BlockCache::device().put(p, n * sizeof(T));
        }
        p = nullptr; n = 0;
    }
    void zero(cudaStream_t s = 0) { if (n) FB_CUDA(cudaMemsetAsync(p, 0, n * sizeof(T), s)); }
    void upload(const T *h, size_t cnt, cudaStream_t s = 0) { FB_CUDA(cudaMemcpyAsync(p, h, cnt * sizeof(T), cudaMemcpyHostToDevice, s)); }
    void download(T *h, size_t cnt, cudaStream_t s = 0) const { FB_CUDA(cudaMemcpyAsync(h, p, cnt * sizeof(T), cudaMemcpyDeviceToHost, s)); }
};

// ---- packed upper-triangle index helpers (0-based) -------------------------
// row-major packed: cell (i,j), j>=i, lives at rm_start(i,n) + (j-i);
// matches the reference's cell order (FastData.ipp / SURVEY appendix A).
__host__ __device__ __forceinline__ int64_t rm_start(int64_t i, int64_t n) { return i * n - (i * (i - 1)) / 2; }
// column-major packed: cell (i,j) lives at cm_start(j) + i
__host__ __device__ __forceinline__ int64_t cm_start(int64_t j) { return (j * (j + 1)) / 2; }

#ifdef __CUDACC__
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ double warp_max(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmax(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}
__device__ __forceinline__ double warp_min(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmin(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}
#endif

} // namespace fb
