// fb_util.cuh -- shared helpers for the fast-binless CUDA library (sm_100a).
#pragma once
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
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
        if (n) FB_CUDA(cudaMalloc((void **)&p, n * sizeof(T)));
    }
    void release() { if (p) cudaFree(p); p = nullptr; n = 0; }
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
