// fb_comm.cu -- NCCL communicator (dlopen) and the equal-cell shard plan. See fb_comm.cuh.
#include "fb_comm.cuh"
#include "fb_gfl.cuh"
#include <dlfcn.h>
#include <cstdlib>
#include <cstring>

namespace fb {

namespace {
// the subset of nccl.h this library calls (NCCL 2.x ABI: enums are ints, ncclUniqueId is 128 opaque bytes)
struct UniqueId { char internal[128]; };
enum { kNcclSum = 0, kNcclMax = 2, kNcclMin = 3, kNcclInt64 = 4, kNcclFloat64 = 8 };
using fn_get_uid = int (*)(UniqueId *);
using fn_init_rank = int (*)(void **, int, UniqueId, int);
using fn_destroy = int (*)(void *);
using fn_group = int (*)();
using fn_send = int (*)(const void *, size_t, int, int, void *, cudaStream_t);
using fn_recv = int (*)(void *, size_t, int, int, void *, cudaStream_t);
using fn_allreduce = int (*)(const void *, void *, size_t, int, int, void *, cudaStream_t);
using fn_bcast = int (*)(const void *, void *, size_t, int, int, void *, cudaStream_t);
using fn_errstr = const char *(*)(int);
using fn_allgather = int (*)(const void *, void *, size_t, int, void *, cudaStream_t);
enum { F_UID, F_INIT, F_DESTROY, F_GSTART, F_GEND, F_SEND, F_RECV, F_ALLREDUCE, F_BCAST, F_ALLGATHER, F_ERRSTR };
const char *kNames[] = {"ncclGetUniqueId", "ncclCommInitRank", "ncclCommDestroy", "ncclGroupStart", "ncclGroupEnd",
                        "ncclSend", "ncclRecv", "ncclAllReduce", "ncclBroadcast", "ncclAllGather", "ncclGetErrorString"};
} // namespace

#define FB_NCCL(call)                                                                                       \
    do {                                                                                                    \
        int _r = (call);                                                                                    \
        if (_r != 0)                                                                                        \
            throw CommError(std::string(#call) + " failed: " + reinterpret_cast<fn_errstr>(fn_[F_ERRSTR])(_r)); \
    } while (0)

namespace { thread_local Comm *tl_current = nullptr; }
Comm &Comm::get() { static Comm process_wide; return tl_current ? *tl_current : process_wide; }
void Comm::set_current(Comm *c) { tl_current = c; }

void Comm::load() {
    if (lib_) return;
    const char *env = std::getenv("FB_NCCL_LIB");
    const char *cands[] = {env, "libnccl.so.2", "libnccl.so"};
    for (const char *c : cands) {
        if (!c || !*c) continue;
        lib_ = dlopen(c, RTLD_NOW | RTLD_GLOBAL);
        if (lib_) break;
    }
    if (!lib_) throw CommError("NCCL not found: set FB_NCCL_LIB to the path of libnccl.so.2");
    for (int i = 0; i <= F_ERRSTR; ++i) {
        fn_[i] = dlsym(lib_, kNames[i]);
        if (!fn_[i]) throw CommError(std::string("NCCL symbol missing: ") + kNames[i]);
    }
}

void Comm::unique_id(void *out128) {
    load();
    UniqueId id;
    FB_NCCL(reinterpret_cast<fn_get_uid>(fn_[F_UID])(&id));
    std::memcpy(out128, &id, sizeof id);
}

void Comm::init(int rank, int world, const void *id128) {
    if (world < 1 || rank < 0 || rank >= world) throw ArgError("bad rank / world size");
    destroy();
    rank_ = rank; world_ = world;
    if (world == 1) return;
    load();
    UniqueId id;
    std::memcpy(&id, id128, sizeof id);
    FB_NCCL(reinterpret_cast<fn_init_rank>(fn_[F_INIT])(&comm_, world, id, rank));
    // NCCL connects lazily: the first all-reduce, the first send/recv with each peer and the first broadcast each
    // cost 0.1-1 s of channel set-up (measured: 0.65 s inside the first solve on 2 GPUs, ~7 s on 8). Pay it here.
    cudaStream_t s = nullptr;
    double *buf = nullptr;
    FB_CUDA(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking));
    FB_CUDA(cudaMalloc(&buf, sizeof(double) * 4 * (size_t)(world + 1)));
    FB_CUDA(cudaMemsetAsync(buf, 0, sizeof(double) * 4 * (size_t)(world + 1), s));
    allreduce_sum(buf, buf + 2, 2, s);
    group_start();
    for (int p = 0; p < world; ++p) {
        if (p == rank) continue;
        send(buf, 1, p, s);
        recv(buf + 4 + 4 * p, 1, p, s);
    }
    group_end();
    group_start();
    for (int p = 0; p < world; ++p) bcast(buf + 4 + 4 * p + 1, 1, p, s);
    group_end();
    FB_CUDA(cudaStreamSynchronize(s));
    cudaFree(buf);
    cudaStreamDestroy(s);
}

void *Comm::ipc_open(const cudaIpcMemHandle_t &h) {
    for (auto &e : ipc_) if (std::memcmp(&e.first, &h, sizeof h) == 0) return e.second;
    void *p = nullptr;
    if (cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) { cudaGetLastError(); return nullptr; }
    ipc_.emplace_back(h, p);
    return p;
}

void Comm::destroy() {
    for (auto &e : ipc_) cudaIpcCloseMemHandle(e.second);
    ipc_.clear();
    if (comm_) { reinterpret_cast<fn_destroy>(fn_[F_DESTROY])(comm_); comm_ = nullptr; }
    rank_ = 0; world_ = 1;
}

void Comm::group_start() { need(); FB_NCCL(reinterpret_cast<fn_group>(fn_[F_GSTART])()); }
void Comm::group_end() { need(); FB_NCCL(reinterpret_cast<fn_group>(fn_[F_GEND])()); }
void Comm::send(const double *p, size_t n, int peer, cudaStream_t s) {
    need();
    FB_NCCL(reinterpret_cast<fn_send>(fn_[F_SEND])(p, n, kNcclFloat64, peer, comm_, s));
}
void Comm::recv(double *p, size_t n, int peer, cudaStream_t s) {
    need();
    FB_NCCL(reinterpret_cast<fn_recv>(fn_[F_RECV])(p, n, kNcclFloat64, peer, comm_, s));
}
void Comm::allreduce_sum(const double *src, double *dst, size_t n, cudaStream_t s) {
    need();
    FB_NCCL(reinterpret_cast<fn_allreduce>(fn_[F_ALLREDUCE])(src, dst, n, kNcclFloat64, kNcclSum, comm_, s));
}
void Comm::allreduce_i64(long long *buf, size_t n, bool take_max, cudaStream_t s) {
    need();
    FB_NCCL(reinterpret_cast<fn_allreduce>(fn_[F_ALLREDUCE])(buf, buf, n, kNcclInt64, take_max ? kNcclMax : kNcclMin, comm_, s));
}
void Comm::allgather(double *buf, size_t n_per_rank, cudaStream_t s) {
    need();
    FB_NCCL(reinterpret_cast<fn_allgather>(fn_[F_ALLGATHER])(buf + (size_t)rank_ * n_per_rank, buf, n_per_rank, kNcclFloat64, comm_, s));
}
void Comm::bcast(double *p, size_t n, int root, cudaStream_t s) {
    need();
    FB_NCCL(reinterpret_cast<fn_bcast>(fn_[F_BCAST])(p, p, n, kNcclFloat64, root, comm_, s));
}

void shard_plan(int n, int md, int world, int *row_begin, int *col_begin) {
    if (n < 1 || world < 1) throw ArgError("bad shard plan arguments");
    if (md <= 0 || md > n) md = n;
    const Grid g = Grid::make(n, md);
    int r = 0, c = 0;
    for (int k = 0; k <= world; ++k) {
        // first line whose start offset reaches k/world of the cells (ceil), so that blocks differ by < one line
        const int64_t target = (g.ncells * (int64_t)k + world - 1) / world;
        while (r < n && g.rm_row_start(r) < target) ++r;
        while (c < n && g.cm_col_start(c) < target) ++c;
        row_begin[k] = (k == world) ? n : r;
        col_begin[k] = (k == world) ? n : c;
    }
    row_begin[0] = col_begin[0] = 0;
}

} // namespace fb
