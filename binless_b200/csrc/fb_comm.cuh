// fb_comm.cuh -- one process per GPU: the process-wide NCCL communicator of the library.
//
// The only multi-GPU step on this path is the trail-block ADMM of ONE fused-lasso solve (SURVEY.md section 8e,
// "trail blocks within one matrix"): the chains of graph_fused_lasso_weight_warm (gfl_graph_fl.c:53-213) are split
// into equal-cell row blocks and column blocks, one pair per rank, and the consensus terms z - u of the cells where a
// row block of one rank crosses a column block of another are exchanged once per iteration (grouped ncclSend/ncclRecv
// over NVLink), plus one small all-reduce of the per-chain residual sums. NCCL is loaded at run time with dlopen
// (libnccl.so.2, the copy already mapped by torch when the host is Python), so the library itself links cudart only
// and still loads on a machine without NCCL.
#pragma once
#include "fb_util.cuh"
#include <vector>

namespace fb {

struct CommError : std::runtime_error { using std::runtime_error::runtime_error; };

// Equal-cell split of the band grid (n bins, band width md; md == n is the full triangle) into `world` row blocks and
// `world` column blocks: row_begin / col_begin get world + 1 entries, block g = [begin[g], begin[g+1]).
void shard_plan(int n, int md, int world, int *row_begin, int *col_begin);

class Comm {
public:
    // the communicator of the calling host thread: the process-wide one (one process per GPU), or the one a rank thread
    // of a single-process multi-GPU handle installed with set_current (fb_create_multi: one host thread per device)
    static Comm &get();
    static void set_current(Comm *c);
    Comm() {}
    Comm(const Comm &) = delete;
    Comm &operator=(const Comm &) = delete;
    bool active() const { return comm_ != nullptr && world_ > 1; }
    int rank() const { return rank_; }
    int world() const { return world_; }
    void unique_id(void *out128);                          // ncclGetUniqueId (rank 0; the host distributes the bytes)
    void init(int rank, int world, const void *id128);     // ncclCommInitRank on the current device
    void destroy();
    void group_start();
    void group_end();
    void send(const double *p, size_t n, int peer, cudaStream_t s);
    void recv(double *p, size_t n, int peer, cudaStream_t s);
    void allreduce_sum(const double *src, double *dst, size_t n, cudaStream_t s);
    void bcast(double *p, size_t n, int root, cudaStream_t s);
    void allreduce_i64(long long *buf, size_t n, bool take_max, cudaStream_t s);   // in place; exact in any order
    void allgather(double *buf, size_t n_per_rank, cudaStream_t s);                // in place: rank r's part at buf + r * n_per_rank
    // CUDA IPC mapping of a peer's allocation, kept for the life of the communicator (opening one costs tens of
    // milliseconds; the solver of every solve asks for the same blocks again); nullptr when it cannot be opened
    void *ipc_open(const cudaIpcMemHandle_t &h);
private:
    void need() const { if (!comm_) throw CommError("no communicator (fb_comm_init has not been called on this thread)"); }
    std::vector<std::pair<cudaIpcMemHandle_t, void *>> ipc_;
    void load();
    void *lib_ = nullptr, *comm_ = nullptr;
    int rank_ = 0, world_ = 1;
    void *fn_[16] = {};
};

} // namespace fb
