// fb_gfl.cuh -- graph-fused-lasso ADMM solver on the triangle / band grid (device-resident).
//
// Replaces, on the GPU, what the reference does in
//   FusedLassoGaussianEstimator<GFLLibrary>            FusedLassoGaussianEstimator.hpp:16-71
//   GFLLibraryTriangleImpl / GFLLibraryTrapezeImpl     GFLLibraryTriangleImpl.cpp, GFLLibraryTrapezeImpl.cpp
//   graph_fused_lasso_weight_warm + update_*           gfl_graph_fl.c:53-365
//   tf_dp_weight                                       gfl_tf.c:184-321
// (paths relative to /root/reference/binless/src). See DESIGN.md for the layout.
#pragma once
#include "fb_util.cuh"
#include <memory>
#include <vector>

namespace fb {

struct AsetPlan;   // fb_prox_aset.cuh

// Geometry of the solved grid: upper-triangular band of width md (md == n is the
// full triangle). Cells (i,j), 0 <= j-i < md.
//   "rm" = row-major packed (row i contiguous), "cm" = column-major packed.
struct Grid {
    int n = 0, md = 0;
    int nfull = 0;          // rows holding md cells
    int64_t ncells = 0;     // cells in the band
    __host__ __device__ __forceinline__ int row_len(int i) const { return min_(md, n - i); }
    __host__ __device__ __forceinline__ int col_len(int j) const { return min_(md, j + 1); }
    __host__ __device__ __forceinline__ int col_first_row(int j) const { return j + 1 - col_len(j); }
    __host__ __device__ __forceinline__ int64_t rm_row_start(int64_t i) const {
        if (i <= nfull) return i * md;
        // rows nfull..i-1 hold n-k cells
        return (int64_t)nfull * md + (i - nfull) * (int64_t)n - (i * (i - 1) - (int64_t)nfull * (nfull - 1)) / 2;
    }
    __host__ __device__ __forceinline__ int64_t cm_col_start(int64_t j) const {
        if (j < md) return (j * (j + 1)) / 2;
        return ((int64_t)md * (md + 1)) / 2 + (j - md) * (int64_t)md;
    }
    __host__ __device__ __forceinline__ int64_t rm(int i, int j) const { return rm_row_start(i) + (j - i); }
    __host__ __device__ __forceinline__ int64_t cm(int i, int j) const { return cm_col_start(j) + (i - col_first_row(j)); }
    __host__ __device__ __forceinline__ int nchains() const { return 2 * (n - 1); }
    // chain q: even -> row q/2, odd -> column n-1-q/2 (equal lengths, long first)
    __host__ __device__ __forceinline__ void chain(int q, int &dir, int &idx, int &len, int64_t &start) const {
        int r = q >> 1;
        if ((q & 1) == 0) { dir = 0; idx = r; len = row_len(r); start = rm_row_start(r); }
        else { dir = 1; idx = n - 1 - r; len = col_len(idx); start = cm_col_start(idx); }
    }
    static __host__ __device__ __forceinline__ int min_(int a, int b) { return a < b ? a : b; }
    static Grid make(int n, int md) {
        Grid g; g.n = n; g.md = md; g.nfull = n - md + 1;
        g.ncells = (int64_t)md * (n - md + 1) + ((int64_t)(md - 1) * md) / 2;
        return g;
    }
};

// Device-resident control block of one ADMM solve.
struct GflCtl {
    double alpha, lam, converge, inflate;
    double presnorm, dresnorm;
    int step, maxsteps, done;
    int rescale;      // pending dual rescale: 0 none, 1 u /= inflate, 2 u *= inflate
    int next_chain;   // work queue head of k_prox (longest chains first)
    int idle;         // set by k_ctl on launches queued after convergence: the trail-block exchange is skipped
    int prox_sel;     // chain prox of the NEXT iteration: 0 taut-string team kernel k_prox, 1 parallel active set k_prox_aset
    int aset_from;    // first step (1-based) that uses k_prox_aset; -1 never
    const double *y, *w;   // inputs of the running solve (the kernels of the captured iteration read them from here)
    int ctas_done;    // CTAs of the running chain-prox kernel that have finished (the last one ends the iteration, fb_ctl.cuh)
};

// What the last CTA of a chain-prox kernel needs to end the iteration itself (on = 0: a separate k_ctl launch follows)
struct GflFold {
    int on, graph, npart;
    int64_t ncells, nz;
    cudaGraphConditionalHandle cond;
};

// Trail-block sharding (one process per GPU): the cells {line in [l0,l1), x in [x0,x1)} of the
// band, packed line by line, that one rank sends to / receives from one peer. dir 0: lines are rows (row-major arrays,
// x = column); dir 1: lines are columns (column-major arrays, x = row).
struct ShardRegion {
    int l0, l1, x0, x1;
    int64_t buf_off;   // first packed cell in the exchange buffer
    int64_t off_idx;   // first entry of this region in the line-offset array
};

// One GFLLibrary strategy object: persistent beta / z / u / alpha (warm start)
// living in HBM. y and w are device pointers to ncells_tri packed (row-major)
// values of ONE dataset in the FULL triangle; for a band they are gathered.
class GflSolver {
public:
    GflSolver(int nbins, unsigned maxdiag, cudaStream_t stream);
    ~GflSolver();
    void reset();                      // cold start (GFLLibraryTriangleImpl::reset)
    // an idle solver continues on another stream (the captured ADMM graph does not depend on the stream it was built on)
    void rebind(cudaStream_t stream) { st_ = stream; }
    bool sharded() const { return sharded_; }
    // runs ADMM to convergence; returns `step` like graph_fused_lasso_weight_warm
    int optimize(const double *d_y_tri, const double *d_w_tri, double lam2, double converge);
    // clamp(+-50) of the solution scattered to the full packed triangle
    // (FusedLassoGaussianEstimator::get + fill_triangle); out has n(n+1)/2 entries
    void get_beta_full(double *d_out_tri);
    // band solves: the same with the nobs-weighted mean of the result subtracted (step_signal's centring) in the same pass
    void get_beta_full_centered(double *d_out_tri, const int32_t *d_nobs_tri, double nobs_total);
    // state import/export in the reference's trail-slot order
    void export_state(double *h_beta, double *h_z, double *h_u) ;
    void import_state(const double *h_beta, const double *h_z, const double *h_u);

    bool trapeze() const { return trapeze_; }
    const Grid &grid() const { return g_; }
    int64_t tsz() const { return 2 * g_.ncells - 2; }
    double alpha = 5.0;                // Settings<GFLLibraryImpl>::get_alpha
    double inflate = 2.0;
    int ninner_max = 30000;
    unsigned counter = 0;
    int last_steps = 0;
    float ms_last = 0.f;               // device time of the last optimize()
    int64_t kernel_launches = 0;

private:
    Grid g_;
    int nbins_;
    bool trapeze_;
    cudaStream_t st_;
    DevBuf<double> beta_, betaT_, zr_, ur_, zc_, uc_, yband_, wband_, fillavg_;
    DevBuf<double> partials_, partials_all_, band_partials_, center_partials_, center_val_;
    // trail-block sharding over the ranks of fb::Comm (inactive: one rank owns every chain)
    bool sharded_ = false;
    int rank_ = 0, world_ = 1;
    int r0_ = 0, r1_ = 0, c0_ = 0, c1_ = 0;          // rows / columns whose chains this rank owns
    std::vector<int> row_begin_, col_begin_;
    std::vector<int64_t> send_cnt_, send_off_, recv_cnt_, recv_off_;   // per rank, in doubles
    DevBuf<int> chain_list_;
    int n_list_ = 0;
    DevBuf<ShardRegion> send_regs_, recv_regs_;
    DevBuf<int64_t> send_lineoff_, recv_lineoff_;
    DevBuf<double> send_buf_, recv_buf_;
    int max_lines_ = 1;
    // FB_SHARD_TIMING development aid: device time per phase of the sharded iteration
    bool shard_timing_ = false;
    std::vector<cudaEvent_t> tev_;
    double tacc_[8] = {};
    int tcount_ = 0;
    void tmark(int k);
    void tflush();
    void setup_shard();
    void exchange(bool zero_u);
    // Direct exchange over peer memory (NVLink): every rank's z arrays have the same layout, so the owner of a chain
    // stores z - u of a crossing cell straight into the peer's array at the same index (one kernel per direction, no
    // pack / ncclSend / ncclRecv / unpack); peers in other processes are reached through CUDA IPC handles, peers in this
    // process (fb_create_multi) through plain pointers. One scalar all-reduce after the stores is the barrier.
    bool p2p_ = false;
    DevBuf<double *> peer_zr_, peer_zc_, peer_part_, peer_beta_;   // per exchange partner (region order): its zr / zc / partials_all_
    bool p2p_ipc_ = false;
    DevBuf<double> bar_;
    void setup_p2p();
    void teardown_p2p();
    void allgather_rows(double *rm_array);
    void allgather_cols(double *cm_array);
    DevBuf<GflCtl> ctl_;
    DevBuf<long long> trace_;         // FB_PROX_TRACE development aid
    GflCtl *h_ctl_ = nullptr;          // pinned
    const double *y_tri_ = nullptr, *w_tri_ = nullptr;
    int n_partials_ = 0;
    int n_sm_ = 148, prox_warps_ = 1, prox_grid_ = 1, prox_stride_ = 0, prox_rlen_ = 0;
    size_t prox_smem_ = 0;
    bool prox_long_ = false;           // reciprocal table shorter than the longest chain
    bool prox_gs_ = false;             // prefix sums of the chains in global scratch instead of shared memory
    DevBuf<double> gscratch_;
    // parallel active-set chain prox (fb_prox_aset.cu): used whenever the longest chain fits one CTA's shared memory;
    // FB_PROX_ASET=0 keeps the taut-string team kernel k_prox (A/B tests)
    std::unique_ptr<AsetPlan> aset_;
    bool use_aset_ = false;
    int aset_from_ = 13;               // cold start: while the string has few knots the team kernel is faster (B200, 2400 bins: up to step ~13)
    bool cold_ = true;                 // no solve yet since reset(): the duals carry no knots
    cudaEvent_t ev0_, ev1_;
    // the whole ADMM loop as ONE graph launch: a WHILE conditional node whose body is one iteration, the condition
    // set on the device by k_ctl (no host polling, no launches queued past convergence)
    bool fold_ctl_ = false;            // FB_CTL_FOLD: see the constructor
    bool use_graph_ = false;
    cudaGraph_t graph_ = nullptr;
    cudaGraphExec_t graph_exec_ = nullptr;
    cudaGraphConditionalHandle cond_ = 0;
    void build_graph();
    void iteration(const double *y, const double *w, int parity);
    void launch_iteration(const double *y, const double *w, bool in_graph);
};

} // namespace fb
