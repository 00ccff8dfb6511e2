// fb_prox_aset.cuh -- parallel active-set chain prox (update_z + update_u + residual sums of one ADMM iteration,
// gfl_graph_fl.c:267-365 with tf_dp_weight, gfl_tf.c:184-321). See fb_prox_aset.cu.
#pragma once
#include "fb_gfl.cuh"
#include "fb_prox_aset_passes.cuh"
#include <vector>

namespace fb {

// One chain of the trail decomposition as a batch sees it.
struct AsChain {
    int64_t g;     // first cell in the direction's packed array (row-major for rows, column-major for columns)
    int L;         // cells
    int q;         // global chain number (Grid::chain): slot of its residual partial sums
    int base;      // first string point of the chain in the batch's shared-memory arrays (multiple of 32)
    int pad;
};
// A batch: consecutive chains of one direction that fit the shared-memory arrays of one CTA together.
struct AsBatch {
    int c0, nch;   // chains [c0, c0 + nch) of the chain table
    int dir;       // 0 rows (beta, z_row, u_row), 1 columns (betaT, z_col, u_col)
    int npts;      // string points in use (multiple of 32)
    int w0;        // first entry of the batch in the word table
};
struct AsetPlan {
    bool usable = false;
    int cap = 0;               // string points per batch (multiple of 32)
    int nbatches = 0, grid = 0;
    int maxpass = 0;
    size_t smem = 0;
    DevBuf<AsChain> chains;
    DevBuf<AsBatch> batches;
    DevBuf<AsWord> words;
    DevBuf<unsigned long long> stats;   // FB_ASET_STATS development aid
};

// Builds the batches for the rows [r0, r1) and columns [c0, c1) of the grid (the whole grid when not sharded).
// usable == false when the longest chain does not fit the shared memory of one CTA (the caller keeps k_prox).
void aset_plan_build(AsetPlan &plan, const Grid &g, int r0, int r1, int c0, int c1, int n_sm);

void aset_prox_launch(const AsetPlan &plan, GflCtl *ctl, const double *beta, const double *betaT, double *zr, double *ur,
                      double *zc, double *uc, double *partials, const GflFold &fold, cudaStream_t st);

} // namespace fb
