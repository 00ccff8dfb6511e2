// fb_gfl.cu -- ADMM kernels of the graph-fused-lasso solve (sm_100a).
//
// One ADMM iteration (gfl_graph_fl.c:143-183) is three launches on one stream:
//   k_beta  closed-form beta update over 32x32 tiles of the triangle, reading the row duals in
//           row-major and the column duals in column-major order (shared-memory tile transpose);
//           writes beta twice: row-major (input of the row chains) and column-major (column chains)
//   k_prox  persistent kernel, two WARPS per chain working from both ends in lock step (static first chain, then a
//           dynamic longest-first queue): stages the prefix sums of beta+u of one chain in shared memory (or, for
//           chains too long for ten of them to fit, in a per-team slot of global scratch), runs the exact 1-D
//           fused-lasso prox as a warp-cooperative taut string (same unique minimiser as tf_dp_weight,
//           gfl_tf.c:184-321: 160 candidate end points per round, slope bounds by warp scans, first tube violation
//           by ballot), then streams z, u += beta - z (gfl_graph_fl.c:306-327) and the residual sums
//           (gfl_graph_fl.c:329-365) with coalesced accesses
//   k_ctl   single block: fixed-order reduction of the per-chain partials, convergence test,
//           varying-penalty rule every 50 steps (gfl_graph_fl.c:163-181)
// The loop itself is ONE graph launch (WHILE conditional node, condition set by k_ctl); with collectives between
// the kernels (trail-block sharding over NCCL, fb_comm.cuh) or FB_NO_GRAPH the host polls the control block instead.
#include "fb_gfl.cuh"
#include "fb_comm.cuh"
#include "fb_taut.cuh"
#include "fb_prox_aset.cuh"
#include "fb_ctl.cuh"
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cstring>
#include <unistd.h>

namespace fb {

namespace {

constexpr int TILE = 32;
constexpr int CTL_THREADS = 1024;
constexpr int POLL_BATCH = 8;
constexpr int PROX_MAX_TEAMS = 10;    // two warps per chain; 640 threads leave 96 registers per thread (instruction bound:
                                      // more teams with fewer registers are slower at 2400 bins, 261 -> 299 us with 12)
constexpr int PROX_GS_TEAMS = 15;     // chains in global scratch pay L1/L2 latency in every round: 15 teams (64 registers,
                                      // named barriers 1..15) are faster there (12 000 bins: 17.4 -> 13.3 ms per iteration)
using namespace taut;

// --------------------------------------------------------------------------
// beta update (update_beta_weight, gfl_graph_fl.c:241-264)
// --------------------------------------------------------------------------
#ifdef FB_BETA_V1
__global__ void __launch_bounds__(TILE * 8)
k_beta(Grid g, const GflCtl *__restrict__ ctl, const double *__restrict__ y, const double *__restrict__ w,
       const double *__restrict__ zr, double *__restrict__ ur, const double *__restrict__ zc,
       double *__restrict__ uc, double *__restrict__ beta, double *__restrict__ betaT,
       const int r0, const int r1, const int c0, const int c1) {
    if (ctl->done) return;
    // the grid covers the band only: tile column = tile row + blockIdx.x (a full-width grid at chromosome scale launched
    // 2.4 million CTAs that returned at once: 2.4 ms per iteration at 50 000 bins with a 2000-wide band)
    const int i0 = blockIdx.y * TILE, j0 = (blockIdx.y + blockIdx.x) * TILE;
    if (j0 >= g.n) return;
    // tile must intersect the band 0 <= j-i < md
    if (j0 + TILE - 1 < i0 || j0 - (i0 + TILE - 1) >= g.md) return;
    // ... and hold a cell of a chain this rank owns (rows [r0, r1) or columns [c0, c1); everything when not sharded)
    if (!(i0 < r1 && i0 + TILE > r0) && !(j0 < c1 && j0 + TILE > c0)) return;
    __shared__ double s_tcol[TILE][TILE + 1]; // z_col - u_col, indexed [j][i]
    __shared__ double s_beta[TILE][TILE + 1]; // indexed [i][j]
    const int tx = threadIdx.x, ty = threadIdx.y;
    const double alpha = ctl->alpha, inflate = ctl->inflate;
    const int mode = ctl->rescale;
    if (!y) { y = ctl->y; w = ctl->w; }   // captured iteration: the inputs of the current solve

    // phase A: column duals, coalesced along i in column-major storage
    for (int jj = ty; jj < TILE; jj += 8) {
        const int j = j0 + jj, i = i0 + tx;
        double t = 0.0;
        if (j < g.n && i <= j && j - i < g.md && j >= 1) {
            const int64_t idx = g.cm(i, j);
            double u = uc[idx];
            if (mode == 1) { u = u / inflate; uc[idx] = u; }
            else if (mode == 2) { u = u * inflate; uc[idx] = u; }
            t = zc[idx] - u;
        }
        s_tcol[jj][tx] = t;
    }
    __syncthreads();
    // phase B: row duals + beta, coalesced along j in row-major storage
    for (int ii = ty; ii < TILE; ii += 8) {
        const int i = i0 + ii, j = j0 + tx;
        double b = 0.0;
        if (j < g.n && i <= j && j - i < g.md) {
            const int64_t idx = g.rm(i, j);
            const bool has_row = (i <= g.n - 2), has_col = (j >= 1);
            double r = 0.0;
            if (has_row) {
                double u = ur[idx];
                if (mode == 1) { u = u / inflate; ur[idx] = u; }
                else if (mode == 2) { u = u * inflate; ur[idx] = u; }
                r += zr[idx] - u;
            }
            if (has_col) r += s_tcol[tx][ii];
            const int nz = (has_row ? 1 : 0) + (has_col ? 1 : 0);
            const double wi = w[idx];
            b = (2.0 * y[idx] * wi + alpha * r) / (2.0 * wi + alpha * nz);
            beta[idx] = b;
        }
        s_beta[ii][tx] = b;
    }
    __syncthreads();
    // phase C: column-major copy of beta, coalesced along i
    for (int jj = ty; jj < TILE; jj += 8) {
        const int j = j0 + jj, i = i0 + tx;
        if (j < g.n && i <= j && j - i < g.md && j >= 1) betaT[g.cm(i, j)] = s_beta[tx][jj];
    }
}
#else
// Every global load of the tile is issued before the first use (24 per thread in flight): ncu showed the first version
// stalled on the long scoreboard 61 % of the time with a handful of loads per warp outstanding.
#ifndef FB_BETA_MINB
#define FB_BETA_MINB 1
#endif
__global__ void __launch_bounds__(TILE * 8, FB_BETA_MINB)
k_beta(Grid g, const GflCtl *__restrict__ ctl, const double *__restrict__ y, const double *__restrict__ w,
       const double *__restrict__ zr, double *__restrict__ ur, const double *__restrict__ zc,
       double *__restrict__ uc, double *__restrict__ beta, double *__restrict__ betaT,
       const int r0, const int r1, const int c0, const int c1) {
    if (ctl->done) return;
    // the grid covers the band only: tile column = tile row + blockIdx.x (a full-width grid at chromosome scale launched
    // 2.4 million CTAs that returned at once: 2.4 ms per iteration at 50 000 bins with a 2000-wide band)
    const int i0 = blockIdx.y * TILE, j0 = (blockIdx.y + blockIdx.x) * TILE;
    if (j0 >= g.n) return;
    // tile must intersect the band 0 <= j-i < md
    if (j0 + TILE - 1 < i0 || j0 - (i0 + TILE - 1) >= g.md) return;
    // ... and hold a cell of a chain this rank owns (rows [r0, r1) or columns [c0, c1); everything when not sharded)
    if (!(i0 < r1 && i0 + TILE > r0) && !(j0 < c1 && j0 + TILE > c0)) return;
    __shared__ double s_tcol[TILE][TILE + 1]; // z_col - u_col, indexed [j][i]
    __shared__ double s_beta[TILE][TILE + 1]; // indexed [i][j]
    const int tx = threadIdx.x, ty = threadIdx.y;
    const double alpha = ctl->alpha, inflate = ctl->inflate;
    const int mode = ctl->rescale;
    if (!y) { y = ctl->y; w = ctl->w; }   // captured iteration: the inputs of the current solve
    constexpr int NQ = TILE / 8;
    // column-major side: cell (i0 + tx, j0 + ty + 8 q); row-major side: cell (i0 + ty + 8 q, j0 + tx)
    int64_t cidx[NQ], ridx[NQ];
    bool cok[NQ], rok[NQ], hrow[NQ];
    double vuc[NQ], vzc[NQ], vur[NQ], vzr[NQ], vw[NQ], vy[NQ];
#pragma unroll
    for (int q = 0; q < NQ; ++q) {
        const int j = j0 + ty + 8 * q, i = i0 + tx;
        cok[q] = j < g.n && i <= j && j - i < g.md && j >= 1;
        cidx[q] = cok[q] ? g.cm(i, j) : 0;
        const int i2 = i0 + ty + 8 * q, j2 = j0 + tx;
        rok[q] = j2 < g.n && i2 <= j2 && j2 - i2 < g.md;
        ridx[q] = rok[q] ? g.rm(i2, j2) : 0;
        hrow[q] = rok[q] && i2 <= g.n - 2;
    }
#pragma unroll
    for (int q = 0; q < NQ; ++q) {
        vuc[q] = cok[q] ? uc[cidx[q]] : 0.0; vzc[q] = cok[q] ? zc[cidx[q]] : 0.0;
        vur[q] = hrow[q] ? ur[ridx[q]] : 0.0; vzr[q] = hrow[q] ? zr[ridx[q]] : 0.0;
        vw[q] = rok[q] ? w[ridx[q]] : 0.0; vy[q] = rok[q] ? y[ridx[q]] : 0.0;
    }
    // phase A: column duals into the transposition tile
#pragma unroll
    for (int q = 0; q < NQ; ++q) {
        double t = 0.0;
        if (cok[q]) {
            double u = vuc[q];
            if (mode == 1) { u = u / inflate; uc[cidx[q]] = u; }
            else if (mode == 2) { u = u * inflate; uc[cidx[q]] = u; }
            t = vzc[q] - u;
        }
        s_tcol[ty + 8 * q][tx] = t;
    }
    __syncthreads();
    // phase B: row duals + beta
#pragma unroll
    for (int q = 0; q < NQ; ++q) {
        const int ii = ty + 8 * q;
        double b = 0.0;
        if (rok[q]) {
            const bool has_col = (j0 + tx >= 1);
            double r = 0.0;
            if (hrow[q]) {
                double u = vur[q];
                if (mode == 1) { u = u / inflate; ur[ridx[q]] = u; }
                else if (mode == 2) { u = u * inflate; ur[ridx[q]] = u; }
                r += vzr[q] - u;
            }
            if (has_col) r += s_tcol[tx][ii];
            const int nz = (hrow[q] ? 1 : 0) + (has_col ? 1 : 0);
            const double wi = vw[q];
            b = (2.0 * vy[q] * wi + alpha * r) / (2.0 * wi + alpha * nz);
            beta[ridx[q]] = b;
        }
        s_beta[ii][tx] = b;
    }
    __syncthreads();
    // phase C: column-major copy of beta, coalesced along i
#pragma unroll
    for (int q = 0; q < NQ; ++q)
        if (cok[q]) betaT[cidx[q]] = s_beta[tx][ty + 8 * q];
}
#endif

// --------------------------------------------------------------------------
// chain prox: one warp per chain, exact taut string. For the chain v_0..v_{L-1} (v = beta + u) and
// tube radius lam' = lam/(alpha/2) it finds the unique minimiser of
//   1/2 sum (x_k - v_k)^2 + lam' sum |x_{k+1} - x_k|
// i.e. exactly what tf_dp_weight(n, y = beta+u, w = alpha/2, lam) returns (update_z, gfl_graph_fl.c:267-304).
// S[k] = sum_{i<k} (v_i - c) lives in shared memory. A segment of the solution starting at k0 from height
// s0 = S[k0] + a (a = 0 at the chain start, +-lam' when it leaves the upper / lower tube edge) may end at k
// with any slope in [lo_k, hi_k] = [(S[k] - lam'_k - s0)/(k-k0), (S[k] + lam'_k - s0)/(k-k0)], lam'_L = 0.
// The string bends where lo_k exceeds the running minimum of hi (it touches the upper edge there) or hi_k
// drops below the running maximum of lo. 32 candidates k are examined per round: running bounds by warp
// scans, first violation by ballot. Finished segments overwrite S in place with their value c + slope.
// --------------------------------------------------------------------------

// --------------------------------------------------------------------------
// Two warps per chain ("team"), one working from each end. The 1-D fused-lasso minimiser is unique, and the taut
// string of the mirrored chain is the mirror image of the taut string, so the LEFT warp determines the knots in
// increasing order while the RIGHT warp determines them in decreasing order (on S'(k) = S[L] - S[L-k], the prefix
// sums of the reversed chain). A knot found by the other warp is a point the string is known to pass through:
// it becomes this warp's fixed end point (offset aEnd = -/+ lam or 0 instead of the tube). The warps advance in
// lock step, one round each between two named barriers, until the finished parts meet.
//
// One round examines 32*B consecutive candidates, lane l owning the B consecutive ones starting at
// p + 1 + l*B (B odd: conflict-free 8-byte shared-memory accesses). Violations against the lane-local running
// bounds are found before the warp scan, those against earlier lanes / earlier rounds after it with B
// independent compares. TAIL = false requires p + 32*B < kEnd (every candidate is an interior point).
// --------------------------------------------------------------------------
__device__ __forceinline__ void team_sync(int bar) { asm volatile("bar.sync %0, 64;" ::"r"(bar) : "memory"); }


// Lock-step taut string of one chain by the two warps of a team. S holds the prefix sums S[0..L] on entry; on
// return cell i of the solution is at S[i] for i < split (written by the left warp) and at S[i+1] for i >= split.
template <bool LONG>
__device__ __forceinline__ int taut_string_team(double *__restrict__ S, volatile int *__restrict__ ti, const int L,
                                                const double lam, const double c, const double *__restrict__ R, const int rl,
                                                const int side, const int lane, const int bar, int &nrounds) {
    Frame f; f.S = S; f.L = L; f.dir = side; f.SL = S[L];
    Seg sg; sg.k0 = 0; sg.s0 = 0.0; seg_reset(sg);
    int dL = 0, dR = L;      // cells [0, dL) and [dR, L) are finished (original coordinates)
    int aL = 0, aR = 0;      // sign of the tube edge each warp's current segment starts from (own frame)
    while (dL < dR) {
        const int kEnd = side ? L - dL : dR;
        const double aEnd = -(double)(side ? aL : aR) * lam;   // the sign flips under mirroring
        if (kEnd <= sg.p) seg_reset(sg);                               // the other warp's knot moved inside the scanned range
        int isup = 0, kend = 0;
        double slope = 0.0;
        const bool viol = taut_round<LONG>(f, R, rl, lam, sg, kEnd, aEnd, lane, isup, slope, kend);
        sg.p += 32 * PBR;
        ++nrounds;
        int status = 0, asign = 0;
        double val = 0.0, s0n = 0.0;
        if (viol) {
            status = 1; asign = isup ? 1 : -1;
            val = c + slope;
            s0n = f.at(kend) + (isup ? lam : -lam);
        } else if (sg.p >= kEnd) { // the fixed end point was examined without violation: the segment runs up to it
            status = 1; kend = kEnd;
            val = c + ((f.at(kEnd) - sg.s0) + aEnd) * recip<LONG>(R, rl, kEnd - sg.k0);
        }
        // mailbox: one packed word per warp, double buffered by round parity, so ONE barrier per round is enough
        // (the fills below touch only cells beyond the knots the other warp may still read)
        const int slot = (nrounds & 1) * 2;
        if (lane == 0) ti[slot + side] = ((side ? L - kend : kend) << 3) | ((asign + 1) << 1) | status;
        team_sync(bar);
        const int ow = ti[slot + 1 - side];
        const int ost = ow & 1, oe = ow >> 3, oa = ((ow >> 1) & 3) - 1;
        int eL = dL, eR = dR;
        if (side ? ost : status) eL = side ? oe : kend;
        if (side ? status : ost) eR = side ? L - kend : oe;
        if (eR < eL) eR = eL;                                    // the left warp wins an overlap
        if (status) {
            if (side == 0) {
#pragma unroll 1
                for (int i = dL + lane; i < eL; i += 32) S[i] = val;
            } else {
#pragma unroll 1
                for (int i = eR + lane; i < dR; i += 32) S[i + 1] = val;
            }
            sg.k0 = kend; sg.s0 = s0n; seg_reset(sg);
            if (side) aR = asign; else aL = asign;
        }
        if (ost) { if (side) aL = oa; else aR = oa; }
        dL = eL; dR = eR;
    }
    team_sync(bar);      // every segment value is in S before either warp writes its half back
    return dL;
}

// GS: the prefix sums of a chain live in a per-team slot of global scratch (L1/L2 resident) instead of shared memory,
// which then holds only the reciprocal table: ten teams per SM whatever the chain length.
template <bool LONG, bool GS>
__global__ void __launch_bounds__(64 * (GS ? PROX_GS_TEAMS : PROX_MAX_TEAMS), 1)
k_prox(Grid g, GflCtl *__restrict__ ctl, const double *__restrict__ beta, const double *__restrict__ betaT,
       double *__restrict__ zr, double *__restrict__ ur, double *__restrict__ zc, double *__restrict__ uc,
       double *__restrict__ partials, int stride, int rlen, long long *__restrict__ trace,
       const int *__restrict__ list, const int nlist, double *__restrict__ gscratch, const GflFold fold) {
    if (ctl->done || ctl->prox_sel) return;   // prox_sel: this iteration belongs to k_prox_aset
    __shared__ int s_last;
    extern __shared__ double smem[];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int team = wid >> 1, side = wid & 1, bar = 1 + team;
    double *__restrict__ R = smem;                                   // R[m] = 1/m, m = 1..rlen-1
    double *__restrict__ S = GS ? gscratch + ((size_t)blockIdx.x * (blockDim.x >> 6) + team) * (size_t)stride
                                : smem + rlen + (size_t)team * stride;
    volatile int *ti = reinterpret_cast<volatile int *>(GS ? smem + rlen + 8 * team : S + (stride - 8)); // team mailbox: 16 ints
    for (int m = threadIdx.x; m < rlen; m += blockDim.x) R[m] = m ? 1.0 / (double)m : 0.0;
    __syncthreads();
    const double alpha = ctl->alpha;
    const double lam = ctl->lam / (alpha / 2.0);
    const int nchains = list ? nlist : g.nchains();   // sharded: the chains of this rank, longest first
    // the first chain of every team is assigned statically (team-major, so each SM gets a mix of lengths): no burst of
    // ~1500 atomics on one address at kernel start; later chains come from the shared counter
    const int nteams_all = gridDim.x * (blockDim.x >> 6);
    if (side == 0 && lane == 0) ti[8] = team * gridDim.x + blockIdx.x;
    team_sync(bar);
    int qi = ti[8];
    while (qi < nchains) {
        const int q = list ? list[qi] : qi;
        int qnext = 0;
        if (side == 0 && lane == 0) qnext = nteams_all + atomicAdd(&ctl->next_chain, 1); // published after this chain: latency hidden
        long long t0 = 0, t1 = 0, t2 = 0;
        int nrounds = 0;
        if (trace) t0 = clock64();
        int dir, idx, L; int64_t start;
        g.chain(q, dir, idx, L, start);
        const double *__restrict__ B = (dir == 0 ? beta : betaT) + start;
        double *__restrict__ Z = (dir == 0 ? zr : zc) + start;
        double *__restrict__ U = (dir == 0 ? ur : uc) + start;
        const int half = (L + 1) >> 1;
        const int lo_ = side ? half : 0, hi_ = side ? L : half;    // the cells this warp stages and writes back
        const bool copy = (lam == 0.0 || L == 1);                   // tf_dp_weight copies y (gfl_tf.c:210-215)
        double c = 0.0;
        if (!copy && side) c = __ldcg(B) + __ldcg(U);
        // stage v = beta + u at S[k+1], coalesced, 8 loads per array in flight
        int ps = lo_;
#pragma unroll 1
        for (; ps + 256 <= hi_; ps += 256) {   // whole batches: no predicates, one address per array
            const double *__restrict__ bp = B + ps + lane, *__restrict__ up = U + ps + lane;
            double *__restrict__ sp = S + ps + lane + 1;
            double vb[8], vu[8];
#pragma unroll
            for (int b = 0; b < 8; ++b) { vb[b] = __ldcg(bp + 32 * b); vu[b] = __ldcg(up + 32 * b); }
#pragma unroll
            for (int b = 0; b < 8; ++b) sp[32 * b] = vb[b] + vu[b];
        }
        if (ps < hi_) {
            double vb[8], vu[8];
#pragma unroll
            for (int b = 0; b < 8; ++b) { const int k = ps + 32 * b + lane; vb[b] = k < hi_ ? __ldcg(B + k) : 0.0; vu[b] = k < hi_ ? __ldcg(U + k) : 0.0; }
#pragma unroll
            for (int b = 0; b < 8; ++b) { const int k = ps + 32 * b + lane; if (k < hi_) S[k + 1] = vb[b] + vu[b]; }
        }
        __syncwarp();
        int split = 0; // cell i of the solution at S[i] for i < split, else at S[i + 1]
        if (!copy) {
            if (!side) c = S[1];
            // in-place prefix sums of (v_i - c) over this warp's half: lane-local runs of PB, one warp scan per 32*PB
            double carry = 0.0;
            int pp = lo_;
#pragma unroll 1
            for (; pp + PW <= hi_; pp += PW) {     // whole blocks: no predicates
                double *__restrict__ sp = S + pp + lane * PB + 1;
                double t[PB];
                double run = 0.0;
#pragma unroll
                for (int b = 0; b < PB; ++b) { run += sp[b] - c; t[b] = run; }
                const double incl = scan_sum_incl(run, lane);
                const double off = (incl - run) + carry;
#pragma unroll
                for (int b = 0; b < PB; ++b) sp[b] = t[b] + off;
                carry += __shfl_sync(FULL, incl, 31);
            }
            if (pp < hi_) {
                const int kb = pp + lane * PB;
                double t[PB];
                double run = 0.0;
#pragma unroll
                for (int b = 0; b < PB; ++b) { const int k = kb + b; run += (k < hi_) ? S[k + 1] - c : 0.0; t[b] = run; }
                const double incl = scan_sum_incl(run, lane);
                const double off = (incl - run) + carry;
#pragma unroll
                for (int b = 0; b < PB; ++b) { const int k = kb + b; if (k < hi_) S[k + 1] = t[b] + off; }
            }
            if (!side && lane == 0) S[0] = 0.0;
            team_sync(bar);
            if (side) { // the right half continues the left half's sums
                const double off = S[half];
#pragma unroll 1
                for (int p = lo_; p < hi_; p += 128) {
                    double t[4];
#pragma unroll
                    for (int b = 0; b < 4; ++b) { const int k = p + 32 * b + lane; t[b] = k < hi_ ? S[k + 1] : 0.0; }
#pragma unroll
                    for (int b = 0; b < 4; ++b) { const int k = p + 32 * b + lane; if (k < hi_) S[k + 1] = t[b] + off; }
                }
            }
            team_sync(bar);
            if (trace) t1 = clock64();
            split = taut_string_team<LONG>(S, ti, L, lam, c, R, rlen, side, lane, bar, nrounds);
            if (trace) t2 = clock64();
        } else {
            team_sync(bar);
        }
        // u += beta - z, residual sums (gfl_graph_fl.c:306-365)
        double ap = 0.0, ad = 0.0;
        int po = lo_;
#pragma unroll 1
        for (; po + 256 <= hi_; po += 256) {   // whole batches
            const double *__restrict__ bp = B + po + lane;
            double *__restrict__ up = U + po + lane, *__restrict__ zp = Z + po + lane;
            const int k0 = po + lane;
            double b8[8], u8[8], z8[8];
#pragma unroll
            for (int b = 0; b < 8; ++b) { b8[b] = __ldcg(bp + 32 * b); u8[b] = __ldcg(up + 32 * b); z8[b] = __ldcg(zp + 32 * b); }
#pragma unroll
            for (int b = 0; b < 8; ++b) {
                const int k = k0 + 32 * b;
                const double z = S[k < split ? k : k + 1];
                const double r = b8[b] - z;
                __stcg(up + 32 * b, u8[b] + r);
                __stcg(zp + 32 * b, z);
                ap += r * r;
                const double dd = alpha * (z - z8[b]);
                ad += dd * dd;
            }
        }
        if (po < hi_) {
            double b8[8], u8[8], z8[8];
#pragma unroll
            for (int b = 0; b < 8; ++b) {
                const int k = po + 32 * b + lane;
                b8[b] = k < hi_ ? __ldcg(B + k) : 0.0; u8[b] = k < hi_ ? __ldcg(U + k) : 0.0; z8[b] = k < hi_ ? __ldcg(Z + k) : 0.0;
            }
#pragma unroll
            for (int b = 0; b < 8; ++b) {
                const int k = po + 32 * b + lane;
                if (k < hi_) {
                    const double z = S[k < split ? k : k + 1];
                    const double r = b8[b] - z;
                    __stcg(U + k, u8[b] + r);
                    __stcg(Z + k, z);
                    ap += r * r;
                    const double dd = alpha * (z - z8[b]);
                    ad += dd * dd;
                }
            }
        }
        ap = warp_sum(ap); ad = warp_sum(ad);
        if (lane == 0) { partials[2 * (2 * q + side)] = ap; partials[2 * (2 * q + side) + 1] = ad; }
        if (trace && lane == 0 && side == 0) {
            unsigned smid; asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
            trace[4 * q] = t1 - t0; trace[4 * q + 1] = t2 - t1; trace[4 * q + 2] = clock64() - t2;
            trace[4 * q + 3] = ((long long)smid << 48) | (long long)nrounds;
        }
        if (side == 0 && lane == 0) ti[8] = qnext;
        team_sync(bar); // both warps are done with S; the next chain is published
        qi = ti[8];
    }
    // the last CTA to finish ends the iteration (residual norms, convergence, penalty): no k_ctl launch
    if (fold.on) {
        __syncthreads();     // every team of this CTA is done with the shared memory
        ctl_fold(ctl, fold, partials, &s_last, smem);
    }
}

// --------------------------------------------------------------------------
// trail-block exchange: pack z - u of the cells another rank's chains cross (with the pending dual rescale applied,
// exactly the expression k_beta evaluates on the owner), or unpack the received terms into z (u = 0 there).
// grid (lines, peers); one block walks whole lines.
// --------------------------------------------------------------------------
template <int DIR, bool PACK>
__global__ void __launch_bounds__(256)
k_exchange(Grid g, const GflCtl *__restrict__ ctl, const ShardRegion *__restrict__ regs, const int64_t *__restrict__ lineoff,
           double *__restrict__ z, double *__restrict__ u, double *__restrict__ buf, const int zero_u) {
    if (ctl->idle) return;
    const ShardRegion r = regs[2 * blockIdx.y + DIR];
    const int mode = ctl->rescale;
    const double inflate = ctl->inflate;
    for (int line = r.l0 + blockIdx.x; line < r.l1; line += gridDim.x) {
        int lo, hi; int64_t base;
        if (DIR == 0) {
            lo = max(r.x0, line); hi = min(r.x1, line + g.row_len(line));
            base = g.rm_row_start(line) + (lo - line);
        } else {
            const int f = g.col_first_row(line);
            lo = max(r.x0, f); hi = min(r.x1, line + 1);
            base = g.cm_col_start(line) + (lo - f);
        }
        const int64_t off = r.buf_off + lineoff[r.off_idx + (line - r.l0)];
        for (int x = threadIdx.x; x < hi - lo; x += blockDim.x) {
            if (PACK) {
                double uu = u[base + x];
                if (mode == 1) uu = uu / inflate; else if (mode == 2) uu = uu * inflate;
                buf[off + x] = z[base + x] - uu;
            } else {
                z[base + x] = buf[off + x];
                if (zero_u) u[base + x] = 0.0;
            }
        }
    }
}

// The same terms stored straight into the peers' arrays (same layout on every rank => same index): grid (lines, peers)
template <int DIR>
__global__ void __launch_bounds__(256)
k_push(Grid g, const GflCtl *__restrict__ ctl, const ShardRegion *__restrict__ regs, const double *__restrict__ z,
       const double *__restrict__ u, double *const *__restrict__ peer) {
    if (ctl->idle) return;
    const ShardRegion r = regs[2 * blockIdx.y + DIR];
    double *__restrict__ dst = peer[blockIdx.y];
    const int mode = ctl->rescale;
    const double inflate = ctl->inflate;
    for (int line = r.l0 + blockIdx.x; line < r.l1; line += gridDim.x) {
        int lo, hi; int64_t base;
        if (DIR == 0) {
            lo = max(r.x0, line); hi = min(r.x1, line + g.row_len(line));
            base = g.rm_row_start(line) + (lo - line);
        } else {
            const int f = g.col_first_row(line);
            lo = max(r.x0, f); hi = min(r.x1, line + 1);
            base = g.cm_col_start(line) + (lo - f);
        }
        for (int x = threadIdx.x; x < hi - lo; x += blockDim.x) {
            double uu = u[base + x];
            if (mode == 1) uu = uu / inflate; else if (mode == 2) uu = uu * inflate;
            dst[base + x] = z[base + x] - uu;
        }
    }
}
// the residual sums of this rank's chains (4 doubles per chain) into every rank's full array: what the all-reduce of the
// zero-padded arrays produced, without adding anything (3.2 MB took NCCL 4 ms per iteration on 8 GPUs at 50 000 bins)
__global__ void __launch_bounds__(256)
k_push_partials(const GflCtl *__restrict__ ctl, const int *__restrict__ list, int nlist, const double *__restrict__ partials,
                double *__restrict__ all_local, double *const *__restrict__ peer, int npeers) {
    if (ctl->idle) return;
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= nlist) return;
    const int q = list[k];
    const double2 a = *reinterpret_cast<const double2 *>(partials + 4 * (size_t)q);
    const double2 b = *reinterpret_cast<const double2 *>(partials + 4 * (size_t)q + 2);
    *reinterpret_cast<double2 *>(all_local + 4 * (size_t)q) = a;
    *reinterpret_cast<double2 *>(all_local + 4 * (size_t)q + 2) = b;
    for (int p = 0; p < npeers; ++p) {
        double *dst = peer[p];
        *reinterpret_cast<double2 *>(dst + 4 * (size_t)q) = a;
        *reinterpret_cast<double2 *>(dst + 4 * (size_t)q + 2) = b;
    }
}

// the solution of this rank's row block into every peer's beta (the all-gather at the end of a solve, by peer stores)
__global__ void __launch_bounds__(256)
k_push_rows(const double *__restrict__ src, int64_t b, int64_t e, double *const *__restrict__ peer, int npeers) {
    for (int64_t i = b + (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < e; i += (int64_t)gridDim.x * blockDim.x) {
        const double v = src[i];
        for (int p = 0; p < npeers; ++p) peer[p][i] = v;
    }
}

// the duals of the cells whose chains another rank owns are zero here (their z holds z - u of the owner)
template <int DIR>
__global__ void __launch_bounds__(256)
k_zero_remote_u(Grid g, const ShardRegion *__restrict__ regs, double *__restrict__ u) {
    const ShardRegion r = regs[2 * blockIdx.y + DIR];
    for (int line = r.l0 + blockIdx.x; line < r.l1; line += gridDim.x) {
        int lo, hi; int64_t base;
        if (DIR == 0) {
            lo = max(r.x0, line); hi = min(r.x1, line + g.row_len(line));
            base = g.rm_row_start(line) + (lo - line);
        } else {
            const int f = g.col_first_row(line);
            lo = max(r.x0, f); hi = min(r.x1, line + 1);
            base = g.cm_col_start(line) + (lo - f);
        }
        for (int x = threadIdx.x; x < hi - lo; x += blockDim.x) u[base + x] = 0.0;
    }
}

// --------------------------------------------------------------------------
// residual norms, convergence, varying penalty (gfl_graph_fl.c:159-181)
// --------------------------------------------------------------------------
template <bool GRAPH>
__global__ void __launch_bounds__(CTL_THREADS)
k_ctl(GflCtl *ctl, const double *__restrict__ partials, int npart, int64_t ncells, int64_t nz, cudaGraphConditionalHandle cond) {
    if (ctl->done) { if (threadIdx.x == 0) { ctl->idle = 1; if (GRAPH) cudaGraphSetConditional(cond, 0); } return; }
    __shared__ double sm[2 * CTL_VT];
    ctl_tail(ctl, partials, npart, ncells, nz, cond, GRAPH, sm);
}

// applies a pending dual rescale after the last iteration of a solve
// The control block to the host by stores from an SM into mapped pinned memory: a cudaMemcpy of these few bytes would
// queue in the device-to-host copy engine behind any bulk download another stream has in flight (the early output
// columns of Model::run_signal: 4 solves x up to 6 ms of waiting).
__global__ void k_export_ctl(const GflCtl *__restrict__ ctl, GflCtl *__restrict__ host_ctl) {
    static_assert(sizeof(GflCtl) % 4 == 0, "word copy");
    const unsigned *src = reinterpret_cast<const unsigned *>(ctl);
    volatile unsigned *dst = reinterpret_cast<volatile unsigned *>(host_ctl);
    for (unsigned k = threadIdx.x; k < sizeof(GflCtl) / 4; k += blockDim.x) dst[k] = src[k];
    __threadfence_system();
}

__global__ void k_rescale_u(double *__restrict__ u, int64_t n, int mode, double inflate) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) u[i] = (mode == 1) ? u[i] / inflate : u[i] * inflate;
}

// band gather / scatter (extract_trapeze, fill_triangle: GFLLibraryTrapezeImpl.cpp:93-131)
__global__ void k_band_gather(Grid g, int nbins, const double *__restrict__ tri, double *__restrict__ band) {
    const int i = blockIdx.y;
    const int len = g.row_len(i);
    const int64_t src = rm_start(i, nbins), dst = g.rm_row_start(i);
    for (int d = blockIdx.x * blockDim.x + threadIdx.x; d < len; d += gridDim.x * blockDim.x) band[dst + d] = tri[src + d];
}
// sum of w*y and w over the cells outside the band, per row partials
__global__ void k_band_outside_sums(Grid g, int nbins, const double *__restrict__ y, const double *__restrict__ w,
                                    double *__restrict__ partial, const int sh_rank, const int sh_world) {
    const int i = blockIdx.x;
    if ((i % sh_world) != sh_rank) {     // a row of another GPU of the communicator: an exact zero in the all-reduce
        if (threadIdx.x == 0) { partial[2 * i] = 0.0; partial[2 * i + 1] = 0.0; }
        return;
    }
    __shared__ double s_a[256], s_b[256];
    double a = 0.0, b = 0.0;
    const int64_t src = rm_start(i, nbins);
    for (int d = g.md + threadIdx.x; i + d < nbins; d += blockDim.x) { a += w[src + d] * y[src + d]; b += w[src + d]; }
    s_a[threadIdx.x] = a; s_b[threadIdx.x] = b;
    __syncthreads();
    for (int o = 128; o > 0; o >>= 1) {
        if ((int)threadIdx.x < o) { s_a[threadIdx.x] += s_a[threadIdx.x + o]; s_b[threadIdx.x] += s_b[threadIdx.x + o]; }
        __syncthreads();
    }
    if (threadIdx.x == 0) { partial[2 * i] = s_a[0]; partial[2 * i + 1] = s_b[0]; }
}
// the per-row sums in a fixed shape: 1024 threads stride over the rows, then a tree (a single thread walking 50 000 rows
// took 3.2 ms per solve at chromosome scale)
__global__ void __launch_bounds__(1024)
k_band_avg(int nrows, const double *__restrict__ partial, double *__restrict__ avg) {
    __shared__ double s_a[1024], s_b[1024];
    double a = 0.0, b = 0.0;
    for (int i = threadIdx.x; i < nrows; i += 1024) { a += partial[2 * i]; b += partial[2 * i + 1]; }
    s_a[threadIdx.x] = a; s_b[threadIdx.x] = b;
    __syncthreads();
    for (int o = 512; o > 0; o >>= 1) {
        if ((int)threadIdx.x < o) { s_a[threadIdx.x] += s_a[threadIdx.x + o]; s_b[threadIdx.x] += s_b[threadIdx.x + o]; }
        __syncthreads();
    }
    if (threadIdx.x == 0) avg[0] = s_a[0] / s_b[0];
}
// Centring of a band solution (step_signal, fast_binless.ipp:33-44: beta -= sum(beta nobs) / sum(nobs)) without a pass over
// the far region: there beta is one value v, so sum(beta nobs) = sum_band clamp(beta) nobs + v (sum(nobs) - sum_band nobs).
// Per (row, quarter) partial sums over the band, a fixed-shape final sum, then ONE pass writes clamp(beta | v) - centre.
__global__ void __launch_bounds__(256)
k_band_center_partial(Grid g, int nbins, const double *__restrict__ beta, const int32_t *__restrict__ nobs_tri, double clampv,
                      double *__restrict__ partial) {
    __shared__ double s_a[256], s_b[256];
    const int i = blockIdx.y;
    const int len = g.row_len(i);
    const int64_t src = g.rm_row_start(i), tri = rm_start(i, nbins);
    double a = 0.0, b = 0.0;
    for (int d = blockIdx.x * blockDim.x + threadIdx.x; d < len; d += gridDim.x * blockDim.x) {
        const double no = (double)nobs_tri[tri + d];
        a += fmin(clampv, fmax(-clampv, beta[src + d])) * no; b += no;
    }
    s_a[threadIdx.x] = a; s_b[threadIdx.x] = b;
    __syncthreads();
    for (int o = 128; o > 0; o >>= 1) {
        if ((int)threadIdx.x < o) { s_a[threadIdx.x] += s_a[threadIdx.x + o]; s_b[threadIdx.x] += s_b[threadIdx.x + o]; }
        __syncthreads();
    }
    if (threadIdx.x == 0) { const int64_t q = (int64_t)i * gridDim.x + blockIdx.x; partial[2 * q] = s_a[0]; partial[2 * q + 1] = s_b[0]; }
}
__global__ void __launch_bounds__(1024)
k_band_center_final(int64_t npart, const double *__restrict__ partial, const double *__restrict__ fill, double clampv, double nobs_total,
                    double *__restrict__ centre) {
    __shared__ double s_a[1024], s_b[1024];
    double a = 0.0, b = 0.0;
    for (int64_t q = threadIdx.x; q < npart; q += 1024) { a += partial[2 * q]; b += partial[2 * q + 1]; }
    s_a[threadIdx.x] = a; s_b[threadIdx.x] = b;
    __syncthreads();
    for (int o = 512; o > 0; o >>= 1) {
        if ((int)threadIdx.x < o) { s_a[threadIdx.x] += s_a[threadIdx.x + o]; s_b[threadIdx.x] += s_b[threadIdx.x + o]; }
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        const double v = fmin(clampv, fmax(-clampv, fill[0]));
        centre[0] = (s_a[0] + v * (nobs_total - s_b[0])) / nobs_total;
    }
}
__global__ void k_beta_full_centered(Grid g, int nbins, const double *__restrict__ beta, const double *__restrict__ avg,
                                     const double *__restrict__ centre, double clampv, double *__restrict__ out) {
    const int i = blockIdx.y;
    const int len = g.row_len(i);
    const int64_t dst = rm_start(i, nbins), src = g.rm_row_start(i);
    const double c = centre[0];
    for (int d = blockIdx.x * blockDim.x + threadIdx.x; i + d < nbins; d += gridDim.x * blockDim.x) {
        double v = (d < len) ? beta[src + d] : avg[0];
        v = fmin(clampv, fmax(-clampv, v));
        out[dst + d] = v - c;
    }
}

// beta (clamped at +-50) scattered to the full triangle; outside the band = avg
__global__ void k_beta_full(Grid g, int nbins, const double *__restrict__ beta, const double *__restrict__ avg,
                            double clampv, double *__restrict__ out) {
    const int i = blockIdx.y;
    const int len = g.row_len(i);
    const int64_t dst = rm_start(i, nbins), src = g.rm_row_start(i);
    for (int d = blockIdx.x * blockDim.x + threadIdx.x; i + d < nbins; d += gridDim.x * blockDim.x) {
        double v = (d < len) ? beta[src + d] : avg[0];
        v = fmin(clampv, fmax(-clampv, v));
        out[dst + d] = v;
    }
}

} // namespace

// ==========================================================================
GflSolver::GflSolver(int nbins, unsigned maxdiag, cudaStream_t stream) : nbins_(nbins), st_(stream) {
    unsigned md = maxdiag < (unsigned)nbins ? maxdiag : (unsigned)nbins; // FusedLassoGaussianEstimator.hpp:33
    if (md > (unsigned)(nbins + 1) / 2 || md <= 1) { trapeze_ = false; md = nbins; } // GFLLibrary.hpp:21-27
    else trapeze_ = true;
    if (nbins < 2) throw ArgError("fused lasso needs at least 2 bins");
    g_ = Grid::make(nbins, (int)md);
    const int64_t N = g_.ncells;
    beta_.alloc(N); betaT_.alloc(N); zr_.alloc(N); ur_.alloc(N); zc_.alloc(N); uc_.alloc(N);
    if (trapeze_) { yband_.alloc(N); wband_.alloc(N); fillavg_.alloc(1); }
    n_partials_ = 2 * g_.nchains(); // one (primal, dual) pair per warp of a team
    partials_.alloc(2 * (size_t)n_partials_);
    if (trapeze_) band_partials_.alloc(2 * (size_t)nbins);
    // k_prox: one warp per chain, the chain's prefix sums in shared memory (8 (Lmax + 1) bytes per warp)
    {
        int dev = 0, smem_max = 0;
        FB_CUDA(cudaGetDevice(&dev));
        FB_CUDA(cudaDeviceGetAttribute(&smem_max, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev));
        FB_CUDA(cudaDeviceGetAttribute(&n_sm_, cudaDevAttrMultiProcessorCount, dev));
        const int lmax = g_.row_len(0);
        prox_stride_ = lmax + 2 + 8;   // S[0..L], one spare slot, 16-int team mailbox
        prox_rlen_ = lmax + 1;
        const int64_t words = (int64_t)smem_max / 8;
        // Short chains: prefix sums and the full reciprocal table in shared memory, ten teams per SM. When fewer than
        // PROX_GS_BELOW teams would fit, the prefix sums move to global scratch (GS) and shared memory keeps the table;
        // beyond ~28 000 cells the table is cut and the quotient computed in place past its end (LONG).
        constexpr int PROX_GS_BELOW = 6;   // measured: 3000 bins (8 teams fit) is faster in shared memory, 6000 (3 fit) in scratch
        int64_t fit = (words - prox_rlen_) / prox_stride_;
        prox_gs_ = fit < PROX_GS_BELOW;
        if (const char *e = std::getenv("FB_PROX_GS")) prox_gs_ = std::atoi(e) != 0 || fit < 1;
        prox_long_ = false;
        if (prox_gs_) {
            const int64_t cap = words - 8 * PROX_GS_TEAMS - 8;
            if (prox_rlen_ > cap) { prox_rlen_ = (int)cap; prox_long_ = true; }
        }
        if (const char *e = std::getenv("FB_PROX_RTAB")) { // test aid: a table of this size (forces GS + LONG)
            const int v = std::atoi(e);
            if (v >= 2 && v < prox_rlen_) { prox_long_ = true; prox_gs_ = true; prox_rlen_ = v; }
        }
        prox_warps_ = prox_gs_ ? PROX_GS_TEAMS : std::min<int64_t>(PROX_MAX_TEAMS, fit);
        prox_warps_ = std::min(prox_warps_, std::max(1, (g_.nchains() + n_sm_ - 1) / n_sm_));
        prox_smem_ = prox_gs_ ? ((size_t)prox_rlen_ + 8 * prox_warps_ + 8) * 8 : ((size_t)prox_warps_ * prox_stride_ + prox_rlen_) * 8;
        prox_smem_ = std::max(prox_smem_, (size_t)2 * CTL_VT * sizeof(double));   // the last CTA reduces the residual sums there (fb_ctl.cuh)
        const void *fn = prox_long_ ? (const void *)k_prox<true, true> : prox_gs_ ? (const void *)k_prox<false, true> : (const void *)k_prox<false, false>;
        FB_CUDA(cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)prox_smem_));
        prox_grid_ = std::min(n_sm_, (g_.nchains() + prox_warps_ - 1) / prox_warps_);
        if (prox_gs_) gscratch_.alloc((size_t)prox_grid_ * prox_warps_ * prox_stride_);
    }
    ctl_.alloc(1);
    if (std::getenv("FB_PROX_TRACE")) { trace_.alloc(4 * (size_t)g_.nchains()); trace_.zero(); }
    h_ctl_ = static_cast<GflCtl *>(BlockCache::host().get(sizeof(GflCtl)));
    FB_CUDA(cudaEventCreate(&ev0_)); FB_CUDA(cudaEventCreate(&ev1_));
    r0_ = c0_ = 0; r1_ = c1_ = g_.n;
    if (Comm::get().active()) setup_shard();
    aset_.reset(new AsetPlan());
    {
        const char *e = std::getenv("FB_PROX_ASET");
        if (!(e && std::atoi(e) == 0) && !trace_.p) aset_plan_build(*aset_, g_, r0_, r1_, c0_, c1_, n_sm_);
        use_aset_ = aset_->usable;
        if (const char *f = std::getenv("FB_ASET_FROM")) aset_from_ = std::max(1, std::atoi(f));
    }
    // FB_CTL_FOLD=1: the last CTA of the chain-prox kernel ends the iteration itself instead of a k_ctl launch. Measured on
    // B200 at C2: 237 us per iteration against 231 with the separate launch (the tail sits on the critical path of the
    // slowest CTA, the 1024-thread k_ctl overlaps the drain of the prox kernel), so it is off by default.
    fold_ctl_ = !sharded_ && std::getenv("FB_CTL_FOLD") && std::atoi(std::getenv("FB_CTL_FOLD")) != 0;
    // the polling loop stays for the sharded solve (collectives between the kernels) and the trace aid
    use_graph_ = !sharded_ && !trace_.p && st_ != nullptr && !std::getenv("FB_NO_GRAPH");
    reset();
}

// Row blocks / column blocks of equal cell count, the chains of this rank (longest first) and, per peer, the two
// regions whose consensus terms cross between the ranks: my rows x its columns (row duals, sent; its rows x my
// columns, received) and my columns x its rows (column duals).
void GflSolver::setup_shard() {
    Comm &cm = Comm::get();
    sharded_ = true; rank_ = cm.rank(); world_ = cm.world();
    shard_timing_ = std::getenv("FB_SHARD_TIMING") != nullptr;
    row_begin_.resize(world_ + 1); col_begin_.resize(world_ + 1);
    shard_plan(g_.n, g_.md, world_, row_begin_.data(), col_begin_.data());
    r0_ = row_begin_[rank_]; r1_ = row_begin_[rank_ + 1];
    c0_ = col_begin_[rank_]; c1_ = col_begin_[rank_ + 1];
    // chains: row r (r <= n-2) is q = 2r, column j (j >= 1) is q = 2(n-1-j)+1 (Grid::chain)
    std::vector<std::pair<int, int>> mine; // (-length, q)
    for (int r = r0_; r < r1_ && r <= g_.n - 2; ++r) mine.push_back({-g_.row_len(r), 2 * r});
    for (int j = std::max(c0_, 1); j < c1_; ++j) mine.push_back({-g_.col_len(j), 2 * (g_.n - 1 - j) + 1});
    std::sort(mine.begin(), mine.end());
    std::vector<int> list(mine.size());
    for (size_t k = 0; k < mine.size(); ++k) list[k] = mine[k].second;
    n_list_ = (int)list.size();
    chain_list_.alloc(std::max<size_t>(1, list.size()));
    if (n_list_) FB_CUDA(cudaMemcpy(chain_list_.p, list.data(), list.size() * sizeof(int), cudaMemcpyHostToDevice));
    prox_grid_ = std::max(1, std::min(prox_grid_, (n_list_ + prox_warps_ - 1) / prox_warps_));   // never above the scratch slots
    // regions
    auto line_cells = [&](int dir, int line, int x0, int x1) -> int {
        int lo, hi;
        if (dir == 0) { lo = std::max(x0, line); hi = std::min(x1, line + g_.row_len(line)); }
        else { lo = std::max(x0, g_.col_first_row(line)); hi = std::min(x1, line + 1); }
        return std::max(0, hi - lo);
    };
    auto build = [&](bool sending, std::vector<ShardRegion> &regs, std::vector<int64_t> &lineoff,
                     std::vector<int64_t> &cnt, std::vector<int64_t> &off) -> int64_t {
        cnt.assign(world_, 0); off.assign(world_, 0);
        int64_t total = 0;
        for (int p = 0; p < world_; ++p) {
            if (p == rank_) continue;
            off[p] = total;
            for (int dir = 0; dir < 2; ++dir) {
                ShardRegion r;
                // lines belong to the SENDER: its rows (dir 0) or columns (dir 1); x runs over the RECEIVER's columns / rows
                const int ls = sending ? rank_ : p, xs = sending ? p : rank_;
                r.l0 = dir == 0 ? row_begin_[ls] : col_begin_[ls]; r.l1 = dir == 0 ? row_begin_[ls + 1] : col_begin_[ls + 1];
                r.x0 = dir == 0 ? col_begin_[xs] : row_begin_[xs]; r.x1 = dir == 0 ? col_begin_[xs + 1] : row_begin_[xs + 1];
                r.buf_off = total; r.off_idx = (int64_t)lineoff.size();
                int64_t acc = 0;
                for (int line = r.l0; line < r.l1; ++line) { lineoff.push_back(acc); acc += line_cells(dir, line, r.x0, r.x1); }
                total += acc;
                max_lines_ = std::max(max_lines_, r.l1 - r.l0);
                regs.push_back(r);
            }
            cnt[p] = total - off[p];
        }
        return total;
    };
    std::vector<ShardRegion> sregs, rregs;
    std::vector<int64_t> soff, roff;
    const int64_t stot = build(true, sregs, soff, send_cnt_, send_off_);
    const int64_t rtot = build(false, rregs, roff, recv_cnt_, recv_off_);
    send_regs_.alloc(sregs.size()); recv_regs_.alloc(rregs.size());
    FB_CUDA(cudaMemcpy(send_regs_.p, sregs.data(), sregs.size() * sizeof(ShardRegion), cudaMemcpyHostToDevice));
    FB_CUDA(cudaMemcpy(recv_regs_.p, rregs.data(), rregs.size() * sizeof(ShardRegion), cudaMemcpyHostToDevice));
    send_lineoff_.alloc(std::max<size_t>(1, soff.size())); recv_lineoff_.alloc(std::max<size_t>(1, roff.size()));
    if (!soff.empty()) FB_CUDA(cudaMemcpy(send_lineoff_.p, soff.data(), soff.size() * sizeof(int64_t), cudaMemcpyHostToDevice));
    if (!roff.empty()) FB_CUDA(cudaMemcpy(recv_lineoff_.p, roff.data(), roff.size() * sizeof(int64_t), cudaMemcpyHostToDevice));
    send_buf_.alloc(std::max<int64_t>(1, stot)); recv_buf_.alloc(std::max<int64_t>(1, rtot));
    send_buf_.zero(st_); recv_buf_.zero(st_);
    partials_.zero(st_);   // chains of other ranks contribute exact zeros to the all-reduce
    partials_all_.alloc(partials_.n);
    setup_p2p();
}

// Exchange of the array addresses: per rank {IPC handle of zr, of zc, raw pointers, pid, device}, broadcast from every
// rank in turn. A peer in this process is written through its raw pointer (peer access enabled), a peer in another
// process through the opened IPC handle. Any failure anywhere (no IPC in this container, no peer access) makes EVERY rank
// fall back to the packed ncclSend / ncclRecv exchange: the decision is all-reduced.
void GflSolver::setup_p2p() {
    Comm &cm = Comm::get();
    struct Info { cudaIpcMemHandle_t hr, hc, hp, hb; unsigned long long pr, pc, pp, pb; long long pid, dev; };
    static_assert(sizeof(Info) % sizeof(double) == 0, "Info is sent as doubles");
    constexpr int ND = sizeof(Info) / sizeof(double);
    const char *env = std::getenv("FB_SHARD_P2P");
    bool ok = !(env && std::atoi(env) == 0);
    std::vector<Info> info(world_);
    std::memset(info.data(), 0, info.size() * sizeof(Info));
    Info &me = info[rank_];
    int dev = 0; cudaGetDevice(&dev);
    me.pr = (unsigned long long)zr_.p; me.pc = (unsigned long long)zc_.p; me.pp = (unsigned long long)partials_all_.p;
    me.pb = (unsigned long long)beta_.p;
    me.pid = (long long)getpid(); me.dev = dev;
    if (ok && (cudaIpcGetMemHandle(&me.hr, zr_.p) != cudaSuccess || cudaIpcGetMemHandle(&me.hc, zc_.p) != cudaSuccess ||
               cudaIpcGetMemHandle(&me.hp, partials_all_.p) != cudaSuccess || cudaIpcGetMemHandle(&me.hb, beta_.p) != cudaSuccess)) {
        cudaGetLastError(); ok = false;
    }
    DevBuf<double> buf((size_t)ND * world_);
    FB_CUDA(cudaMemcpyAsync(buf.p + (size_t)ND * rank_, &me, sizeof(Info), cudaMemcpyHostToDevice, st_));
    cm.group_start();
    for (int p = 0; p < world_; ++p) cm.bcast(buf.p + (size_t)ND * p, ND, p, st_);
    cm.group_end();
    FB_CUDA(cudaMemcpyAsync(info.data(), buf.p, info.size() * sizeof(Info), cudaMemcpyDeviceToHost, st_));
    FB_CUDA(cudaStreamSynchronize(st_));
    std::vector<double *> pr, pc, pq, pbeta;
    for (int p = 0; p < world_ && ok; ++p) {
        if (p == rank_) continue;
        double *r = nullptr, *c = nullptr, *q = nullptr, *bt = nullptr;
        if (info[p].pid == me.pid) {     // same process (fb_create_multi): one address space
            int can = 0;
            if (cudaDeviceCanAccessPeer(&can, dev, (int)info[p].dev) != cudaSuccess || !can) { cudaGetLastError(); ok = false; break; }
            const cudaError_t e = cudaDeviceEnablePeerAccess((int)info[p].dev, 0);
            if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) { cudaGetLastError(); ok = false; break; }
            cudaGetLastError();
            r = (double *)info[p].pr; c = (double *)info[p].pc; q = (double *)info[p].pp; bt = (double *)info[p].pb;
        } else {
            void *a = cm.ipc_open(info[p].hr), *b = a ? cm.ipc_open(info[p].hc) : nullptr;   // kept open by the communicator
            void *d = (a && b) ? cm.ipc_open(info[p].hp) : nullptr, *e = d ? cm.ipc_open(info[p].hb) : nullptr;
            if (!a || !b || !d || !e) { ok = false; break; }
            p2p_ipc_ = true;
            r = (double *)a; c = (double *)b; q = (double *)d; bt = (double *)e;
        }
        pr.push_back(r); pc.push_back(c); pq.push_back(q); pbeta.push_back(bt);
    }
    // agree: the sum of the failure flags over the ranks
    bar_.alloc(4);
    const double fail[2] = {ok ? 0.0 : 1.0, 0.0};
    FB_CUDA(cudaMemcpyAsync(bar_.p, fail, sizeof fail, cudaMemcpyHostToDevice, st_));
    cm.allreduce_sum(bar_.p, bar_.p + 2, 2, st_);
    double tot[2] = {0, 0};
    FB_CUDA(cudaMemcpyAsync(tot, bar_.p + 2, sizeof tot, cudaMemcpyDeviceToHost, st_));
    FB_CUDA(cudaStreamSynchronize(st_));
    if ((std::getenv("FB_SHARD_VERBOSE") || shard_timing_) && rank_ == 0)
        std::fprintf(stderr, "[fb shard] exchange of the crossing terms: %s\n",
                     tot[0] != 0.0 ? "packed ncclSend / ncclRecv" : (p2p_ipc_ ? "peer stores (CUDA IPC)" : "peer stores (same process)"));
    if (tot[0] != 0.0) { teardown_p2p(); p2p_ = false; return; }
    peer_zr_.alloc(std::max<size_t>(1, pr.size())); peer_zc_.alloc(std::max<size_t>(1, pc.size()));
    peer_part_.alloc(std::max<size_t>(1, pq.size())); peer_beta_.alloc(std::max<size_t>(1, pbeta.size()));
    if (!pr.empty()) {
        FB_CUDA(cudaMemcpy(peer_zr_.p, pr.data(), pr.size() * sizeof(double *), cudaMemcpyHostToDevice));
        FB_CUDA(cudaMemcpy(peer_zc_.p, pc.data(), pc.size() * sizeof(double *), cudaMemcpyHostToDevice));
        FB_CUDA(cudaMemcpy(peer_part_.p, pq.data(), pq.size() * sizeof(double *), cudaMemcpyHostToDevice));
        FB_CUDA(cudaMemcpy(peer_beta_.p, pbeta.data(), pbeta.size() * sizeof(double *), cudaMemcpyHostToDevice));
    }
    p2p_ = true;
}

void GflSolver::teardown_p2p() {}   // the IPC mappings belong to the communicator (Comm::ipc_open)

void GflSolver::tmark(int k) {
    if (!shard_timing_) return;
    if (tev_.empty()) { tev_.resize(8); for (auto &e : tev_) cudaEventCreate(&e); }
    cudaEventRecord(tev_[k], st_);
}
void GflSolver::tflush() { // after a stream synchronisation: add the phases of the LAST iteration (a sample per poll batch)
    if (!shard_timing_ || tev_.empty()) return;
    for (int k = 0; k < 7; ++k) { float ms = 0.f; if (cudaEventElapsedTime(&ms, tev_[k], tev_[k + 1]) == cudaSuccess) tacc_[k] += ms; }
    ++tcount_;
}

// z - u of the crossing cells, both directions, to and from every peer
void GflSolver::exchange(bool zero_u) {
    Comm &cm = Comm::get();
    dim3 grid((unsigned)std::min(max_lines_, 4096), (unsigned)(world_ - 1));
    if (p2p_) {
        if (zero_u) {
            // start of a solve: no peer may store into this rank's arrays before its own resets / state imports (queued
            // on this stream) are done, so everybody waits for everybody first
            cm.allreduce_sum(bar_.p, bar_.p + 2, 1, st_);
            k_zero_remote_u<0><<<grid, 256, 0, st_>>>(g_, recv_regs_.p, ur_.p);
            k_zero_remote_u<1><<<grid, 256, 0, st_>>>(g_, recv_regs_.p, uc_.p);
            kernel_launches += 2;
        }
        tmark(5);
        k_push<0><<<grid, 256, 0, st_>>>(g_, ctl_.p, send_regs_.p, zr_.p, ur_.p, peer_zr_.p);
        k_push<1><<<grid, 256, 0, st_>>>(g_, ctl_.p, send_regs_.p, zc_.p, uc_.p, peer_zc_.p);
        kernel_launches += 2;
        tmark(6);
        // every rank's stores precede its contribution: when the sum arrives, all the terms this rank reads next have landed
        cm.allreduce_sum(bar_.p, bar_.p + 2, 1, st_);
        tmark(7);
        return;
    }
    k_exchange<0, true><<<grid, 256, 0, st_>>>(g_, ctl_.p, send_regs_.p, send_lineoff_.p, zr_.p, ur_.p, send_buf_.p, 0);
    k_exchange<1, true><<<grid, 256, 0, st_>>>(g_, ctl_.p, send_regs_.p, send_lineoff_.p, zc_.p, uc_.p, send_buf_.p, 0);
    tmark(5);
    cm.group_start();
    for (int p = 0; p < world_; ++p) {
        if (p == rank_) continue;
        if (send_cnt_[p]) cm.send(send_buf_.p + send_off_[p], (size_t)send_cnt_[p], p, st_);
        if (recv_cnt_[p]) cm.recv(recv_buf_.p + recv_off_[p], (size_t)recv_cnt_[p], p, st_);
    }
    cm.group_end();
    tmark(6);
    k_exchange<0, false><<<grid, 256, 0, st_>>>(g_, ctl_.p, recv_regs_.p, recv_lineoff_.p, zr_.p, ur_.p, recv_buf_.p, zero_u ? 1 : 0);
    k_exchange<1, false><<<grid, 256, 0, st_>>>(g_, ctl_.p, recv_regs_.p, recv_lineoff_.p, zc_.p, uc_.p, recv_buf_.p, zero_u ? 1 : 0);
    kernel_launches += 4;
    tmark(7);
}

// every rank ends with the whole array: the row blocks (row-major array) or column blocks (column-major array)
void GflSolver::allgather_rows(double *a) {
    Comm &cm = Comm::get();
    cm.group_start();
    for (int p = 0; p < world_; ++p) {
        const int64_t b = g_.rm_row_start(row_begin_[p]), e = g_.rm_row_start(row_begin_[p + 1]);
        if (e > b) cm.bcast(a + b, (size_t)(e - b), p, st_);
    }
    cm.group_end();
}
void GflSolver::allgather_cols(double *a) {
    Comm &cm = Comm::get();
    cm.group_start();
    for (int p = 0; p < world_; ++p) {
        const int64_t b = g_.cm_col_start(col_begin_[p]), e = g_.cm_col_start(col_begin_[p + 1]);
        if (e > b) cm.bcast(a + b, (size_t)(e - b), p, st_);
    }
    cm.group_end();
}

GflSolver::~GflSolver() {
    cudaStreamSynchronize(st_); // buffers go back to the block cache only when nothing uses them any more
    // (peer stores: every exchange ends with a barrier all-reduce that this rank's stream has passed, so no peer is still
    // storing into these arrays; a collective in a destructor would hang the survivors of a failed rank)
    if (shard_timing_ && tcount_) {
        static const char *nm_nccl[7] = {"k_beta", "k_prox", "allreduce", "k_ctl", "pack", "send/recv", "unpack"};
        static const char *nm_p2p[7] = {"k_beta", "k_prox", "allreduce", "k_ctl", "-", "push (peer stores)", "barrier"};
        const char **nm = p2p_ ? nm_p2p : nm_nccl;
        std::string line = "[fb shard timing] rank " + std::to_string(rank_) + " n=" + std::to_string(g_.n) + ", mean over " +
                           std::to_string(tcount_) + " sampled iterations (ms):";
        char buf[64];
        for (int k = 0; k < 7; ++k) { std::snprintf(buf, sizeof buf, " %s %.3f", nm[k], tacc_[k] / tcount_); line += buf; }
        line += "\n";
        std::fputs(line.c_str(), stderr);     // one write: the ranks share the terminal
    }
    for (auto &e : tev_) cudaEventDestroy(e);
    if (graph_exec_) cudaGraphExecDestroy(graph_exec_);
    if (graph_) cudaGraphDestroy(graph_);
    if (h_ctl_) BlockCache::host().put(h_ctl_, sizeof(GflCtl));
    cudaEventDestroy(ev0_); cudaEventDestroy(ev1_);
}

void GflSolver::reset() {
    counter = 0;
    alpha = 5.0;
    cold_ = true;
    beta_.zero(st_); zr_.zero(st_); ur_.zero(st_); zc_.zero(st_); uc_.zero(st_);
    betaT_.zero(st_);
}

void GflSolver::launch_iteration(const double *y, const double *w, bool in_graph) {
    const int nt = (g_.n + TILE - 1) / TILE;
    dim3 tb(TILE, 8), tg((unsigned)std::min(nt, (g_.md + TILE - 2) / TILE + 2), (unsigned)nt);   // tile columns that can touch the band
    tmark(0);
    k_beta<<<tg, tb, 0, st_>>>(g_, ctl_.p, y, w, zr_.p, ur_.p, zc_.p, uc_.p, beta_.p, betaT_.p, r0_, r1_, c0_, c1_);
    tmark(1);
    const int *list = sharded_ ? chain_list_.p : nullptr;
    // both chain-prox kernels are launched; the one ctl->prox_sel does not name returns at once (k_ctl picks per step:
    // the taut-string search is faster while the string has few knots, the parallel active set afterwards)
    // not sharded: the last CTA of the chain-prox kernel ends the iteration itself (no k_ctl launch); sharded: the
    // partial sums are all-reduced first, k_ctl follows
    GflFold fold{};
    fold.on = fold_ctl_ ? 1 : 0; fold.graph = in_graph ? 1 : 0; fold.npart = n_partials_;
    fold.ncells = g_.ncells; fold.nz = 2 * g_.ncells - 2; fold.cond = cond_;
    if (use_aset_) aset_prox_launch(*aset_, ctl_.p, beta_.p, betaT_.p, zr_.p, ur_.p, zc_.p, uc_.p, partials_.p, fold, st_);
#define FB_PROX_LAUNCH(LONG_, GS_)                                                                                              \
    k_prox<LONG_, GS_><<<prox_grid_, 64 * prox_warps_, prox_smem_, st_>>>(g_, ctl_.p, beta_.p, betaT_.p, zr_.p, ur_.p, zc_.p, uc_.p,  \
                                                                         partials_.p, prox_stride_, prox_rlen_, trace_.p, list, n_list_, gscratch_.p, fold)
    if (prox_long_) FB_PROX_LAUNCH(true, true);
    else if (prox_gs_) FB_PROX_LAUNCH(false, true);
    else FB_PROX_LAUNCH(false, false);
#undef FB_PROX_LAUNCH
    // sharded: the residual sums of every chain on every rank (exactly one rank contributes a non-zero term per
    // slot, so the sum is exact and k_ctl reduces the same numbers in the same order everywhere)
    tmark(2);
    if (sharded_ && p2p_) {
        // every rank stores the sums of its own chains into everybody's array (peer stores), then a scalar all-reduce as
        // the barrier: the same numbers in the same slots as the all-reduce of the zero-padded arrays, nothing added
        if (n_list_) k_push_partials<<<(n_list_ + 255) / 256, 256, 0, st_>>>(ctl_.p, chain_list_.p, n_list_, partials_.p, partials_all_.p,
                                                                          peer_part_.p, world_ - 1);
        Comm::get().allreduce_sum(bar_.p, bar_.p + 2, 1, st_);
    } else if (sharded_) Comm::get().allreduce_sum(partials_.p, partials_all_.p, 2 * (size_t)n_partials_, st_);
    tmark(3);
    const double *parts = sharded_ ? partials_all_.p : partials_.p;
    if (fold.on) { tmark(4); return; }
    if (in_graph) k_ctl<true><<<1, CTL_THREADS, 0, st_>>>(ctl_.p, parts, n_partials_, g_.ncells, 2 * g_.ncells - 2, cond_);
    else k_ctl<false><<<1, CTL_THREADS, 0, st_>>>(ctl_.p, parts, n_partials_, g_.ncells, 2 * g_.ncells - 2, cond_);
    tmark(4);
    if (sharded_) exchange(false);
}

void GflSolver::iteration(const double *y, const double *w, int parity) {
    (void)parity;
    launch_iteration(y, w, false);
    kernel_launches += (use_aset_ ? 4 : 3) - (fold_ctl_ ? 1 : 0);
}

// One iteration captured into the body of a WHILE conditional node (CUDA >= 12.4). The kernels read y / w from the
// control block, so the same executable graph serves every solve of this object.
void GflSolver::build_graph() {
    FB_CUDA(cudaGraphCreate(&graph_, 0));
    FB_CUDA(cudaGraphConditionalHandleCreate(&cond_, graph_, 1, cudaGraphCondAssignDefault));
    cudaGraphNodeParams np = {};
    np.type = cudaGraphNodeTypeConditional;
    np.conditional.handle = cond_;
    np.conditional.type = cudaGraphCondTypeWhile;
    np.conditional.size = 1;
    cudaGraphNode_t node;
    FB_CUDA(cudaGraphAddNode(&node, graph_, nullptr, 0, &np));
    cudaGraph_t body = np.conditional.phGraph_out[0];
    FB_CUDA(cudaStreamBeginCaptureToGraph(st_, body, nullptr, nullptr, 0, cudaStreamCaptureModeThreadLocal));
    launch_iteration(nullptr, nullptr, true);
    FB_CUDA(cudaStreamEndCapture(st_, nullptr));
    FB_CUDA(cudaGraphInstantiate(&graph_exec_, graph_, 0));
}

int GflSolver::optimize(const double *d_y_tri, const double *d_w_tri, double lam2, double converge) {
    FB_CUDA(cudaEventRecord(ev0_, st_));
    const double *y = d_y_tri, *w = d_w_tri;
    if (trapeze_) {
        dim3 gg(4, nbins_);
        k_band_gather<<<gg, 256, 0, st_>>>(g_, nbins_, d_y_tri, yband_.p);
        k_band_gather<<<gg, 256, 0, st_>>>(g_, nbins_, d_w_tri, wband_.p);
        y = yband_.p; w = wband_.p;
        kernel_launches += 2;
    }
    y_tri_ = d_y_tri; w_tri_ = d_w_tri;
    GflCtl c;
    c.alpha = alpha; c.lam = lam2; c.converge = converge; c.inflate = inflate;
    c.presnorm = c.dresnorm = 0; c.step = 1; c.maxsteps = ninner_max;
    c.done = !(c.step < c.maxsteps && (converge + 1) > converge);
    c.rescale = 0; c.next_chain = 0; c.idle = 0; c.ctas_done = 0;
    c.aset_from = use_aset_ ? (cold_ ? aset_from_ : 0) : -1;
    c.prox_sel = (c.aset_from >= 0 && c.step >= c.aset_from) ? 1 : 0;
    c.y = y; c.w = w;
    *h_ctl_ = c;
    FB_CUDA(cudaMemcpyAsync(ctl_.p, h_ctl_, sizeof(GflCtl), cudaMemcpyHostToDevice, st_));
    if (sharded_) exchange(true);   // whatever state the ranks start from: remote slots hold z - u of their owner, u = 0
    int parity = 0;
    bool done = c.done != 0;
    if (!done && use_graph_) {
        if (!graph_exec_) build_graph();
        FB_CUDA(cudaGraphLaunch(graph_exec_, st_));
        k_export_ctl<<<1, 32, 0, st_>>>(ctl_.p, h_ctl_);
        FB_CUDA(cudaStreamSynchronize(st_));
        if (!h_ctl_->done) throw CudaError("the captured ADMM loop ended without setting done");
        kernel_launches += ((use_aset_ ? 4 : 3) - (fold_ctl_ ? 1 : 0)) * (int64_t)(h_ctl_->step - 1);
        done = true;
    }
    while (!done) {
        // sharded solves of large grids: an iteration (with its exchange) costs milliseconds, the poll microseconds,
        // and iterations queued past convergence would still run their collectives
        int batch = (sharded_ && g_.ncells > 2000000) ? 1 : POLL_BATCH;
        if (const char *e = std::getenv("FB_SHARD_POLL")) batch = std::max(1, std::atoi(e));   // development aid
        for (int b = 0; b < batch; ++b) { iteration(y, w, parity); parity ^= 1; }
        FB_LAUNCH_CHECK();
        k_export_ctl<<<1, 32, 0, st_>>>(ctl_.p, h_ctl_);
        FB_CUDA(cudaStreamSynchronize(st_));
        if (!h_ctl_->done) tflush();
        done = h_ctl_->done != 0;
    }
    if (!c.done) {
        if (h_ctl_->rescale) { // rescale decided on the very last step still applies to u
            const int64_t N = g_.ncells;
            k_rescale_u<<<(unsigned)((N + 255) / 256), 256, 0, st_>>>(ur_.p, N, h_ctl_->rescale, inflate);
            k_rescale_u<<<(unsigned)((N + 255) / 256), 256, 0, st_>>>(uc_.p, N, h_ctl_->rescale, inflate);
            kernel_launches += 2;
        }
        alpha = h_ctl_->alpha;
        last_steps = h_ctl_->step;
    } else last_steps = 1;
    counter += last_steps;
    cold_ = false;
    if (sharded_ && p2p_) {                  // every rank continues with the whole solution: row blocks by peer stores + barrier
        const int64_t b = g_.rm_row_start(row_begin_[rank_]), e = g_.rm_row_start(row_begin_[rank_ + 1]);
        if (e > b) k_push_rows<<<(unsigned)std::min<int64_t>(4 * 148, (e - b + 255) / 256), 256, 0, st_>>>(beta_.p, b, e, peer_beta_.p, world_ - 1);
        Comm::get().allreduce_sum(bar_.p, bar_.p + 2, 1, st_);
        kernel_launches += 1;
    } else if (sharded_) allgather_rows(beta_.p);
    if (trapeze_) { // average of the far region (fill_triangle)
        // the rows of the far region are dealt to the ranks; the per-row sums of the others arrive as x + 0 (exact)
        k_band_outside_sums<<<nbins_, 256, 0, st_>>>(g_, nbins_, d_y_tri, d_w_tri, band_partials_.p, sharded_ ? rank_ : 0, sharded_ ? world_ : 1);
        if (sharded_) Comm::get().allreduce_sum(band_partials_.p, band_partials_.p, 2 * (size_t)nbins_, st_);
        k_band_avg<<<1, 1024, 0, st_>>>(nbins_, band_partials_.p, fillavg_.p);
        kernel_launches += 2;
    }
    FB_CUDA(cudaEventRecord(ev1_, st_));
    FB_CUDA(cudaEventSynchronize(ev1_));
    FB_CUDA(cudaEventElapsedTime(&ms_last, ev0_, ev1_));
    if (trace_.p) { // development aid (FB_PROX_TRACE=file): per-chain cycle counts of the last executed k_prox
        std::vector<long long> h(trace_.n);
        FB_CUDA(cudaMemcpy(h.data(), trace_.p, trace_.n * sizeof(long long), cudaMemcpyDeviceToHost));
        if (FILE *f = std::fopen(std::getenv("FB_PROX_TRACE"), "wb")) { std::fwrite(h.data(), sizeof(long long), h.size(), f); std::fclose(f); }
    }
    return last_steps;
}

void GflSolver::get_beta_full(double *d_out_tri) {
    dim3 gg(4, nbins_);
    k_beta_full<<<gg, 256, 0, st_>>>(g_, nbins_, beta_.p, trapeze_ ? fillavg_.p : beta_.p, 50.0, d_out_tri);
    kernel_launches += 1;
    FB_LAUNCH_CHECK();
}

// get() and the centring of step_signal in one pass (band solves only; nobs_total = sum of nobs over the dataset, exact)
void GflSolver::get_beta_full_centered(double *d_out_tri, const int32_t *d_nobs_tri, double nobs_total) {
    if (!trapeze_) throw StateError("get_beta_full_centered is for band solves");
    if (!center_partials_.p) { center_partials_.alloc(2 * (size_t)4 * nbins_); center_val_.alloc(1); }
    dim3 gg(4, nbins_);
    k_band_center_partial<<<gg, 256, 0, st_>>>(g_, nbins_, beta_.p, d_nobs_tri, 50.0, center_partials_.p);
    k_band_center_final<<<1, 1024, 0, st_>>>((int64_t)4 * nbins_, center_partials_.p, fillavg_.p, 50.0, nobs_total, center_val_.p);
    k_beta_full_centered<<<gg, 256, 0, st_>>>(g_, nbins_, beta_.p, fillavg_.p, center_val_.p, 50.0, d_out_tri);
    kernel_launches += 3;
    FB_LAUNCH_CHECK();
}

// reference slot order: row chains (rm index p < N-1), then column chains
// (cm index >= 1 maps to slot N-1 + cm-1)
void GflSolver::export_state(double *h_beta, double *h_z, double *h_u) {
    const int64_t N = g_.ncells;
    if (sharded_ && (h_z || h_u)) { // collect the duals of every rank's chains
        allgather_rows(zr_.p); allgather_rows(ur_.p); allgather_cols(zc_.p); allgather_cols(uc_.p);
    }
    FB_CUDA(cudaStreamSynchronize(st_));
    if (h_beta) FB_CUDA(cudaMemcpy(h_beta, beta_.p, N * sizeof(double), cudaMemcpyDeviceToHost));
    if (h_z) {
        FB_CUDA(cudaMemcpy(h_z, zr_.p, (N - 1) * sizeof(double), cudaMemcpyDeviceToHost));
        FB_CUDA(cudaMemcpy(h_z + (N - 1), zc_.p + 1, (N - 1) * sizeof(double), cudaMemcpyDeviceToHost));
    }
    if (h_u) {
        FB_CUDA(cudaMemcpy(h_u, ur_.p, (N - 1) * sizeof(double), cudaMemcpyDeviceToHost));
        FB_CUDA(cudaMemcpy(h_u + (N - 1), uc_.p + 1, (N - 1) * sizeof(double), cudaMemcpyDeviceToHost));
    }
}
void GflSolver::import_state(const double *h_beta, const double *h_z, const double *h_u) {
    const int64_t N = g_.ncells;
    FB_CUDA(cudaStreamSynchronize(st_));
    if (h_beta) FB_CUDA(cudaMemcpy(beta_.p, h_beta, N * sizeof(double), cudaMemcpyHostToDevice));
    if (h_z) {
        cold_ = false;
        FB_CUDA(cudaMemcpy(zr_.p, h_z, (N - 1) * sizeof(double), cudaMemcpyHostToDevice));
        FB_CUDA(cudaMemcpy(zc_.p + 1, h_z + (N - 1), (N - 1) * sizeof(double), cudaMemcpyHostToDevice));
    }
    if (h_u) {
        FB_CUDA(cudaMemcpy(ur_.p, h_u, (N - 1) * sizeof(double), cudaMemcpyHostToDevice));
        FB_CUDA(cudaMemcpy(uc_.p + 1, h_u + (N - 1), (N - 1) * sizeof(double), cudaMemcpyHostToDevice));
    }
}

} // namespace fb
