// fb_prox_lane.cu -- "split and verify" chain prox: the z / u update of one ADMM iteration
// (update_z with tf_dp_weight, update_u, primal_resnorm, dual_resnorm: gfl_graph_fl.c:267-365, gfl_tf.c:184-321).
//
// The 1-D fused-lasso minimiser of a chain is the derivative of the taut string through the tube S[k] -+ lam around the
// prefix sums S of v = beta + u, from (0, 0) to (L, S[L]). A point where the string is KNOWN to touch the tube (position
// and side) cuts the problem into independent ones with fixed end points, and the previous ADMM iterate tells where
// those points probably are: its jumps. So, per batch of chains held in shared memory:
//   A  stream beta, u, z of the chains in: v and its per-chain prefix sums S; the jumps of the old z become knot
//      hypotheses, the largest jump of every window of LP_W cells a SPLIT
//   B  every range between two consecutive splits is an independent exact problem: ONE LANE runs the sequential taut
//      string on it (lanes fetch ranges from a shared queue as they finish); ranges longer than LP_LONG cells are solved
//      by a whole warp with the candidate-parallel search of fb_taut.cuh
//   C  a split is correct iff the string bends towards the tube edge it touches there; then the glued string satisfies
//      the KKT conditions of a strictly convex problem, i.e. it IS the minimiser. Failed splits are removed, the merged
//      ranges solved again, until every remaining split passes (each pass removes at least one split)
//   D  stream out: z = c + slope of the piece a cell lies on, u += beta - z, the residual sums per chain
// The result does not depend on the hypothesis, only the amount of sequential work does (oracle/lane_prox.c is the scalar
// model, tests/test_lane_prox_model.py pins it to the reference's tf_dp_weight). Every chain is computed from its own
// inputs only, in a fixed order: results do not depend on how chains are batched or which rank owns them.
#include "fb_prox_lane.cuh"
#include "fb_taut.cuh"
#include <algorithm>
#include <cstdlib>

namespace fb {

namespace {
using namespace taut;

#ifndef FB_LP_W
#define FB_LP_W 8
#endif
constexpr int LP_NT = 64, LP_NW = LP_NT / 32;   // two partial-sum slots per chain, like the team kernel
constexpr int LP_W = FB_LP_W;                   // cells per split window (power of two <= 32)
constexpr int LP_LONG = 64;                     // ranges longer than this are solved by a whole warp
constexpr int LP_RT = 256;                      // reciprocal table R[m] = 1/m
constexpr int LP_COOP = 1024;                   // chains longer than this are streamed by all warps of the CTA
enum { C_QN = 0, C_QHEAD, C_NLONG, C_LHEAD, C_BATCH, C_COUNT = 8 };

struct Sm {
    double *S, *R, *Cc;
    long long *G;
    int *Lc, *Bc, *Qc;
    unsigned *K, *UP, *SPLIT, *FIXM, *DIRTY, *FAILED;
    unsigned short *q, *lq;
    int *ci;
};

__host__ __device__ inline size_t lp_align8(size_t x) { return (x + 7) & ~size_t(7); }
__host__ __device__ inline int lp_qcap(int cap) { return cap / LP_W + 2 * (cap / 32) + 8; }
__host__ __device__ inline size_t lp_smem_bytes(int cap) {
    const int maxch = cap / 32, nw = cap / 32;
    size_t b = (size_t)cap * 8 + LP_RT * 8 + (size_t)maxch * 8 + (size_t)maxch * 8 + (size_t)maxch * 12;
    b = lp_align8(b) + (size_t)nw * 4 * 6;
    b = lp_align8(b) + (size_t)lp_qcap(cap) * 2 * 2;
    b = lp_align8(b) + C_COUNT * 4;
    return lp_align8(b);
}
__device__ __forceinline__ Sm carve(unsigned char *p, int cap) {
    const int maxch = cap / 32, nw = cap / 32;
    Sm s;
    s.S = reinterpret_cast<double *>(p); p += (size_t)cap * 8;
    s.R = reinterpret_cast<double *>(p); p += LP_RT * 8;
    s.Cc = reinterpret_cast<double *>(p); p += (size_t)maxch * 8;
    s.G = reinterpret_cast<long long *>(p); p += (size_t)maxch * 8;
    s.Lc = reinterpret_cast<int *>(p); p += (size_t)maxch * 4;
    s.Bc = reinterpret_cast<int *>(p); p += (size_t)maxch * 4;
    s.Qc = reinterpret_cast<int *>(p); p += (size_t)maxch * 4;
    p = reinterpret_cast<unsigned char *>(lp_align8(reinterpret_cast<size_t>(p)));
    s.K = reinterpret_cast<unsigned *>(p); p += (size_t)nw * 4;
    s.UP = reinterpret_cast<unsigned *>(p); p += (size_t)nw * 4;
    s.SPLIT = reinterpret_cast<unsigned *>(p); p += (size_t)nw * 4;
    s.FIXM = reinterpret_cast<unsigned *>(p); p += (size_t)nw * 4;
    s.DIRTY = reinterpret_cast<unsigned *>(p); p += (size_t)nw * 4;
    s.FAILED = reinterpret_cast<unsigned *>(p); p += (size_t)nw * 4;
    p = reinterpret_cast<unsigned char *>(lp_align8(reinterpret_cast<size_t>(p)));
    s.q = reinterpret_cast<unsigned short *>(p); p += (size_t)lp_qcap(cap) * 2;
    s.lq = reinterpret_cast<unsigned short *>(p); p += (size_t)lp_qcap(cap) * 2;
    p = reinterpret_cast<unsigned char *>(lp_align8(reinterpret_cast<size_t>(p)));
    s.ci = reinterpret_cast<int *>(p);
    return s;
}

// first set bit at a position > p / last set bit at a position <= p (both exist: every chain starts and ends with one)
__device__ __forceinline__ int next_bit(const unsigned *__restrict__ M, int p) {
    int wi = p >> 5;
    unsigned m = M[wi] & (0xfffffffeu << (p & 31));
    while (!m) m = M[++wi];
    return (wi << 5) + __ffs(m) - 1;
}
__device__ __forceinline__ int prev_bit_le(const unsigned *__restrict__ M, int p) {
    int wi = p >> 5;
    unsigned m = M[wi] & (0xffffffffu >> (31 - (p & 31)));
    while (!m) m = M[--wi];
    return (wi << 5) + 31 - __clz(m);
}
__device__ __forceinline__ bool bit_at(const unsigned *__restrict__ M, int p) { return (M[p >> 5] >> (p & 31)) & 1u; }
// offset of the string from S at a knot: 0 at a chain end, +lam on the upper tube edge, -lam on the lower one
__device__ __forceinline__ double side_off(const Sm &s, int p, double lam) {
    return bit_at(s.FIXM, p) ? 0.0 : (bit_at(s.UP, p) ? lam : -lam);
}
__device__ __forceinline__ double recip_t(const double *__restrict__ R, int m) { return m < LP_RT ? R[m] : 1.0 / (double)m; }

// knots strictly inside (a, b) of an earlier pass go before the range is solved again
__device__ __forceinline__ void clear_inside(const Sm &s, int a, int b) {
    for (int wi = a >> 5; wi <= (b >> 5); ++wi) {
        unsigned inside = 0xffffffffu;
        if (wi == (a >> 5)) inside &= 0xfffffffeu << (a & 31);
        if (wi == (b >> 5)) inside &= ~(0xffffffffu << (b & 31));
        if (s.K[wi] & inside) atomicAnd(&s.K[wi], ~inside);
        if (s.UP[wi] & inside) atomicAnd(&s.UP[wi], ~inside);
    }
}
__device__ __forceinline__ void set_knot(const Sm &s, int p, bool up) {
    atomicOr(&s.K[p >> 5], 1u << (p & 31));
    if (up) atomicOr(&s.UP[p >> 5], 1u << (p & 31));
}

// B: the lanes of a warp solve ranges from the queue, one range per lane at a time, one candidate per step
__device__ __forceinline__ void lane_pass(const Sm &s, const double lam, const int lane) {
    const double INF = __longlong_as_double(0x7ff0000000000000LL);
    const int qn = s.ci[C_QN];
    bool act = false, drained = false;
    int b = 0, k = 0, k0 = 0, kH = 0, kL = 0;
    double y0 = 0.0, H = INF, Lw = -INF, aEnd = 0.0;
    for (;;) {
        const unsigned idle = __ballot_sync(FULL, !act);
        if (idle && !drained) {
            const int want = __popc(idle);
            int first = 0;
            if (lane == 0) first = atomicAdd(&s.ci[C_QHEAD], want);
            first = __shfl_sync(FULL, first, 0);
            if (first + want >= qn) drained = true;
            if (!act) {
                const int idx = first + __popc(idle & ((1u << lane) - 1u));
                if (idx < qn) {
                    const int a = s.q[idx];
                    b = next_bit(s.SPLIT, a);
                    clear_inside(s, a, b);
                    if (b - a > LP_LONG) s.lq[atomicAdd(&s.ci[C_NLONG], 1)] = (unsigned short)a;
                    else {
                        k0 = a; y0 = s.S[a] + side_off(s, a, lam); aEnd = side_off(s, b, lam);
                        k = a + 1; H = INF; Lw = -INF; kH = kL = a; act = true;
                    }
                }
            }
        }
        if (!__any_sync(FULL, act)) { if (drained) break; continue; }
        if (act) {
            const double q = s.S[k] - y0;
            const double r = s.R[k - k0];
            const bool last = (k == b);
            const double h = (q + (last ? aEnd : lam)) * r;
            const double l = (q + (last ? aEnd : -lam)) * r;
            const bool vu = l > H, vd = h < Lw;
            if (vu || vd) {       // the segment from k0 ends where the violated bound was attained
                const int kk = vu ? kH : kL;
                set_knot(s, kk, vu);
                k0 = kk; y0 = s.S[kk] + (vu ? lam : -lam);
                H = INF; Lw = -INF; kH = kL = kk; k = kk + 1;
            } else {
                if (h <= H) { H = h; kH = k; }
                if (l >= Lw) { Lw = l; kL = k; }
                if (last) act = false; else ++k;
            }
        }
    }
}

// B (long ranges): one warp, 32 * PBR candidates per round
__device__ __forceinline__ void coop_range(const Sm &s, const double lam, const int a, const int b, const int lane) {
    Frame f; f.S = s.S; f.L = 0; f.dir = 0; f.SL = 0.0;
    Seg sg; sg.k0 = a; sg.s0 = s.S[a] + side_off(s, a, lam); seg_reset(sg);
    const double aEnd = side_off(s, b, lam);
    for (;;) {
        int isup = 0, kend = 0;
        double slope = 0.0;
        const bool viol = taut_round<true>(f, s.R, LP_RT, lam, sg, b, aEnd, lane, isup, slope, kend);
        sg.p += 32 * PBR;
        if (viol) {
            if (lane == 0) set_knot(s, kend, isup != 0);
            sg.k0 = kend; sg.s0 = s.S[kend] + (isup ? lam : -lam); seg_reset(sg);
        } else if (sg.p >= b) break;
    }
}

// C: does the string bend towards the tube edge it touches at the split p?
__device__ __forceinline__ bool bend_ok(const Sm &s, const int p, const double lam) {
    const int a = prev_bit_le(s.K, p - 1), b = next_bit(s.K, p);
    const bool up = bit_at(s.UP, p);
    const double yk = s.S[p] + (up ? lam : -lam);
    const double ya = s.S[a] + side_off(s, a, lam), yb = s.S[b] + side_off(s, b, lam);
    const double bend = (yb - yk) / (double)(b - p) - (yk - ya) / (double)(p - a);
    return up ? bend > 0.0 : bend < 0.0;
}

__global__ void __launch_bounds__(LP_NT)
k_prox_lane(GflCtl *__restrict__ ctl, const LpChain *__restrict__ chains, const LpBatch *__restrict__ batches,
            const int nbatches, const int cap, const double *__restrict__ beta, const double *__restrict__ betaT,
            double *__restrict__ zr, double *__restrict__ ur, double *__restrict__ zc, double *__restrict__ uc,
            double *__restrict__ partials) {
    if (ctl->done) return;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const Sm s = carve(smem_raw, cap);
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    for (int m = tid; m < LP_RT; m += LP_NT) s.R[m] = m ? 1.0 / (double)m : 0.0;
    const double alpha = ctl->alpha;
    const double lam = ctl->lam / (alpha / 2.0);
    const bool copy = (lam == 0.0);          // tf_dp_weight copies y (gfl_tf.c:210-215)
    int bi = blockIdx.x;
    while (bi < nbatches) {
        const LpBatch bt = batches[bi];
        if (tid == 0) s.ci[C_BATCH] = gridDim.x + atomicAdd(&ctl->next_chain, 1);   // read after the batch: latency hidden
        const double *__restrict__ Bv = bt.dir == 0 ? beta : betaT;
        double *__restrict__ Zv = bt.dir == 0 ? zr : zc;
        double *__restrict__ Uv = bt.dir == 0 ? ur : uc;
        const int nwu = bt.npts >> 5;
        for (int c = tid; c < bt.nch; c += LP_NT) {
            const LpChain ch = chains[bt.c0 + c];
            s.G[c] = ch.g; s.Lc[c] = ch.L; s.Bc[c] = ch.base; s.Qc[c] = ch.q;
        }
        __syncthreads();
        if (!copy) {
            // ---- A1: stage v = beta + u at S[base + i + 1]; knot hypotheses and splits from the old z ----
            for (int c = 0; c < bt.nch; ++c) {
                const int L = s.Lc[c];
                const bool coop = L > LP_COOP;
                if (!coop && (c % LP_NW) != wid) continue;
                const int base = s.Bc[c];
                const double *__restrict__ B = Bv + s.G[c];
                const double *__restrict__ U = Uv + s.G[c];
                const double *__restrict__ Z = Zv + s.G[c];
                const int ngr = L >> 5;        // groups 0..ngr cover the points 0..L
                for (int g = coop ? wid : 0; g <= ngr; g += coop ? LP_NW : 1) {
                    const int i = (g << 5) + lane;
                    const bool in = i < L;
                    double b = 0.0, u = 0.0, zp = 0.0;
                    if (in) { b = __ldcg(B + i); u = __ldcg(U + i); zp = __ldcg(Z + i); }
                    double zl = __shfl_up_sync(FULL, zp, 1);
                    if (lane == 0) zl = (i > 0 && in) ? __ldcg(Z + i - 1) : zp;
                    if (in) s.S[base + i + 1] = b + u;
                    const bool knot = in && i >= 1 && zp != zl;
                    const unsigned jb = __float_as_uint(fabsf((float)(zp - zl)));
                    unsigned key = knot ? ((((jb >> 5) + 1u) << 5) | (unsigned)lane) : 0u;
                    unsigned best = key;
#pragma unroll
                    for (int o = 1; o < LP_W; o <<= 1) best = max(best, __shfl_xor_sync(FULL, best, o));
                    const bool split = knot && key == best;
                    unsigned kb = __ballot_sync(FULL, split);
                    const unsigned ub = __ballot_sync(FULL, split && zp > zl);
                    unsigned fix = 0u, endb = 0u;
                    if (g == 0) fix |= 1u;
                    if (g == ngr) { endb = 1u << (L & 31); fix |= endb; }
                    kb |= fix;
                    if (lane == 0) {
                        const int wb = (base >> 5) + g;
                        s.K[wb] = kb; s.UP[wb] = ub; s.SPLIT[wb] = kb; s.FIXM[wb] = fix;
                        s.DIRTY[wb] = kb & ~endb; s.FAILED[wb] = 0u;
                    }
                }
            }
            if (tid == 0) { s.ci[C_QN] = 0; s.ci[C_QHEAD] = 0; s.ci[C_NLONG] = 0; s.ci[C_LHEAD] = 0; }
            __syncthreads();
            // ---- A2: per chain, in-place prefix sums of v - c (c = v_0), one warp per chain ----
            for (int c = wid; c < bt.nch; c += LP_NW) {
                const int L = s.Lc[c];
                double *__restrict__ S = s.S + s.Bc[c];
                const double cc = S[1];
                __syncwarp();
                double carry = 0.0;
                int pp = 0;
#pragma unroll 1
                for (; pp + PW <= L; pp += PW) {
                    double *__restrict__ sp = S + pp + lane * PB + 1;
                    double t[PB];
                    double run = 0.0;
#pragma unroll
                    for (int b = 0; b < PB; ++b) { run += sp[b] - cc; t[b] = run; }
                    const double incl = scan_sum_incl(run, lane);
                    const double off = (incl - run) + carry;
#pragma unroll
                    for (int b = 0; b < PB; ++b) sp[b] = t[b] + off;
                    carry += __shfl_sync(FULL, incl, 31);
                }
                if (pp < L) {
                    const int kb = pp + lane * PB;
                    double t[PB];
                    double run = 0.0;
#pragma unroll
                    for (int b = 0; b < PB; ++b) { const int k = kb + b; run += (k < L) ? S[k + 1] - cc : 0.0; t[b] = run; }
                    const double incl = scan_sum_incl(run, lane);
                    const double off = (incl - run) + carry;
#pragma unroll
                    for (int b = 0; b < PB; ++b) { const int k = kb + b; if (k < L) S[k + 1] = t[b] + off; }
                }
                if (lane == 0) { S[0] = 0.0; s.Cc[c] = cc; }
            }
            __syncthreads();
            // ---- B / C: solve the dirty ranges, verify the splits, merge around the failed ones, repeat ----
            for (;;) {
                for (int wi = tid; wi < nwu; wi += LP_NT) {
                    unsigned m = s.DIRTY[wi];
                    if (m) {
                        int pos = atomicAdd(&s.ci[C_QN], __popc(m));
                        while (m) { s.q[pos++] = (unsigned short)((wi << 5) + __ffs(m) - 1); m &= m - 1u; }
                        s.DIRTY[wi] = 0u;
                    }
                }
                __syncthreads();
                lane_pass(s, lam, lane);
                __syncthreads();
                {
                    const int nlong = s.ci[C_NLONG];
                    for (;;) {
                        int idx = 0;
                        if (lane == 0) idx = atomicAdd(&s.ci[C_LHEAD], 1);
                        idx = __shfl_sync(FULL, idx, 0);
                        if (idx >= nlong) break;
                        const int a = s.lq[idx];
                        coop_range(s, lam, a, next_bit(s.SPLIT, a), lane);
                    }
                }
                __syncthreads();
                int failed_any = 0;
                for (int wi = tid; wi < nwu; wi += LP_NT) {
                    unsigned m = s.SPLIT[wi] & ~s.FIXM[wi], fb = 0u;
                    while (m) {
                        const int bpos = __ffs(m) - 1;
                        m &= m - 1u;
                        if (!bend_ok(s, (wi << 5) + bpos, lam)) fb |= 1u << bpos;
                    }
                    if (fb) { s.FAILED[wi] = fb; failed_any = 1; }
                }
                if (!__syncthreads_or(failed_any)) break;
                for (int wi = tid; wi < nwu; wi += LP_NT) s.SPLIT[wi] &= ~s.FAILED[wi];
                if (tid == 0) { s.ci[C_QN] = 0; s.ci[C_QHEAD] = 0; s.ci[C_NLONG] = 0; s.ci[C_LHEAD] = 0; }
                __syncthreads();
                for (int wi = tid; wi < nwu; wi += LP_NT) {
                    unsigned m = s.FAILED[wi];
                    while (m) {
                        const int p = (wi << 5) + __ffs(m) - 1;
                        m &= m - 1u;
                        const int ps = prev_bit_le(s.SPLIT, p);      // the surviving split on the left starts the merged range
                        atomicOr(&s.DIRTY[ps >> 5], 1u << (ps & 31));
                    }
                    s.FAILED[wi] = 0u;
                }
                __syncthreads();
            }
        }
        // ---- D: z = c + slope of the piece, u += beta - z, residual sums (gfl_graph_fl.c:306-365) ----
        for (int c = 0; c < bt.nch; ++c) {
            const int L = s.Lc[c];
            const bool coop = L > LP_COOP;
            if (!coop && (c % LP_NW) != wid) continue;
            const int base = s.Bc[c];
            const double *__restrict__ B = Bv + s.G[c];
            double *__restrict__ U = Uv + s.G[c];
            double *__restrict__ Z = Zv + s.G[c];
            const double cc = copy ? 0.0 : s.Cc[c];
            double ap = 0.0, ad = 0.0;
            const int ngr = (L - 1) >> 5;
            for (int g = coop ? wid : 0; g <= ngr; g += coop ? LP_NW : 1) {
                const int i = (g << 5) + lane;
                if (i < L) {
                    const double b = __ldcg(B + i), u = __ldcg(U + i), zo = __ldcg(Z + i);
                    double z;
                    if (copy) z = b + u;
                    else {
                        const int pa = prev_bit_le(s.K, base + i), pb = next_bit(s.K, base + i);
                        const double ya = s.S[pa] + side_off(s, pa, lam), yb = s.S[pb] + side_off(s, pb, lam);
                        z = cc + (yb - ya) * recip_t(s.R, pb - pa);
                    }
                    const double r = b - z;
                    __stcg(U + i, u + r);
                    __stcg(Z + i, z);
                    ap += r * r;
                    const double dd = alpha * (z - zo);
                    ad += dd * dd;
                }
            }
            ap = warp_sum(ap); ad = warp_sum(ad);
            const int q = s.Qc[c];
            if (lane == 0) {
                if (coop) { partials[2 * (LP_NW * q + wid)] = ap; partials[2 * (LP_NW * q + wid) + 1] = ad; }
                else {
                    partials[2 * (LP_NW * q)] = ap; partials[2 * (LP_NW * q) + 1] = ad;
                    for (int w = 1; w < LP_NW; ++w) { partials[2 * (LP_NW * q + w)] = 0.0; partials[2 * (LP_NW * q + w) + 1] = 0.0; }
                }
            }
        }
        __syncthreads();
        bi = s.ci[C_BATCH];
        __syncthreads();
    }
}

} // namespace

// ==========================================================================
void lane_plan_build(LanePlan &plan, const Grid &g, int r0, int r1, int c0, int c1, int n_sm) {
    plan.usable = false;
    int dev = 0, smem_max = 0;
    FB_CUDA(cudaGetDevice(&dev));
    FB_CUDA(cudaDeviceGetAttribute(&smem_max, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev));
    const int lmax = g.row_len(0);
    int cap = std::max(2048, ((lmax + 1 + 31) / 32) * 32);
    if (const char *e = std::getenv("FB_LP_CAP")) cap = std::max(((lmax + 1 + 31) / 32) * 32, (std::atoi(e) / 32) * 32);
    if (cap > 65535 - 64 || lp_smem_bytes(cap) > (size_t)smem_max / 2) return;   // at least two CTAs per SM
    std::vector<LpChain> chains;
    std::vector<LpBatch> batches;
    std::vector<int64_t> cells;
    auto add_dir = [&](int dir, int lo, int hi) {
        LpBatch cur{(int)chains.size(), 0, dir, 0};
        int64_t cur_cells = 0;
        auto flush = [&]() {
            if (cur.nch) { batches.push_back(cur); cells.push_back(cur_cells); }
            cur = LpBatch{(int)chains.size(), 0, dir, 0}; cur_cells = 0;
        };
        for (int x = lo; x < hi; ++x) {
            LpChain ch;
            if (dir == 0) { ch.L = g.row_len(x); ch.g = g.rm_row_start(x); ch.q = 2 * x; }
            else { ch.L = g.col_len(x); ch.g = g.cm_col_start(x); ch.q = 2 * (g.n - 1 - x) + 1; }
            const int need = ((ch.L + 1 + 31) / 32) * 32;
            if (cur.npts + need > cap) flush();
            ch.base = cur.npts; ch.pad = 0;
            chains.push_back(ch);
            cur.npts += need; cur.nch += 1; cur_cells += ch.L;
        }
        flush();
    };
    add_dir(0, r0, std::min(r1, g.n - 1));      // rows 0..n-2 are chains
    add_dir(1, std::max(c0, 1), c1);            // columns 1..n-1
    if (batches.empty()) { plan.nbatches = 0; plan.usable = true; plan.cap = cap; plan.smem = lp_smem_bytes(cap); plan.grid = 1; return; }
    // most cells first: the dynamic queue then ends with the small batches
    std::vector<int> order(batches.size());
    for (size_t k = 0; k < order.size(); ++k) order[k] = (int)k;
    std::stable_sort(order.begin(), order.end(), [&](int a, int b) { return cells[a] > cells[b]; });
    std::vector<LpBatch> sorted(batches.size());
    for (size_t k = 0; k < order.size(); ++k) sorted[k] = batches[order[k]];
    plan.chains.alloc(chains.size()); plan.batches.alloc(sorted.size());
    FB_CUDA(cudaMemcpy(plan.chains.p, chains.data(), chains.size() * sizeof(LpChain), cudaMemcpyHostToDevice));
    FB_CUDA(cudaMemcpy(plan.batches.p, sorted.data(), sorted.size() * sizeof(LpBatch), cudaMemcpyHostToDevice));
    plan.cap = cap; plan.nbatches = (int)sorted.size(); plan.smem = lp_smem_bytes(cap);
    FB_CUDA(cudaFuncSetAttribute(k_prox_lane, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)plan.smem));
    int per_sm = 1;
    FB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_prox_lane, LP_NT, plan.smem));
    plan.grid = std::max(1, std::min(plan.nbatches, n_sm * std::max(1, per_sm)));
    plan.usable = true;
}

void lane_prox_launch(const LanePlan &plan, GflCtl *ctl, const double *beta, const double *betaT, double *zr, double *ur,
                      double *zc, double *uc, double *partials, cudaStream_t st) {
    if (!plan.nbatches) return;
    k_prox_lane<<<plan.grid, LP_NT, plan.smem, st>>>(ctl, plan.chains.p, plan.batches.p, plan.nbatches, plan.cap, beta, betaT,
                                                     zr, ur, zc, uc, partials);
}

} // namespace fb
