// fb_prox_aset.cu -- parallel active-set chain prox: the z / u update of one ADMM iteration
// (update_z with tf_dp_weight, update_u, primal_resnorm, dual_resnorm: gfl_graph_fl.c:267-365, gfl_tf.c:184-321).
//
// The 1-D fused-lasso minimiser of a chain is the derivative of the taut string through the tube S[k] -+ lam around the
// prefix sums S of v = beta + u, from (0, 0) to (L, S[L]): the polygon through its KNOTS (points where it touches the
// tube: position, side). A knot set is the solution iff (KKT conditions of a strictly convex problem)
//   (i)  between two knots the polygon stays inside the tube at every point, and
//   (ii) at every knot it bends towards the tube edge it touches.
// Both tests are independent per point / per knot and so are the repairs: the point of a piece that is furthest
// outside becomes a knot, a knot that bends the wrong way goes. Starting from the knots of the PREVIOUS ADMM iterate
// (its jumps), all pieces and knots of all chains of a batch are treated at once, pass after pass, until nothing
// changes; pieces whose end knots did not change are not looked at again. No step is sequential along a chain: the
// reference's dynamic programme (and the taut-string search of k_prox) walk every chain from one end to the other.
// On the iterates of real solves 2-4 passes settle a chain (12 from a cold start); the result does not depend on the
// starting set. Should a batch not settle within `maxpass` passes (rounding-level ties could in principle flip a knot
// for ever; never seen), its chains are solved by the sequential taut string, one lane per chain.
// tests/test_aset_prox_model.py pins a scalar C model of these passes to the reference's tf_dp_weight.
//
// Per batch of chains held in shared memory (one CTA):
//   A  stream beta, u, z in: v, the per-chain prefix sums S, the knot masks K / UP from the jumps of the old z
//   P  passes:  P0 apply the pending knot changes; per 32-point word the last knot before / first knot after it
//               P1 slope of every piece with a changed end knot (one thread per word walks its knots)
//               P2 bend test of the knots next to such a piece; tube test of its points (one thread per point), the
//                  worst point of a piece by atomicMax on a per-piece key
//               P3 the worst points become knots
//   D  stream out: z = c + slope of the piece a cell lies on, u += beta - z, the residual sums per chain
// Every chain is computed from its own inputs only, in a fixed order: results do not depend on how chains are batched
// or which rank owns them.
#include "fb_prox_aset.cuh"
#include <algorithm>
#include <cstdio>
#include <cstdlib>

namespace fb {

namespace {
constexpr unsigned FULL = 0xffffffffu;
constexpr int AS_NT = 256, AS_NW = AS_NT / 32;
constexpr int AS_RT = 256;                      // reciprocal table R[m] = 1/m
constexpr int AS_PB = 9, AS_PW = 32 * AS_PB;    // blocked prefix sums: AS_PB consecutive points per lane (odd: no bank conflicts)
constexpr int AS_SPLIT = 2 * AS_PW;             // chains longer than this are scanned by all warps of the CTA
constexpr int AS_UN = 2;                        // words whose global loads are in flight together per warp
constexpr double AS_EPS = 0x1p-46;              // rounding margin of the tube test (see pass_points)
constexpr int NONE16 = 0xffff;
enum { C_BATCH = 0, C_FLAG, C_COUNT = 4 };

struct Sm {
    double *S, *SLP, *R, *Cc, *SPAN;
    long long *G;
    unsigned *PM, *K, *UP, *FIXM, *ENDM, *DIRTY, *REM, *ADD, *CH;
    int *Lc, *Bc, *Qc, *ci;
    unsigned short *PREV, *NEXT, *CNT, *GC;
};

__host__ __device__ inline size_t as_align8(size_t x) { return (x + 7) & ~size_t(7); }
__host__ __device__ inline size_t as_smem_bytes(int cap) {
    const size_t nw = (size_t)cap / 32;
    size_t b = (size_t)cap * 16 + AS_RT * 8 + nw * 8 + AS_NW * 8 + nw * 8;   // S, SLP, R, Cc, SPAN, G
    b += (size_t)cap * 4 + nw * 4 * 8 + nw * 4 * 3 + C_COUNT * 4;             // PM, 8 masks, Lc Bc Qc, ci
    b = as_align8(b) + nw * 2 * 4;                                             // PREV NEXT CNT GC
    return as_align8(b);
}
__device__ __forceinline__ Sm carve(unsigned char *p, int cap) {
    const size_t nw = (size_t)cap / 32;
    Sm s;
    s.S = reinterpret_cast<double *>(p); p += (size_t)cap * 8;
    s.SLP = reinterpret_cast<double *>(p); p += (size_t)cap * 8;
    s.R = reinterpret_cast<double *>(p); p += AS_RT * 8;
    s.Cc = reinterpret_cast<double *>(p); p += nw * 8;
    s.SPAN = reinterpret_cast<double *>(p); p += AS_NW * 8;
    s.G = reinterpret_cast<long long *>(p); p += nw * 8;
    s.PM = reinterpret_cast<unsigned *>(p); p += (size_t)cap * 4;
    s.K = reinterpret_cast<unsigned *>(p); p += nw * 4;
    s.UP = reinterpret_cast<unsigned *>(p); p += nw * 4;
    s.FIXM = reinterpret_cast<unsigned *>(p); p += nw * 4;
    s.ENDM = reinterpret_cast<unsigned *>(p); p += nw * 4;
    s.DIRTY = reinterpret_cast<unsigned *>(p); p += nw * 4;
    s.REM = reinterpret_cast<unsigned *>(p); p += nw * 4;
    s.ADD = reinterpret_cast<unsigned *>(p); p += nw * 4;
    s.CH = reinterpret_cast<unsigned *>(p); p += nw * 4;
    s.Lc = reinterpret_cast<int *>(p); p += nw * 4;
    s.Bc = reinterpret_cast<int *>(p); p += nw * 4;
    s.Qc = reinterpret_cast<int *>(p); p += nw * 4;
    s.ci = reinterpret_cast<int *>(p); p += C_COUNT * 4;
    p = reinterpret_cast<unsigned char *>(as_align8(reinterpret_cast<size_t>(p)));
    s.PREV = reinterpret_cast<unsigned short *>(p); p += nw * 2;
    s.NEXT = reinterpret_cast<unsigned short *>(p); p += nw * 2;
    s.CNT = reinterpret_cast<unsigned short *>(p); p += nw * 2;
    s.GC = reinterpret_cast<unsigned short *>(p);
    return s;
}

__device__ __forceinline__ double recip_t(const double *__restrict__ R, int m) { return m < AS_RT ? R[m] : 1.0 / (double)m; }
// offset of the string from S at a knot: 0 at a chain end, +lam on the upper tube edge, -lam on the lower one
__device__ __forceinline__ double off_bits(unsigned fixw, unsigned upw, int bit, double lam) {
    return ((fixw >> bit) & 1u) ? 0.0 : (((upw >> bit) & 1u) ? lam : -lam);
}
__device__ __forceinline__ double off_at(const Sm &s, int p, double lam) { return off_bits(s.FIXM[p >> 5], s.UP[p >> 5], p & 31, lam); }

__device__ __forceinline__ double scan_sum_incl(double v, int lane) {
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const double t = __shfl_up_sync(FULL, v, o); if (lane >= o) v += t; }
    return v;
}

// in-place prefix sums of (x - cc) over the n points S[1..n] (S already offset), continuing from `carry`; returns the new carry
__device__ __forceinline__ double scan_span(double *__restrict__ S, const int n, const double cc, double carry, const int lane) {
    int pp = 0;
#pragma unroll 1
    for (; pp + AS_PW <= n; pp += AS_PW) {
        double *__restrict__ sp = S + pp + lane * AS_PB + 1;
        double t[AS_PB];
        double run = 0.0;
#pragma unroll
        for (int b = 0; b < AS_PB; ++b) { run += sp[b] - cc; t[b] = run; }
        const double incl = scan_sum_incl(run, lane);
        const double o = (incl - run) + carry;
#pragma unroll
        for (int b = 0; b < AS_PB; ++b) sp[b] = t[b] + o;
        carry += __shfl_sync(FULL, incl, 31);
    }
    if (pp < n) {
        const int kb = pp + lane * AS_PB;
        double t[AS_PB];
        double run = 0.0;
#pragma unroll
        for (int b = 0; b < AS_PB; ++b) { const int k = kb + b; run += (k < n) ? S[k + 1] - cc : 0.0; t[b] = run; }
        const double incl = scan_sum_incl(run, lane);
        const double o = (incl - run) + carry;
#pragma unroll
        for (int b = 0; b < AS_PB; ++b) { const int k = kb + b; if (k < n) S[k + 1] = t[b] + o; }
        carry += __shfl_sync(FULL, incl, 31);
    }
    return carry;
}

// P0 (one warp): K <- (K & ~REM) | ADD; per word the last knot before it, the first knot after it and the number of
// changed knots before it. Returns (all lanes) whether any knot changed.
__device__ __forceinline__ bool pass_begin(const Sm &s, const int nwu, const int lane) {
    int carry_prev = -1, cnt_carry = 0;
    unsigned anych = 0u;
    const int nblk = (nwu + 31) >> 5;
    for (int blk = 0; blk < nblk; ++blk) {
        const int w = (blk << 5) + lane;
        const bool valid = w < nwu;
        unsigned k = 0u, ch = 0u;
        if (valid) {
            const unsigned rem = s.REM[w], add = s.ADD[w];
            k = (s.K[w] & ~rem) | add; ch = rem | add;
            s.K[w] = k; s.CH[w] = ch; s.REM[w] = 0u; s.ADD[w] = 0u;
        }
        int m = k ? (w << 5) + 31 - __clz(k) : -1;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const int t = __shfl_up_sync(FULL, m, o); if (lane >= o) m = max(m, t); }
        int ex = __shfl_up_sync(FULL, m, 1);
        if (lane == 0) ex = -1;
        ex = max(ex, carry_prev);
        int c = __popc(ch), ci = c;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const int t = __shfl_up_sync(FULL, ci, o); if (lane >= o) ci += t; }
        if (valid) { s.PREV[w] = (unsigned short)(ex < 0 ? NONE16 : ex); s.CNT[w] = (unsigned short)(cnt_carry + ci - c); }
        carry_prev = max(carry_prev, __shfl_sync(FULL, m, 31));
        cnt_carry += __shfl_sync(FULL, ci, 31);
        anych |= ch;
    }
    int carry_next = NONE16;
    for (int blk = nblk - 1; blk >= 0; --blk) {
        const int w = (blk << 5) + lane;
        const bool valid = w < nwu;
        const unsigned k = valid ? s.K[w] : 0u;
        int m = k ? (w << 5) + __ffs(k) - 1 : NONE16;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const int t = __shfl_down_sync(FULL, m, o); if (lane + o < 32) m = min(m, t); }
        int ex = __shfl_down_sync(FULL, m, 1);
        if (lane == 31) ex = NONE16;
        ex = min(ex, carry_next);
        if (valid) s.NEXT[w] = (unsigned short)ex;
        carry_next = min(carry_next, __shfl_sync(FULL, m, 0));
    }
    return __any_sync(FULL, anych != 0u);
}

// P1 (one thread per word): slope of every piece starting in word w whose end knots changed (all pieces when `all`)
__device__ __forceinline__ void pass_slopes(const Sm &s, const int w, const double lam, const bool all) {
    const unsigned kall = s.K[w], upw = s.UP[w], fixw = s.FIXM[w], chw = s.CH[w];
    const int cntw = s.CNT[w];
    unsigned m = kall & ~s.ENDM[w], dbits = 0u;
    while (m) {
        const int bit = __ffs(m) - 1;
        m &= m - 1u;
        const int p = (w << 5) + bit;
        const unsigned hm = bit == 31 ? 0u : (kall & (0xfffffffeu << bit));
        const int b = hm ? (w << 5) + __ffs(hm) - 1 : (int)s.NEXT[w];
        bool dirty = all;
        if (!all) {
            const int wb = b >> 5, bb = b & 31;
            const int nle = (int)s.CNT[wb] + __popc(s.CH[wb] & (0xffffffffu >> (31 - bb)));
            const int nlt = cntw + __popc(chw & ((1u << bit) - 1u));
            dirty = nle > nlt;
        }
        if (dirty) {
            const double ya = s.S[p] + off_bits(fixw, upw, bit, lam);
            const double yb = s.S[b] + off_at(s, b, lam);
            s.SLP[p] = (yb - ya) * recip_t(s.R, b - p);
            dbits |= 1u << bit;
        }
    }
    s.DIRTY[w] = dbits;
}

// P2a (one thread per word): the knots of word w next to a piece whose slope changed must bend towards their tube edge
__device__ __forceinline__ void pass_bends(const Sm &s, const int w) {
    const unsigned kall = s.K[w], upw = s.UP[w], dw = s.DIRTY[w];
    unsigned m = kall & ~s.FIXM[w], rem = 0u;
    while (m) {
        const int bit = __ffs(m) - 1;
        m &= m - 1u;
        const unsigned lm = kall & ((1u << bit) - 1u);
        const int a = lm ? (w << 5) + 31 - __clz(lm) : (int)s.PREV[w];
        const unsigned da = lm ? (dw >> (a & 31)) & 1u : (s.DIRTY[a >> 5] >> (a & 31)) & 1u;
        if (!(da | ((dw >> bit) & 1u))) continue;
        const double bend = s.SLP[(w << 5) + bit] - s.SLP[a];
        const bool ok = ((upw >> bit) & 1u) ? bend > 0.0 : bend < 0.0;
        if (!ok) rem |= 1u << bit;
    }
    s.REM[w] = rem;
}

// P2b (one warp per word, one lane per point): the points of the pieces whose slope changed against the tube. A point
// counts as outside only beyond the rounding error of the expression itself: without the margin a knot whose bend is
// zero up to rounding would be removed and added again for ever.
__device__ __forceinline__ void pass_points(const Sm &s, const int g, const double lam, const int lane) {
    const unsigned kw = s.K[g], dw = s.DIRTY[g];
    const int pprev = (kw & 1u) ? -1 : (int)s.PREV[g];       // the piece entering the word from the left
    bool dprev = false;
    double offprev = 0.0;
    if (pprev >= 0) { dprev = (s.DIRTY[pprev >> 5] >> (pprev & 31)) & 1u; offprev = off_at(s, pprev, lam); }
    if (!dw && !dprev) return;
    const unsigned endw = s.ENDM[g];
    const int limit = endw ? __ffs(endw) - 1 : 32;              // points beyond the chain end do not exist
    if (lane >= limit || ((kw >> lane) & 1u)) return;
    const unsigned lm = kw & ((1u << lane) - 1u);
    int pa; double oa; bool dirty;
    if (lm) {
        const int bit = 31 - __clz(lm);
        pa = (g << 5) + bit; oa = off_bits(s.FIXM[g], s.UP[g], bit, lam); dirty = (dw >> bit) & 1u;
    } else { pa = pprev; oa = offprev; dirty = dprev; }
    if (!dirty) return;
    const int p = (g << 5) + lane;
    const double sp = s.S[p];
    const double y = (s.S[pa] + oa) + s.SLP[pa] * (double)(p - pa);
    const double d = y - sp;
    const double e = fabs(d) - lam;
    if (e > 0.0 && e > AS_EPS * (fabs(sp) + fabs(y) + lam)) {
        // key: the violation as a float (truncated: any point outside is a valid new knot), the side, the offset
        unsigned bits = __float_as_uint((float)e);
        const unsigned key = (bits & 0xffffc000u) | (d > 0.0 ? 0x2000u : 0u) | (unsigned)(p - pa);
        atomicMax(&s.PM[pa], key);
    }
}

// P3 (one thread per word): the worst point of every piece that starts in word w becomes a knot
__device__ __forceinline__ void pass_adds(const Sm &s, const int w) {
    unsigned m = s.K[w] & s.DIRTY[w];
    while (m) {
        const int bit = __ffs(m) - 1;
        m &= m - 1u;
        const int p = (w << 5) + bit;
        const unsigned key = s.PM[p];
        if (key) {
            s.PM[p] = 0u;
            const int q = p + (int)(key & 0x1fffu);
            const unsigned qb = 1u << (q & 31);
            atomicOr(&s.ADD[q >> 5], qb);
            if (key & 0x2000u) atomicOr(&s.UP[q >> 5], qb); else atomicAnd(&s.UP[q >> 5], ~qb);
        }
    }
}

// Fallback (one lane per chain): the sequential taut string from the chain start; rewrites the knot masks of the chain
__device__ void chain_sequential(const Sm &s, const int base, const int L, const double lam) {
    const double INF = __longlong_as_double(0x7ff0000000000000LL);
    for (int w = base >> 5; w <= (base + L) >> 5; ++w) { s.K[w] = s.FIXM[w]; s.UP[w] = 0u; s.REM[w] = 0u; s.ADD[w] = 0u; }
    const double *__restrict__ S = s.S + base;
    int k0 = 0, k = 1, kH = 0, kL = 0;
    double y0 = S[0], H = INF, Lw = -INF;
    for (;;) {
        const double q = S[k] - y0, r = 1.0 / (double)(k - k0);
        const bool last = (k == L);
        const double h = (q + (last ? 0.0 : lam)) * r, l = (q + (last ? 0.0 : -lam)) * r;
        const bool vu = l > H, vd = h < Lw;
        if (vu || vd) {
            const int kk = vu ? kH : kL;
            const int p = base + kk;
            s.K[p >> 5] |= 1u << (p & 31);
            if (vu) s.UP[p >> 5] |= 1u << (p & 31);
            k0 = kk; y0 = S[kk] + (vu ? lam : -lam);
            H = INF; Lw = -INF; kH = kL = kk; k = kk + 1;
        } else {
            if (h <= H) { H = h; kH = k; }
            if (l >= Lw) { Lw = l; kL = k; }
            if (last) break;
            ++k;
        }
    }
}

__global__ void __launch_bounds__(AS_NT)
k_prox_aset(GflCtl *__restrict__ ctl, const AsChain *__restrict__ chains, const AsBatch *__restrict__ batches,
            const int nbatches, const int cap, const int maxpass, const double *__restrict__ beta, const double *__restrict__ betaT,
            double *__restrict__ zr, double *__restrict__ ur, double *__restrict__ zc, double *__restrict__ uc,
            double *__restrict__ partials, unsigned long long *__restrict__ stats) {
    if (ctl->done) return;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const Sm s = carve(smem_raw, cap);
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    for (int m = tid; m < AS_RT; m += AS_NT) s.R[m] = m ? 1.0 / (double)m : 0.0;
    const double alpha = ctl->alpha;
    const double lam = ctl->lam / (alpha / 2.0);
    const bool copy = (lam == 0.0);          // tf_dp_weight copies y (gfl_tf.c:210-215)
    int bi = blockIdx.x;
    while (bi < nbatches) {
        const AsBatch bt = batches[bi];
        if (tid == 0) s.ci[C_BATCH] = gridDim.x + atomicAdd(&ctl->next_chain, 1);   // read after the batch: latency hidden
        const double *__restrict__ Bv = bt.dir == 0 ? beta : betaT;
        double *__restrict__ Zv = bt.dir == 0 ? zr : zc;
        double *__restrict__ Uv = bt.dir == 0 ? ur : uc;
        const int nwu = bt.npts >> 5;
        long long tc0 = 0, tc1 = 0, tc2 = 0, tc3 = 0; int npass = 0, fell = 0;
        if (stats) tc0 = clock64();
        for (int c = tid; c < bt.nch; c += AS_NT) {
            const AsChain ch = chains[bt.c0 + c];
            s.G[c] = ch.g; s.Lc[c] = ch.L; s.Bc[c] = ch.base; s.Qc[c] = ch.q;
            for (int w = ch.base >> 5; w <= (ch.base + ch.L) >> 5; ++w) s.GC[w] = (unsigned short)c;
        }
        __syncthreads();
        if (!copy) {
            // ---- A1: stage v = beta + u at S[point + 1]; the jumps of the old z are the first knot set ----
#pragma unroll 1
            for (int g0 = wid; g0 < nwu; g0 += AS_NW * AS_UN) {
                double bq[AS_UN], uq[AS_UN], zq[AS_UN], z0[AS_UN];
#pragma unroll
                for (int x = 0; x < AS_UN; ++x) {
                    const int g = g0 + x * AS_NW;
                    bq[x] = uq[x] = zq[x] = z0[x] = 0.0;
                    if (g < nwu) {
                        const int c = s.GC[g];
                        const int i = (g << 5) - s.Bc[c] + lane;
                        const long long o = s.G[c] + i;
                        if (i < s.Lc[c]) {
                            bq[x] = __ldcg(Bv + o); uq[x] = __ldcg(Uv + o); zq[x] = __ldcg(Zv + o);
                            if (lane == 0 && i > 0) z0[x] = __ldcg(Zv + o - 1);
                        }
                    }
                }
#pragma unroll
                for (int x = 0; x < AS_UN; ++x) {
                    const int g = g0 + x * AS_NW;
                    if (g < nwu) {
                        const int c = s.GC[g];
                        const int base = s.Bc[c], L = s.Lc[c];
                        const int i = (g << 5) - base + lane;
                        const bool in = i < L;
                        const double zp = zq[x];
                        double zl = __shfl_up_sync(FULL, zp, 1);
                        if (lane == 0) zl = z0[x];
                        const double v = bq[x] + uq[x];
                        if (in) s.S[(g << 5) + lane + 1] = v;
                        if (i == 0) s.Cc[c] = v;
                        const bool knot = in && i >= 1 && zp != zl;
                        const unsigned kb = __ballot_sync(FULL, knot);
                        const unsigned ub = __ballot_sync(FULL, knot && zp > zl);
                        s.PM[(g << 5) + lane] = 0u;
                        if (lane == 0) {
                            const unsigned fixs = (g == (base >> 5)) ? 1u : 0u;
                            const unsigned ende = (((base + L) >> 5) == g) ? 1u << ((base + L) & 31) : 0u;
                            s.K[g] = kb | fixs | ende; s.UP[g] = ub; s.FIXM[g] = fixs | ende; s.ENDM[g] = ende;
                            s.DIRTY[g] = 0u; s.REM[g] = 0u; s.ADD[g] = 0u; s.CH[g] = 0u;
                        }
                    }
                }
            }
            __syncthreads();
            if (stats) tc1 = clock64();
            // ---- A2: per chain, in-place prefix sums of v - c (c = v_0) ----
            {
                int nshort = 0;
                for (int c = 0; c < bt.nch; ++c) {
                    const int L = s.Lc[c];
                    double *__restrict__ S = s.S + s.Bc[c];
                    const double cc = s.Cc[c];
                    if (L > AS_SPLIT) {        // all warps: contiguous spans of whole scan blocks (a function of L only)
                        const int nblk = (L + AS_PW - 1) / AS_PW, per = (nblk + AS_NW - 1) / AS_NW;
                        const int p0 = min(L, wid * per * AS_PW), p1 = min(L, (wid + 1) * per * AS_PW);
                        const double tot = p1 > p0 ? scan_span(S + p0, p1 - p0, cc, 0.0, lane) : 0.0;
                        if (lane == 0) s.SPAN[wid] = tot;
                        __syncthreads();
                        double o = 0.0;
                        for (int x = 0; x < wid; ++x) o += s.SPAN[x];
                        if (wid > 0) for (int k = p0 + lane; k < p1; k += 32) S[k + 1] += o;
                        if (tid == 0) S[0] = 0.0;
                        __syncthreads();
                    } else {
                        if ((nshort % AS_NW) == wid) {
                            scan_span(S, L, cc, 0.0, lane);
                            if (lane == 0) S[0] = 0.0;
                        }
                        ++nshort;
                    }
                }
            }
            __syncthreads();
            if (stats) tc2 = clock64();
            // ---- P: passes until no knot changes ----
            bool first = true;
            for (;;) {
                if (wid == 0) {
                    const bool any = pass_begin(s, nwu, lane);
                    if (lane == 0) s.ci[C_FLAG] = (first || any) ? 1 : 0;
                }
                __syncthreads();
                if (!s.ci[C_FLAG]) break;
                if (npass >= maxpass) { fell = 1; break; }
                ++npass;
                if (tid < nwu) pass_slopes(s, tid, lam, first);
                __syncthreads();
                if (tid < nwu) pass_bends(s, tid);
                for (int g = wid; g < nwu; g += AS_NW) pass_points(s, g, lam, lane);
                __syncthreads();
                if (tid < nwu) pass_adds(s, tid);
                __syncthreads();
                first = false;
            }
            if (fell) {     // not settled: the sequential taut string, one lane per chain; then every slope again
                if (tid < bt.nch) chain_sequential(s, s.Bc[tid], s.Lc[tid], lam);
                __syncthreads();
                if (wid == 0) pass_begin(s, nwu, lane);
                __syncthreads();
                if (tid < nwu) pass_slopes(s, tid, lam, true);
                __syncthreads();
            }
        }
        if (stats) tc3 = clock64();
        // ---- D: z = c + slope of the piece, u += beta - z, residual sums (gfl_graph_fl.c:306-365) ----
        double *__restrict__ WP = reinterpret_cast<double *>(s.PM);     // per word: the two partial sums (PM is free now)
#pragma unroll 1
        for (int g0 = wid; g0 < nwu; g0 += AS_NW * AS_UN) {
            double bq[AS_UN], uq[AS_UN], zq[AS_UN];
#pragma unroll
            for (int x = 0; x < AS_UN; ++x) {
                const int g = g0 + x * AS_NW;
                bq[x] = uq[x] = zq[x] = 0.0;
                if (g < nwu) {
                    const int c = s.GC[g];
                    const int i = (g << 5) - s.Bc[c] + lane;
                    const long long o = s.G[c] + i;
                    if (i < s.Lc[c]) { bq[x] = __ldcg(Bv + o); uq[x] = __ldcg(Uv + o); zq[x] = __ldcg(Zv + o); }
                }
            }
#pragma unroll
            for (int x = 0; x < AS_UN; ++x) {
                const int g = g0 + x * AS_NW;
                if (g < nwu) {
                    const int c = s.GC[g];
                    const int i = (g << 5) - s.Bc[c] + lane;
                    const bool in = i < s.Lc[c];
                    double z = bq[x] + uq[x];
                    if (!copy) {
                        // cell i lies between the points p = 32 g + lane and p + 1: its piece starts at the last knot <= p
                        const unsigned lm = s.K[g] & (0xffffffffu >> (31 - lane));
                        const int pa = !in ? 0 : lm ? (g << 5) + 31 - __clz(lm) : (int)s.PREV[g];
                        z = s.Cc[c] + s.SLP[pa];
                    }
                    double ap = 0.0, ad = 0.0;
                    if (in) {
                        const long long o = s.G[c] + i;
                        const double r = bq[x] - z;
                        __stcg(Uv + o, uq[x] + r);
                        __stcg(Zv + o, z);
                        ap = r * r;
                        const double dd = alpha * (z - zq[x]);
                        ad = dd * dd;
                    }
                    ap = warp_sum(ap); ad = warp_sum(ad);
                    if (lane == 0) { WP[4 * g] = ap; WP[4 * g + 1] = ad; }
                }
            }
        }
        __syncthreads();
        for (int c = tid; c < bt.nch; c += AS_NT) {      // per chain, its words in order
            const int base = s.Bc[c], L = s.Lc[c];
            double ap = 0.0, ad = 0.0;
            for (int w = base >> 5; w <= (base + L - 1) >> 5; ++w) { ap += WP[4 * w]; ad += WP[4 * w + 1]; }
            const int q = s.Qc[c];
            partials[4 * q] = ap; partials[4 * q + 1] = ad; partials[4 * q + 2] = 0.0; partials[4 * q + 3] = 0.0;
        }
        if (stats && tid == 0) {   // FB_ASET_STATS development aid: cycles per phase summed over batches
            const long long tc4 = clock64();
            atomicAdd(&stats[0], (unsigned long long)(tc1 - tc0)); atomicAdd(&stats[1], (unsigned long long)(tc2 - tc1));
            atomicAdd(&stats[2], (unsigned long long)(tc3 - tc2)); atomicAdd(&stats[3], (unsigned long long)(tc4 - tc3));
            atomicAdd(&stats[4], (unsigned long long)npass); atomicAdd(&stats[5], 1ull);
            atomicMax(&stats[6], (unsigned long long)npass); atomicAdd(&stats[7], (unsigned long long)fell);
        }
        bi = s.ci[C_BATCH];
        __syncthreads();
    }
}

} // namespace

// ==========================================================================
void aset_plan_build(AsetPlan &plan, const Grid &g, int r0, int r1, int c0, int c1, int n_sm) {
    plan.usable = false;
    int dev = 0, smem_max = 0;
    FB_CUDA(cudaGetDevice(&dev));
    FB_CUDA(cudaDeviceGetAttribute(&smem_max, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev));
    const int lmax = g.row_len(0);
    int cap = std::max(1024, ((lmax + 1 + 31) / 32) * 32);
    if (const char *e = std::getenv("FB_ASET_CAP")) cap = std::max(((lmax + 1 + 31) / 32) * 32, (std::atoi(e) / 32) * 32);
    if (cap > 8160 || as_smem_bytes(cap) > (size_t)smem_max / 2) return;   // 13-bit offsets in the piece keys; two CTAs per SM
    std::vector<AsChain> chains;
    std::vector<AsBatch> batches;
    std::vector<int64_t> cells;
    auto add_dir = [&](int dir, int lo, int hi) {
        AsBatch cur{(int)chains.size(), 0, dir, 0};
        int64_t cur_cells = 0;
        auto flush = [&]() {
            if (cur.nch) { batches.push_back(cur); cells.push_back(cur_cells); }
            cur = AsBatch{(int)chains.size(), 0, dir, 0}; cur_cells = 0;
        };
        for (int x = lo; x < hi; ++x) {
            AsChain ch;
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
    plan.maxpass = 64;
    if (const char *e = std::getenv("FB_ASET_MAXPASS")) plan.maxpass = std::max(0, std::atoi(e));   // test aid: 0 = always the fallback
    if (batches.empty()) { plan.nbatches = 0; plan.usable = true; plan.cap = cap; plan.smem = as_smem_bytes(cap); plan.grid = 1; return; }
    // most cells first: the dynamic queue then ends with the small batches
    std::vector<int> order(batches.size());
    for (size_t k = 0; k < order.size(); ++k) order[k] = (int)k;
    std::stable_sort(order.begin(), order.end(), [&](int a, int b) { return cells[a] > cells[b]; });
    std::vector<AsBatch> sorted(batches.size());
    for (size_t k = 0; k < order.size(); ++k) sorted[k] = batches[order[k]];
    plan.chains.alloc(chains.size()); plan.batches.alloc(sorted.size());
    FB_CUDA(cudaMemcpy(plan.chains.p, chains.data(), chains.size() * sizeof(AsChain), cudaMemcpyHostToDevice));
    FB_CUDA(cudaMemcpy(plan.batches.p, sorted.data(), sorted.size() * sizeof(AsBatch), cudaMemcpyHostToDevice));
    plan.cap = cap; plan.nbatches = (int)sorted.size(); plan.smem = as_smem_bytes(cap);
    FB_CUDA(cudaFuncSetAttribute(k_prox_aset, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)plan.smem));
    int per_sm = 1;
    FB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_prox_aset, AS_NT, plan.smem));
    plan.grid = std::max(1, std::min(plan.nbatches, n_sm * std::max(1, per_sm)));
    if (std::getenv("FB_ASET_STATS")) plan.stats.alloc(8);
    plan.usable = true;
}

void aset_prox_launch(const AsetPlan &plan, GflCtl *ctl, const double *beta, const double *betaT, double *zr, double *ur,
                      double *zc, double *uc, double *partials, cudaStream_t st) {
    if (!plan.nbatches) return;
    if (plan.stats.p) FB_CUDA(cudaMemsetAsync(plan.stats.p, 0, 8 * sizeof(unsigned long long), st));
    k_prox_aset<<<plan.grid, AS_NT, plan.smem, st>>>(ctl, plan.chains.p, plan.batches.p, plan.nbatches, plan.cap, plan.maxpass,
                                                     beta, betaT, zr, ur, zc, uc, partials, plan.stats.p);
    if (plan.stats.p) {   // development aid: one line per launch
        unsigned long long h[8];
        FB_CUDA(cudaMemcpyAsync(h, plan.stats.p, sizeof h, cudaMemcpyDeviceToHost, st));
        FB_CUDA(cudaStreamSynchronize(st));
        if (h[5])
            std::fprintf(stderr, "[fb aset stats] batches %llu grid %d: cycles/batch load %.0f scan %.0f passes %.0f out %.0f | passes mean %.2f max %llu, fallbacks %llu\n",
                         h[5], plan.grid, (double)h[0] / h[5], (double)h[1] / h[5], (double)h[2] / h[5], (double)h[3] / h[5], (double)h[4] / h[5], h[6], h[7]);
    }
}

} // namespace fb
