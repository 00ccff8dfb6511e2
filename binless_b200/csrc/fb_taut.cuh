// fb_taut.cuh -- warp-cooperative taut-string search of the team chain-prox kernel k_prox (fb_gfl.cu: two warps per
// chain from both ends).
// For the chain v_0..v_{L-1} and tube radius lam the 1-D fused-lasso minimiser (tf_dp_weight, gfl_tf.c:184-321) is the
// derivative of the taut string through the tube S[k] -+ lam around the prefix sums S.
#pragma once
#include "fb_util.cuh"

namespace fb {
namespace taut {

constexpr unsigned FULL = 0xffffffffu;

__device__ __forceinline__ double scan_sum_incl(double v, int lane) {
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const double t = __shfl_up_sync(FULL, v, o); if (lane >= o) v += t; }
    return v;
}

#ifndef FB_PBR
#define FB_PBR 5
#endif
constexpr int PBR = FB_PBR;     // 32 * PBR candidates per round (the average segment is ~45 cells); odd: conflict-free
constexpr int PB = 7, PW = 32 * PB;   // run length of the blocked prefix-sum pass

__device__ __forceinline__ double dmin_(double a, double b) { return b < a ? b : a; }
__device__ __forceinline__ double dmax_(double a, double b) { return b > a ? b : a; }
__device__ __forceinline__ double scan_min_fast(double v, int lane) {
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const double t = __shfl_up_sync(FULL, v, o); if (lane >= o) v = dmin_(v, t); }
    return v;
}
__device__ __forceinline__ double scan_max_fast(double v, int lane) {
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const double t = __shfl_up_sync(FULL, v, o); if (lane >= o) v = dmax_(v, t); }
    return v;
}

// latest b (< limit) with x[b] == v, or -1
__device__ __forceinline__ int latest_equal(const double (&x)[PBR], double v, int limit) {
    constexpr int B = PBR;
    int r = -1;
#pragma unroll
    for (int b = 0; b < B; ++b) if (b < limit && x[b] == v) r = b;
    return r;
}

// The chain as one warp sees it: frame coordinate k in [0, L], value sgn * S[idx(k)] + off.
struct Frame {
    const double *S;   // prefix sums of the chain, S[0..L]
    int L, dir;        // dir 0: left warp (k -> S[k]); dir 1: right warp (k -> S[L] - S[L-k])
    double SL;
    __device__ __forceinline__ double at(int k) const { return dir ? SL - S[L - k] : S[k]; }
};

// Search state of one segment (frame coordinates)
struct Seg {
    int k0; double s0;          // start of the segment being searched: the string passes through (k0, s0)
    int p;                      // candidates up to p have been examined
    double H, Lw; int kH, kL;   // running min of hi / max of lo and where they were attained (latest on ties)
};

// Returns true when a tube violation was found: the segment then ends at kend with value c + slope, leaving the
// upper (isup) or lower tube edge. Otherwise the carried bounds are updated. One instantiation only, loops that
// are not on the critical path are not unrolled: the team kernel is instruction-cache bound otherwise.
constexpr int B = PBR;
// 1/m: from the shared-memory table when m < rl, else the same correctly rounded quotient computed in place
// (LONG: beyond ~28 000 cells the table no longer fits in shared memory and covers the first rl distances only)
template <bool LONG>
__device__ __forceinline__ double recip(const double *__restrict__ R, const int rl, const int m) {
    if (LONG) return m < rl ? R[m] : 1.0 / (double)m;
    return R[m];
}
template <bool LONG>
__device__ __forceinline__ bool taut_round(const Frame &f, const double *__restrict__ R, const int rl, const double lam, Seg &g,
                                           const int kEnd, const double aEnd, const int lane,
                                           int &isup, double &slope, int &kend) {
    const double INF = __longlong_as_double(0x7ff0000000000000LL);
    const int kb = g.p + 1 + lane * B;
    const double sgn = f.dir ? -1.0 : 1.0;
    const double off = f.dir ? f.SL - g.s0 : -g.s0;
    double hi[B], lo[B];
    double hmin = INF, lmax = -INF;
    unsigned m = 0;
    if (g.p + 32 * B < kEnd) { // every candidate is an interior point of the chain
        const int m0 = kb - g.k0;
        double sv[B];
        if (f.dir) {
            const double *__restrict__ Sq = f.S + (f.L - kb);
#pragma unroll
            for (int b = 0; b < B; ++b) sv[b] = Sq[-b];
        } else {
            const double *__restrict__ Sq = f.S + kb;
#pragma unroll
            for (int b = 0; b < B; ++b) sv[b] = Sq[b];
        }
#pragma unroll
        for (int b = 0; b < B; ++b) {
            const double base = __fma_rn(sgn, sv[b], off);
            const double ri = recip<LONG>(R, rl, m0 + b);
            const double h = (base + lam) * ri, l = (base - lam) * ri;
            hi[b] = h; lo[b] = l;
            if ((l > hmin) || (h < lmax)) m |= 1u << b;
            hmin = dmin_(hmin, h); lmax = dmax_(lmax, l);
        }
    } else {                   // the fixed end point kEnd is among the candidates; those beyond it do not exist
#pragma unroll
        for (int b = 0; b < B; ++b) {
            const int k = kb + b;
            const int kc = k < kEnd ? k : kEnd;
            const double base = __fma_rn(sgn, f.S[f.dir ? f.L - kc : kc], off);
            const double ri = recip<LONG>(R, rl, kc - g.k0);
            const bool fixed = (k == kEnd);      // a point of the string, not a tube
            double h = (base + (fixed ? aEnd : lam)) * ri;
            double l = (base - (fixed ? -aEnd : lam)) * ri;
            if (k > kEnd) { h = INF; l = -INF; }
            hi[b] = h; lo[b] = l;
            if ((l > hmin) || (h < lmax)) m |= 1u << b;
            hmin = dmin_(hmin, h); lmax = dmax_(lmax, l);
        }
    }
    const double Hs = scan_min_fast(hmin, lane), Ls = scan_max_fast(lmax, lane);
    double Hx = __shfl_up_sync(FULL, Hs, 1), Lx = __shfl_up_sync(FULL, Ls, 1);
    if (lane == 0) { Hx = INF; Lx = -INF; }
    Hx = dmin_(Hx, g.H); Lx = dmax_(Lx, g.Lw);
#pragma unroll
    for (int b = 0; b < B; ++b)
        if ((lo[b] > Hx) || (hi[b] < Lx)) m |= 1u << b;
    const unsigned mask = __ballot_sync(FULL, m != 0u);
    if (mask == 0u) {
        const double nH = __shfl_sync(FULL, Hs, 31), nL = __shfl_sync(FULL, Ls, 31);
        if (nH <= g.H) {
            g.H = nH;
            const int src = 31 - __clz(__ballot_sync(FULL, hmin == nH));
            g.kH = __shfl_sync(FULL, kb + latest_equal(hi, nH, B), src);
        }
        if (nL >= g.Lw) {
            g.Lw = nL;
            const int src = 31 - __clz(__ballot_sync(FULL, lmax == nL));
            g.kL = __shfl_sync(FULL, kb + latest_equal(lo, nL, B), src);
        }
        return false;
    }
    const int fb = __ffs(mask) - 1;
    // lane fb: first violating candidate, the bound it violates and where that bound was attained
    int up = 0, ke = -1;
    double sl = 0.0;
    if (lane == fb) {
        const int bv = __ffs(m) - 1;
        double pmH = INF, pmL = -INF, lv = 0.0;
#pragma unroll
        for (int b = 0; b < B; ++b) {
            if (b < bv) { pmH = dmin_(pmH, hi[b]); pmL = dmax_(pmL, lo[b]); }
            if (b == bv) lv = lo[b];
        }
        const double bH = dmin_(Hx, pmH), bL = dmax_(Lx, pmL);
        up = lv > bH;
        if (up) { sl = bH; if (pmH <= Hx) ke = kb + latest_equal(hi, sl, bv); }
        else    { sl = bL; if (pmL >= Lx) ke = kb + latest_equal(lo, sl, bv); }
    }
    up = __shfl_sync(FULL, up, fb);
    sl = __shfl_sync(FULL, sl, fb);
    ke = __shfl_sync(FULL, ke, fb);
    if (ke < 0) { // attained in an earlier lane of this round, else in an earlier round
        const unsigned eq = __ballot_sync(FULL, lane < fb && (up ? hmin : lmax) == sl);
        if (eq) {
            const int src = 31 - __clz(eq);
            ke = __shfl_sync(FULL, kb + (up ? latest_equal(hi, sl, B) : latest_equal(lo, sl, B)), src);
        } else ke = up ? g.kH : g.kL;
    }
    isup = up; slope = sl; kend = ke;
    return true;
}

__device__ __forceinline__ void seg_reset(Seg &g) {
    const double INF = __longlong_as_double(0x7ff0000000000000LL);
    g.p = g.k0; g.H = INF; g.Lw = -INF; g.kH = g.k0; g.kL = g.k0;
}

} // namespace taut
} // namespace fb
