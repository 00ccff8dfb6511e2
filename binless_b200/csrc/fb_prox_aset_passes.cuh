// fb_prox_aset_passes.cuh -- the shared-memory data structures and the per-word / per-point passes of the parallel
// active-set chain prox (k_prox_aset, fb_prox_aset.cu). Kept free of warp intrinsics so that the very same source is
// compiled for the host by tests/test_aset_passes_host.py (shim: tests/aset_host/shim.h), which runs the passes against
// the reference's tf_dp_weight under AddressSanitizer.
#pragma once
#include <cstdint>

namespace fb {

// One 32-point word of a batch (static: the kernel never searches the chain table).
struct alignas(16) AsWord {
    int64_t goff;  // cell of the direction's packed array that lane 0 of the word holds
    short nvalid;  // cells of the chain in this word (0..32)
    short cidx;    // chain of the word, relative to the batch
    short endbit;  // bit of the chain's last string point if it lies in this word, else -1
    short jw;      // word of the chain (0 = the word that starts with the chain's first point)
};

namespace aset {
constexpr int AS_NT = 256, AS_NW = AS_NT / 32;
constexpr int AS_RT = 256;                      // reciprocal table R[m] = 1/m
constexpr int AS_PB = 9, AS_PW = 32 * AS_PB;    // blocked prefix sums: AS_PB consecutive points per lane (odd: no bank conflicts)
constexpr int AS_SPLIT = 2 * AS_PW;             // chains longer than this are scanned by all warps of the CTA
#ifndef FB_AS_UN
#define FB_AS_UN 2
#endif
constexpr int AS_UN = FB_AS_UN;                 // words whose global loads are in flight together per warp
constexpr double AS_EPS = 0x1p-46;              // rounding margin of the tube test (see pass_points)
enum { C_BATCH = 0, C_COUNT = 4 };
constexpr int AS_TWN = 8;                       // 32-word blocks of the active-word masks (256 words)
enum { WF_ANY = 1, WF_ENTER = 2 };              // word flags: a piece of the word is to be tested / the piece entering it

// The shared memory of one CTA: every array of a batch of at most CAP string points, at compile-time offsets (the
// kernel is instantiated per capacity class; the host test uses a struct of pointers with the same member names).
template <int CAP>
struct SmT {
    static constexpr int NWD = CAP / 32;
    double S[CAP];              // prefix sums of v - c per string point (in D: the residual sums per chain)
    double SLP[CAP];            // at a knot: slope of the piece that starts there
    double R[AS_RT];            // 1 / m
    double Cc[NWD];             // per chain: c = v_0
    double SPAN[AS_NW];
    double WOFF[NWD];           // per word: string offset at the knot before the word
    AsWord WD[NWD];
    unsigned PM[CAP];           // at a knot: key of the worst point of the piece that starts there
    unsigned K[NWD], UP[NWD], FIXM[NWD], DIRTY[NWD], REM[NWD], CH[NWD];
    int Lc[NWD], Bc[NWD], Qc[NWD], ci[C_COUNT];
    unsigned TW[2][AS_TWN];     // words a pass has to look at (this pass / the next one), one bit per word
    unsigned short WPREV[NWD], WFL[NWD];
    unsigned char ENDB[NWD];    // per word: bit of the chain's last point, 32 = not in this word
};

__device__ __forceinline__ double recip_t(const double *R, int m) { return m < AS_RT ? R[m] : 1.0 / (double)m; }
// the fixed points of a word: the first point of a chain, its last point
__device__ __forceinline__ unsigned fix_bits(const AsWord &d) { return (d.jw == 0 ? 1u : 0u) | (d.endbit >= 0 ? 1u << d.endbit : 0u); }
// offset of the string from S at a knot: 0 at a chain end, +lam on the upper tube edge, -lam on the lower one
__device__ __forceinline__ double off_bits(unsigned fixw, unsigned upw, int bit, double lam) {
    return ((fixw >> bit) & 1u) ? 0.0 : (((upw >> bit) & 1u) ? lam : -lam);
}
template <class SM>
__device__ __forceinline__ double off_at(const SM &s, int p, double lam) { return off_bits(s.FIXM[p >> 5], s.UP[p >> 5], p & 31, lam); }

// P1 (one thread per word): slope of every piece that starts in word w and has a changed knot at either end or
// inside (all pieces when `all`); what the points of the word need to know about the piece entering it from the left
template <class SM>
__device__ __forceinline__ void pass_slopes(SM &s, const int w, const double lam, const bool all) {
    const unsigned kall = s.K[w], upw = s.UP[w], chw = s.CH[w];
    const AsWord d = s.WD[w];
    const unsigned fixw = s.FIXM[w];
    unsigned m = kall & ~(d.endbit >= 0 ? 1u << d.endbit : 0u), dbits = 0u;
    // the knot after the last one of this word and the changes on the way (first needed by the piece that crosses)
    int nb = -1; unsigned chmid = 0u;
    while (m) {
        const int bit = __ffs(m) - 1;
        m &= m - 1u;
        const int p = (w << 5) + bit;
        const unsigned above = bit == 31 ? 0u : (0xfffffffeu << bit);
        const unsigned hm = kall & above;
        int b; bool dirty = all;
        if (hm) {
            const int bb = __ffs(hm) - 1;
            b = (w << 5) + bb;
            if (!all) dirty = (chw & (0xffffffffu << bit) & (0xffffffffu >> (31 - bb))) != 0u;
        } else {
            if (nb < 0) {
                int x = w + 1;
                unsigned k = s.K[x];
                while (!k) { chmid |= s.CH[x]; k = s.K[++x]; }
                const int bb = __ffs(k) - 1;
                nb = (x << 5) + bb;
                chmid |= s.CH[x] & (0xffffffffu >> (31 - bb));
            }
            b = nb;
            if (!all) dirty = ((chw & (0xffffffffu << bit)) | chmid) != 0u;
        }
        if (dirty) {
            const double ya = s.S[p] + off_bits(fixw, upw, bit, lam);
            const double yb = s.S[b] + off_at(s, b, lam);
            s.SLP[p] = (yb - ya) * recip_t(&s.R[0], b - p);
            dbits |= 1u << bit;
        }
    }
    s.DIRTY[w] = dbits;
    // the piece entering the word: its first knot, the string offset there, whether it is to be tested
    unsigned fl = dbits ? WF_ANY : 0u;
    if (d.jw > 0) {      // also when the word starts with a knot: its bend test needs the knot before it
        int x = w - 1;
        unsigned k = s.K[x], chl = 0u;
        while (!k) { chl |= s.CH[x]; k = s.K[--x]; }
        const int pb = 31 - __clz(k);
        s.WPREV[w] = (unsigned short)((x << 5) + pb);
        s.WOFF[w] = off_bits(s.FIXM[x], s.UP[x], pb, lam);
        bool dirty = all;
        if (!all) {
            chl |= s.CH[x] & (0xffffffffu << pb);
            if (kall) chl |= chw & (0xffffffffu >> (31 - (__ffs(kall) - 1)));
            else {      // no knot in this word either: the piece ends where the search above ended (or search now)
                if (nb < 0) {
                    int y = w + 1;
                    unsigned k2 = s.K[y];
                    chmid = 0u;
                    while (!k2) { chmid |= s.CH[y]; k2 = s.K[++y]; }
                    chmid |= s.CH[y] & (0xffffffffu >> (31 - (__ffs(k2) - 1)));
                }
                chl |= chw | chmid;
            }
            dirty = chl != 0u;
        }
        if (dirty) fl |= WF_ANY | WF_ENTER;
    }
    s.WFL[w] = (unsigned short)fl;
}

// P2a (one thread per word): the knots of word w next to a piece whose slope changed must bend towards their tube edge
template <class SM>
__device__ __forceinline__ void pass_bends(SM &s, const int w) {
    const unsigned fl = s.WFL[w];
    s.CH[w] = 0u;
    if (!fl) { s.REM[w] = 0u; return; }
    const unsigned kall = s.K[w], upw = s.UP[w], dw = s.DIRTY[w];
    unsigned m = kall & ~s.FIXM[w], rem = 0u;
    while (m) {
        const int bit = __ffs(m) - 1;
        m &= m - 1u;
        const unsigned lm = kall & ((1u << bit) - 1u);
        const int a = lm ? (w << 5) + 31 - __clz(lm) : (int)s.WPREV[w];
        const unsigned da = lm ? (dw >> (a & 31)) & 1u : (fl >> 1) & 1u;
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
template <class SM>
__device__ __forceinline__ void pass_points(SM &s, const int g, const double lam, const int lane, const unsigned lt) {
    const unsigned fl = s.WFL[g];
    if (!fl) return;
    const unsigned kw = s.K[g], dw = s.DIRTY[g], eb = s.ENDB[g];
    // straight-line: lanes without a point to test (knots, points beyond the chain end, clean pieces) compute on point 0
    const unsigned lm = kw & lt;
    const int bit = 31 - __clz(lm | 1u);
    const bool inw = lm != 0u;
    const int pa_ = inw ? (g << 5) + bit : (int)s.WPREV[g];
    const double oa = inw ? off_bits(s.FIXM[g], s.UP[g], bit, lam) : s.WOFF[g];
    const bool dirty = inw ? ((dw >> bit) & 1u) != 0u : (fl & WF_ENTER) != 0u;
    const bool act = dirty && lane < (int)eb && !((kw >> lane) & 1u);
    const int pa = act ? pa_ : 0, p = act ? (g << 5) + lane : 0;
    const double sp = s.S[p];
    const double y = (s.S[pa] + oa) + s.SLP[pa] * (double)(p - pa);
    const double dd = y - sp;
    const double e = fabs(dd) - lam;
    if (act && e > 0.0 && e > AS_EPS * (fabs(sp) + fabs(y) + lam)) {
        // key: the violation as a float (truncated: any point outside is a valid new knot), the side, the offset
        const unsigned bits = __float_as_uint((float)e);
        const unsigned key = (bits & 0xffffc000u) | (dd > 0.0 ? 0x2000u : 0u) | (unsigned)(p - pa);
        atomicMax(&s.PM[pa], key);
    }
}

// last knot at a position <= x / first knot at a position > x (a chain starts and ends with a knot that never goes;
// knots that other threads add or remove meanwhile only move the answer to another knot of the old or the new set)
template <class SM>
__device__ __forceinline__ int knot_le(const SM &s, const int x) {
    int w = x >> 5;
    unsigned m = s.K[w] & (0xffffffffu >> (31 - (x & 31)));
    while (!m) m = s.K[--w];
    return (w << 5) + 31 - __clz(m);
}
template <class SM>
__device__ __forceinline__ int knot_gt(const SM &s, const int x) {
    int w = x >> 5;
    unsigned m = (x & 31) == 31 ? 0u : (s.K[w] & (0xfffffffeu << (x & 31)));
    while (!m) m = s.K[++w];
    return (w << 5) + __ffs(m) - 1;
}

// P3 (one thread per word): knots that bend the wrong way go, the worst point of every tested piece that starts in
// word w becomes a knot. The words from the knot before the first change to the knot after the last one are what the
// next pass has to look at (TW[nxt]): every piece with a changed end lies inside the union of these spans, whatever
// the order in which the threads change K.
template <class SM>
__device__ __forceinline__ void pass_apply(SM &s, const int w, const int nxt) {
    const unsigned rem = s.REM[w];
    int lo = 1 << 30, hi = -1;
    bool lo_rem = false;
    if (rem) { lo = (w << 5) + __ffs(rem) - 1; hi = (w << 5) + 31 - __clz(rem); lo_rem = true; }
    unsigned m = s.DIRTY[w];
    while (m) {
        const int bit = __ffs(m) - 1;
        m &= m - 1u;
        const int p = (w << 5) + bit;
        const unsigned key = s.PM[p];
        if (key) {
            s.PM[p] = 0u;
            const int q = p + (int)(key & 0x1fffu);
            const unsigned qb = 1u << (q & 31);
            if (key & 0x2000u) atomicOr(&s.UP[q >> 5], qb); else atomicAnd(&s.UP[q >> 5], ~qb);
            atomicOr(&s.K[q >> 5], qb); atomicOr(&s.CH[q >> 5], qb);
            if (p < lo) { lo = p; lo_rem = false; }
            if (q > hi) hi = q;
        }
    }
    if (hi < 0) return;
    const int a = lo_rem ? knot_le(s, lo - 1) : lo;
    const int b = knot_gt(s, hi);
    if (rem) { atomicAnd(&s.K[w], ~rem); atomicOr(&s.CH[w], rem); }
    const int wa = a >> 5, wb = b >> 5;
    for (int blk = wa >> 5; blk <= (wb >> 5); ++blk) {
        const int l = max(wa, blk << 5) & 31, h = min(wb, (blk << 5) + 31) & 31;
        atomicOr(&s.TW[nxt][blk], (0xffffffffu << l) & (0xffffffffu >> (31 - h)));
    }
}

// Fallback (one lane per chain): the sequential taut string from the chain start; rewrites the knot masks of the chain
template <class SM>
__device__ __forceinline__ void chain_sequential(SM &s, const int base, const int L, const double lam) {
    const double INF = __longlong_as_double(0x7ff0000000000000LL);
    for (int w = base >> 5; w <= (base + L) >> 5; ++w) { s.K[w] = s.FIXM[w]; s.UP[w] = 0u; s.REM[w] = 0u; s.CH[w] = 0u; }
    const double *S = &s.S[base];
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


} // namespace aset
} // namespace fb
