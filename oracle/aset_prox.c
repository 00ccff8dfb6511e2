/* aset_prox.c -- TEST / DESIGN INFRASTRUCTURE (not on the product path).
 *
 * Scalar C model of the parallel active-set chain prox of k_prox_aset (binless_b200/csrc/fb_prox_aset.cu), with the data
 * structures the kernel uses: prefix sums S[0..L], the knots of the string as two bit masks (position, side).
 *
 * Solves   min_x 1/2 sum_k (x_k - v_k)^2 + lam sum_k |x_{k+1} - x_k|   (what tf_dp_weight(n, v, w = 1, lam),
 * /root/reference/binless/src/gfl_tf.c:184-321, returns; update_z calls it with w = alpha/2, gfl_graph_fl.c:267-304).
 *
 * The minimiser is the derivative of the taut string through the tube S[k] -+ lam around the prefix sums S of v, from
 * (0, 0) to (L, S[L]): the polygon through its KNOTS (points where it touches the tube: position k, side s, height
 * S[k] + s lam). A knot set is the solution iff (KKT of a strictly convex problem)
 *   (i)  the polygon stays inside the tube at every point between two knots, and
 *   (ii) at every knot it bends towards the tube edge it touches (slope right - slope left has the sign of s).
 * Both tests are independent per point / per knot, and so are the repairs:
 *   (i)  fails on a piece  -> the point of the piece that is furthest outside becomes a knot (on the edge it crossed);
 *   (ii) fails at a knot   -> the knot is removed.
 * All pieces and knots are treated at once; pieces whose two end knots did not change are not looked at again. The
 * starting knot set is the previous ADMM iterate's (its jumps); after a few passes nothing changes any more and the
 * polygon IS the solution, whatever the starting set was (only the number of passes depends on it). Should the passes
 * not settle within `maxpass` the caller solves the chain with the sequential taut string instead (never seen on the
 * iterates of real solves; rounding-level ties could in principle make a knot flip for ever).
 * tests/test_aset_prox_model.py checks the model against the reference's tf_dp_weight for arbitrary hypotheses. */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

typedef struct {
    long long passes;          /* passes over all chains (the last one of a chain only confirms) */
    long long point_tests;     /* points tested against the tube, all passes */
    long long knot_tests;      /* bend tests */
    long long added, removed;
    long long chains, fallbacks;
    int max_passes;
} AsetStats;

#define ASET_EPS 0x1p-46
typedef struct { uint32_t *K, *UP, *DIRTY; int L, words; const double *S; double lam; } Ch;

static int getb(const uint32_t *m, int k) { return (int)((m[k >> 5] >> (k & 31)) & 1u); }
static void setb(uint32_t *m, int k) { m[k >> 5] |= 1u << (k & 31); }
static void clrb(uint32_t *m, int k) { m[k >> 5] &= ~(1u << (k & 31)); }
static int prev_le(const uint32_t *m, int k) {              /* last set bit at a position <= k (bit 0 is always set) */
    int w = k >> 5;
    uint32_t x = m[w] & (0xffffffffu >> (31 - (k & 31)));
    while (!x) x = m[--w];
    return (w << 5) + 31 - __builtin_clz(x);
}
static int next_gt(const uint32_t *m, int k) {              /* first set bit at a position > k (bit L is always set) */
    int w = k >> 5;
    uint32_t x = (k & 31) == 31 ? 0u : (m[w] & (0xfffffffeu << (k & 31)));
    while (!x) x = m[++w];
    return (w << 5) + __builtin_ctz(x);
}
static double height(const Ch *c, int k) {
    if (k == 0 || k == c->L) return c->S[k];
    return c->S[k] + (getb(c->UP, k) ? c->lam : -c->lam);
}

/* out[L]; returns the number of passes, -1 when maxpass passes did not settle */
int aset_prox(int L, const double *v, double lam, const double *z_prev, double *out, int maxpass, AsetStats *st) {
    AsetStats local; if (!st) { memset(&local, 0, sizeof local); st = &local; }
    ++st->chains;
    if (L <= 0) return 0;
    if (L == 1 || lam == 0.0) { memcpy(out, v, sizeof(double) * (size_t)L); return 0; }    /* gfl_tf.c:210-215 */
    const int words = (L >> 5) + 1;
    double *S = (double *)malloc(sizeof(double) * (size_t)(L + 1));
    uint32_t *mask = (uint32_t *)calloc((size_t)5 * words, sizeof(uint32_t));
    uint32_t *REM = mask + 3 * words, *ADD = mask + 4 * words;
    uint32_t *pm = (uint32_t *)calloc((size_t)L + 1, sizeof(uint32_t));    /* per piece start: key of the worst point */
    Ch c = {mask, mask + words, mask + 2 * words, L, words, S, lam};
    const double c0 = v[0];
    S[0] = 0.0;
    for (int k = 0; k < L; ++k) S[k + 1] = S[k] + (v[k] - c0);
    setb(c.K, 0); setb(c.K, L);
    for (int k = 1; k < L; ++k)
        if (z_prev[k] != z_prev[k - 1]) { setb(c.K, k); if (z_prev[k] > z_prev[k - 1]) setb(c.UP, k); }
    for (int k = 0; k < L; ++k) if (getb(c.K, k)) setb(c.DIRTY, k);       /* a piece is named by its left knot */
    int passes = 0, settled = 0;
    while (passes < maxpass) {
        ++passes; ++st->passes;
        int changed = 0;
        /* (ii) bends at the knots next to a dirty piece */
        for (int k = 1; k < L; ++k) {
            if (!getb(c.K, k)) continue;
            const int a = prev_le(c.K, k - 1), b = next_gt(c.K, k);
            if (!getb(c.DIRTY, a) && !getb(c.DIRTY, k)) continue;
            ++st->knot_tests;
            const double yk = height(&c, k);
            const double bend = (height(&c, b) - yk) / (double)(b - k) - (yk - height(&c, a)) / (double)(k - a);
            if (!(getb(c.UP, k) ? bend > 0.0 : bend < 0.0)) { setb(REM, k); changed = 1; }
        }
        /* (i) points of the dirty pieces against the tube */
        for (int a = 0; a < L;) {
            const int b = next_gt(c.K, a);
            if (getb(c.DIRTY, a)) {
                const double ya = height(&c, a), slope = (height(&c, b) - ya) / (double)(b - a);
                for (int k = a + 1; k < b; ++k) {
                    ++st->point_tests;
                    const double y = ya + slope * (double)(k - a);
                    const double eu = y - (S[k] + lam), ed = (S[k] - lam) - y;
                    const double e = eu > ed ? eu : ed;
                    /* a point counts as outside only beyond the rounding error of this very expression: without the margin
                     * a knot whose bend is zero up to rounding would be removed and added again for ever */
                    if (e > ASET_EPS * (fabs(S[k]) + fabs(y) + lam)) {
                        /* key: the violation as a float, truncated, the side in bit 12, the point below */
                        const float ef = (float)e;
                        uint32_t bits; memcpy(&bits, &ef, 4);
                        if (bits < 0x00800000u) bits = 0x00800000u;     /* denormal / zero after the conversion */
                        const uint32_t key = (bits & 0xffffe000u) | (eu > ed ? 0x1000u : 0u) | (uint32_t)(k - a > 4095 ? 4095 : k - a);
                        if (key > pm[a]) pm[a] = key;
                    }
                }
            }
            a = b;
        }
        /* update */
        memset(c.DIRTY, 0, sizeof(uint32_t) * (size_t)words);
        for (int a = 0; a < L;) {
            const int b = next_gt(c.K, a);
            if (pm[a]) {
                int d = (int)(pm[a] & 0xfffu);
                if (d == 4095) {   /* the offset did not fit: take the first point outside (any point outside is a valid move) */
                    const double ya = height(&c, a), slope = (height(&c, b) - ya) / (double)(b - a);
                    for (d = 1; d < b - a; ++d) {
                        const double y = ya + slope * (double)d, t = ASET_EPS * (fabs(S[a + d]) + fabs(y) + lam);
                        if (y - (S[a + d] + lam) > t || (S[a + d] - lam) - y > t) break;
                    }
                    const double y = ya + slope * (double)d;
                    pm[a] = (pm[a] & ~0x1000u) | ((y - (S[a + d] + lam) > (S[a + d] - lam) - y) ? 0x1000u : 0u);
                }
                setb(ADD, a + d);
                if (pm[a] & 0x1000u) setb(c.UP, a + d); else clrb(c.UP, a + d);
                pm[a] = 0; changed = 1; ++st->added;
            }
            a = b;
        }
        if (!changed) { settled = 1; break; }
        /* dirty: the pieces that got a new end point */
        for (int k = 1; k < L; ++k) {
            if (getb(REM, k)) { clrb(c.K, k); ++st->removed; }
        }
        for (int k = 1; k < L; ++k) {
            if (getb(REM, k)) { setb(c.DIRTY, prev_le(c.K, k)); }           /* the merged piece */
            if (getb(ADD, k)) { setb(c.K, k); }
        }
        for (int k = 1; k < L; ++k)
            if (getb(ADD, k)) { setb(c.DIRTY, prev_le(c.K, k - 1)); setb(c.DIRTY, k); }
        memset(REM, 0, sizeof(uint32_t) * (size_t)words); memset(ADD, 0, sizeof(uint32_t) * (size_t)words);
    }
    if (passes > st->max_passes) st->max_passes = passes;
    if (settled) {
        for (int a = 0; a < L;) {
            const int b = next_gt(c.K, a);
            const double x = c0 + (height(&c, b) - height(&c, a)) / (double)(b - a);
            for (int k = a; k < b; ++k) out[k] = x;
            a = b;
        }
    } else ++st->fallbacks;
    free(S); free(mask); free(pm);
    return settled ? passes : -1;
}
