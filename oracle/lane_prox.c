/* lane_prox.c -- TEST / DESIGN INFRASTRUCTURE (not on the product path).
 *
 * Scalar C model of the "split and verify" chain prox of k_prox_lane (binless_b200/csrc/fb_gfl.cu), with the data
 * structures the kernel uses: prefix sums S[0..L], the knots of the string as two L-bit masks (position, side), and a
 * list of SPLITS -- hypothesised knots taken from the previous ADMM iterate, at most one per window of W cells (the one
 * with the largest jump).
 *
 * Solves   min_x 1/2 sum_k (x_k - v_k)^2 + lam sum_k |x_{k+1} - x_k|   (what tf_dp_weight(n, v, w = 1, lam),
 * /root/reference/binless/src/gfl_tf.c:184-321, returns; update_z calls it with w = alpha/2, gfl_graph_fl.c:267-304).
 *
 * The minimiser is the derivative of the taut string through the tube S[k] -+ lam from (0, 0) to (L, S[L]). A point
 * where the string is KNOWN to touch the tube (position k, side s: height S[k] + s lam) cuts the problem into two
 * independent ones with fixed end points. So:
 *   1. every range between two consecutive splits is solved on its own, exactly, by ONE lane (sequential taut string
 *      with fixed end points; the lanes of a warp work on 32 ranges at once);
 *   2. a split is CORRECT iff the string bends towards the tube edge it touches there (slope right - slope left has
 *      the sign of s): with that, the glued string satisfies the KKT conditions of a strictly convex problem, so it IS
 *      the solution;
 *   3. a split that fails is removed, the two ranges around it are merged and solved again; repeat until every
 *      remaining split passes (each round removes at least one split, so this ends; with no split left the whole
 *      chain is one exact sequential solve).
 * The result does not depend on the hypothesis; only the amount of sequential work does.
 * tests/test_lane_prox_model.py checks the model against the reference's tf_dp_weight for arbitrary hypotheses. */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

typedef struct {
    long long steps_total;   /* candidates examined by all lanes (lane steps) */
    long long warp_steps;    /* sum over warps (32 consecutive ranges of a pass) of the longest range: lock-step cost */
    long long solves;        /* range solves */
    long long splits, failed;/* splits tried / removed */
    long long cert_dirty;    /* cells of the pieces whose certificate failed (W = 0 mode) */
    int passes, max_range;
} LaneStats;

typedef struct { uint64_t *knot, *up; int L, words; const double *S; double lam; } Chain;

static int get(const uint64_t *m, int k) { return (int)((m[k >> 6] >> (k & 63)) & 1u); }
static void set(uint64_t *m, int k) { m[k >> 6] |= (uint64_t)1 << (k & 63); }
static void clr(uint64_t *m, int k) { m[k >> 6] &= ~((uint64_t)1 << (k & 63)); }

static int prev_knot(const Chain *c, int k) {   /* nearest knot at or below k, 0 = chain start */
    for (int w = k >> 6; w >= 0; --w) {
        uint64_t m = c->knot[w];
        if (w == (k >> 6)) m &= (k & 63) == 63 ? ~(uint64_t)0 : (((uint64_t)1 << ((k & 63) + 1)) - 1);
        if (m) return (w << 6) + 63 - __builtin_clzll(m);
    }
    return 0;
}
static int next_knot(const Chain *c, int k) {   /* nearest knot strictly above k, L = chain end */
    for (int w = (k + 1) >> 6; w < c->words; ++w) {
        uint64_t m = c->knot[w];
        if (w == ((k + 1) >> 6)) m &= ~(uint64_t)0 << ((k + 1) & 63);
        if (m) { const int r = (w << 6) + __builtin_ctzll(m); return r < c->L ? r : c->L; }
    }
    return c->L;
}
static double height(const Chain *c, int k) {
    if (k <= 0 || k >= c->L) return c->S[k <= 0 ? 0 : c->L];
    return c->S[k] + (get(c->up, k) ? c->lam : -c->lam);
}

/* one lane: exact taut string between the fixed points a and b; sets the knots found strictly inside; returns the
 * number of candidates examined (the lane's step count) */
static int solve_range(Chain *c, int a, int b) {
    int k0 = a, steps = 0;
    double y0 = height(c, a);
    const double yb = height(c, b);
    for (;;) {
        double hi = INFINITY, lo = -INFINITY;
        int khi = k0, klo = k0, found = 0, fk = 0, fs = 0;
        for (int k = k0 + 1; k <= b; ++k) {
            double h, l;
            const double m = (double)(k - k0);
            ++steps;
            if (k == b) h = l = (yb - y0) / m;
            else { h = (c->S[k] + c->lam - y0) / m; l = (c->S[k] - c->lam - y0) / m; }
            if (l > hi) { found = 1; fk = khi; fs = 1; break; }
            if (h < lo) { found = 1; fk = klo; fs = 0; break; }
            if (h <= hi) { hi = h; khi = k; }
            if (l >= lo) { lo = l; klo = k; }
        }
        if (!found) return steps;
        set(c->knot, fk);
        if (fs) set(c->up, fk); else clr(c->up, fk);
        k0 = fk; y0 = height(c, fk);
    }
}

/* does the string bend towards the tube edge it touches at the knot k? */
static int bend_ok(const Chain *c, int k) {
    const int a = prev_knot(c, k - 1), b = next_knot(c, k);
    const double yk = height(c, k);
    const double bend = (height(c, b) - yk) / (b - k) - (yk - height(c, a)) / (k - a);
    return (get(c->up, k) ? bend : -bend) > 0;
}

/* out[L]; W = window length of the split selection (cells); returns the number of passes */
int lane_prox(int L, const double *v, double lam, const double *z_prev, double *out, int W, LaneStats *st) {
    LaneStats local; if (!st) { memset(&local, 0, sizeof local); st = &local; }
    if (L == 1 || lam == 0) { memcpy(out, v, sizeof(double) * (size_t)L); return 0; }
    const int certify = W <= 0;      /* W = 0: every hypothesised knot is a split and the certificate goes first */
    if (W < 1) W = 1;
    const int words = (L >> 6) + 1;
    double *S = (double *)malloc(sizeof(double) * (size_t)(L + 1));
    uint64_t *mask = (uint64_t *)calloc((size_t)2 * words, sizeof(uint64_t));
    Chain c = {mask, mask + words, L, words, S, lam};
    const double c0 = v[0];
    S[0] = 0;
    for (int k = 0; k < L; ++k) S[k + 1] = S[k] + (v[k] - c0);
    /* splits: per window of W cells the hypothesised knot with the largest jump (ties: the last one) */
    int *sp = (int *)malloc(sizeof(int) * (size_t)(L / W + 4));
    unsigned char *dirty = (unsigned char *)malloc((size_t)(L / W + 4));
    int ns = 0;
    sp[ns++] = 0;
    for (int w0 = 0; w0 < L; w0 += W) {
        int best = -1; double bj = 0;
        for (int k = (w0 > 1 ? w0 : 1); k < w0 + W && k < L; ++k) {
            const double j = fabs(z_prev[k] - z_prev[k - 1]);
            if (j > 0 && j >= bj) { bj = j; best = k; }
        }
        if (best > 0) { sp[ns++] = best; set(c.knot, best); if (z_prev[best] > z_prev[best - 1]) set(c.up, best); }
    }
    sp[ns] = L;     /* ranges t = 0..ns-1: [sp[t], sp[t+1]] */
    st->splits += ns - 1;
    for (int t = 0; t < ns; ++t) dirty[t] = 1;
    int passes = 0;
    if (certify) {
        /* certificate first: every piece of the hypothesis is tested cell by cell (is the straight line inside the tube?)
         * and every knot for its bend; only the pieces that fail are solved */
        unsigned char *fail = (unsigned char *)calloc((size_t)ns + 1, 1);
        int nfail = 0, ndirty = 0;
        for (int t = 0; t < ns; ++t) {
            const int a = sp[t], b = sp[t + 1];
            const double ya = height(&c, a), yb = height(&c, b);
            const double slope = (yb - ya) / (b - a);
            dirty[t] = 0;
            for (int k = a + 1; k < b; ++k) {
                const double y = ya + slope * (k - a);
                if (y > S[k] + lam || y < S[k] - lam) { dirty[t] = 1; break; }
            }
        }
        for (int t = 1; t < ns; ++t) if (!bend_ok(&c, sp[t])) { fail[t] = 1; ++nfail; }
        st->failed += nfail;
        int m = 1;
        for (int t = 1; t < ns; ++t) {
            if (fail[t]) { clr(c.knot, sp[t]); dirty[m - 1] = 1; if (dirty[t]) dirty[m - 1] = 1; }
            else { sp[m] = sp[t]; dirty[m] = dirty[t]; ++m; }
        }
        sp[m] = L; ns = m;
        free(fail);
        for (int t = 0; t < ns; ++t) if (dirty[t]) { ++ndirty; st->cert_dirty += sp[t + 1] - sp[t]; }
        if (!ndirty) goto done;
    }
    for (;;) {
        ++passes;
        /* solve the dirty ranges: lane i of the pass takes the i-th dirty range */
        int lane = 0, wmax = 0;
        for (int t = 0; t < ns; ++t) {
            if (!dirty[t]) continue;
            const int a = sp[t], b = sp[t + 1];
            for (int k = a + 1; k < b; ++k) if (get(c.knot, k)) clr(c.knot, k);
            const int steps = solve_range(&c, a, b);
            st->steps_total += steps; ++st->solves;
            if (b - a > st->max_range) st->max_range = b - a;
            if (steps > wmax) wmax = steps;
            if (++lane == 32) { st->warp_steps += wmax; lane = 0; wmax = 0; }
        }
        st->warp_steps += wmax;
        /* verify the splits next to a range that was just solved (all judged on the same string) */
        int nfail = 0;
        unsigned char *fail = (unsigned char *)calloc((size_t)ns + 1, 1);
        for (int t = 1; t < ns; ++t)
            if ((dirty[t - 1] || dirty[t]) && !bend_ok(&c, sp[t])) { fail[t] = 1; ++nfail; }
        if (!nfail) { free(fail); break; }
        st->failed += nfail;
        /* drop the failed splits; a range that swallowed one is dirty */
        int m = 1;
        dirty[0] = 0;
        for (int t = 1; t < ns; ++t) {
            if (fail[t]) { clr(c.knot, sp[t]); dirty[m - 1] = 1; }
            else { sp[m] = sp[t]; dirty[m] = 0; ++m; }
        }
        sp[m] = L; ns = m;
        free(fail);
    }
done:
    st->passes += passes;
    for (int a = 0; a < L;) {
        const int b = next_knot(&c, a);
        const double x = c0 + (height(&c, b) - height(&c, a)) / (b - a);
        for (int k = a; k < b; ++k) out[k] = x;
        a = b;
    }
    free(S); free(mask); free(sp); free(dirty);
    return passes;
}
