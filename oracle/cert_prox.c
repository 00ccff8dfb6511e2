/* cert_prox.c -- TEST / DESIGN INFRASTRUCTURE (not on the product path).
 *
 * Scalar C model of the certificate-first chain prox planned for the next round, written
 * with the data structures the CUDA kernel will use: the prefix sums S[0..L], the hypothesised knots as two L-bit masks
 * (position, side), a per-cell certificate that finds its piece by bit scans, a "bad piece" mask keyed by the piece's
 * left knot, maximal runs of bad pieces searched exactly between verified knots as fixed end points.
 * Solves   min_x 1/2 sum_k (x_k - v_k)^2 + lam sum_k |x_{k+1} - x_k|   (what tf_dp_weight(n, v, w = 1, lam),
 * /root/reference/binless/src/gfl_tf.c:184-321, returns; update_z calls it with w = alpha/2, gfl_graph_fl.c:267-304).
 * tests/test_certificate_prototype.py checks it against the reference's tf_dp_weight for arbitrary hypotheses.
 *
 * A point k (1 <= k <= L-1) of the string lies between cells k-1 and k; side +1: the string touches S[k] + lam there and
 * bends upwards (x increases), side -1: S[k] - lam, bends downwards. */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

typedef struct { uint64_t *knot, *up, *bad, *drop; int L, words; const double *S; double lam; } Chain;

static int get(const uint64_t *m, int k) { return (int)((m[k >> 6] >> (k & 63)) & 1u); }
static void set(uint64_t *m, int k) { m[k >> 6] |= (uint64_t)1 << (k & 63); }
static void clr(uint64_t *m, int k) { m[k >> 6] &= ~((uint64_t)1 << (k & 63)); }

/* nearest knot at or below k (0 = chain start) / strictly above k (L = chain end): word scans, as lanes will do */
static int prev_knot(const Chain *c, int k) {
    for (int w = k >> 6; w >= 0; --w) {
        uint64_t m = c->knot[w];
        if (w == (k >> 6)) m &= (k & 63) == 63 ? ~(uint64_t)0 : (((uint64_t)1 << ((k & 63) + 1)) - 1);
        if (m) return (w << 6) + 63 - __builtin_clzll(m);
    }
    return 0;
}
static int next_knot(const Chain *c, int k) {
    for (int w = (k + 1) >> 6; w < c->words; ++w) {
        uint64_t m = c->knot[w];
        if (w == ((k + 1) >> 6)) m &= ~(uint64_t)0 << ((k + 1) & 63);
        if (m) { const int r = (w << 6) + __builtin_ctzll(m); return r < c->L ? r : c->L; }
    }
    return c->L;
}
/* height of the string at a knot, the chain start (0, 0) or the chain end (L, S[L]) */
static double height(const Chain *c, int k) {
    if (k <= 0 || k >= c->L) return c->S[k <= 0 ? 0 : c->L];
    return c->S[k] + (get(c->up, k) ? c->lam : -c->lam);
}

/* exact sequential search between the fixed points a and b (their heights as above): sets the knots found inside */
static void search(Chain *c, int a, int b) {
    int k0 = a;
    double y0 = height(c, a);
    const double yb = height(c, b);
    for (;;) {
        double hi = INFINITY, lo = -INFINITY;
        int khi = k0, klo = k0, found = 0, fk = 0, fs = 0;
        for (int k = k0 + 1; k <= b; ++k) {
            double h, l;
            if (k == b) h = l = (yb - y0) / (k - k0);
            else { h = (c->S[k] + c->lam - y0) / (k - k0); l = (c->S[k] - c->lam - y0) / (k - k0); }
            if (l > hi) { found = 1; fk = khi; fs = 1; break; }
            if (h < lo) { found = 1; fk = klo; fs = 0; break; }
            if (h <= hi) { hi = h; khi = k; }
            if (l >= lo) { lo = l; klo = k; }
        }
        if (!found) return;
        set(c->knot, fk);
        if (fs) set(c->up, fk); else clr(c->up, fk);
        k0 = fk; y0 = height(c, fk);
    }
}

/* out[L]; returns the number of cells handed to the sequential search; *passes = certificate passes */
int cert_prox(int L, const double *v, double lam, const double *z_prev, double *out, int *passes) {
    *passes = 0;
    if (L == 1 || lam == 0) { memcpy(out, v, sizeof(double) * (size_t)L); return 0; }
    const int words = (L >> 6) + 1;
    double *S = (double *)malloc(sizeof(double) * (size_t)(L + 1));
    uint64_t *mask = (uint64_t *)calloc((size_t)4 * words, sizeof(uint64_t));
    Chain c = {mask, mask + words, mask + 2 * words, mask + 3 * words, L, words, S, lam};
    const double c0 = v[0];
    S[0] = 0;
    for (int k = 0; k < L; ++k) S[k + 1] = S[k] + (v[k] - c0);
    for (int k = 1; k < L; ++k)                       /* hypothesis: the jumps of the previous iterate */
        if (z_prev[k] != z_prev[k - 1]) { set(c.knot, k); if (z_prev[k] > z_prev[k - 1]) set(c.up, k); }
    int searched = 0;
    for (;;) {
        ++*passes;
        int any = 0;
        memset(c.bad, 0, sizeof(uint64_t) * (size_t)words);
        /* per cell (one thread each on the GPU): is the line of my piece inside the tube at my position? */
        for (int k = 1; k < L; ++k) {
            if (get(c.knot, k)) continue;
            const int a = prev_knot(&c, k), b = next_knot(&c, k);
            const double ya = height(&c, a), yb = height(&c, b);
            const double y = ya + (yb - ya) / (b - a) * (k - a);
            const double tol = 1e-12 * (1.0 + fabs(S[k]));
            if (y > S[k] + lam + tol || y < S[k] - lam - tol) { set(c.bad, a); any = 1; }   /* piece keyed by its left knot */
        }
        /* per knot: does the string bend towards the tube edge it touches? otherwise both adjacent pieces are dirty and
         * the knot is dropped -- after every knot has been judged on the same hypothesis (c.drop collects them) */
        memset(c.drop, 0, sizeof(uint64_t) * (size_t)words);
        for (int k = 1; k < L; ++k) {
            if (!get(c.knot, k)) continue;
            const int a = prev_knot(&c, k - 1), b = next_knot(&c, k);
            const double yk = height(&c, k);
            const double bend = (height(&c, b) - yk) / (b - k) - (yk - height(&c, a)) / (k - a);
            if (!((get(c.up, k) ? bend : -bend) > 0)) { set(c.bad, a); set(c.bad, k); set(c.drop, k); any = 1; }
        }
        if (!any) break;
        /* maximal runs of dirty pieces -> one exact search each (independent: one warp pair each on the GPU) */
        int a = 0;
        while (a < L) {
            int b = next_knot(&c, a);
            if (!get(c.bad, a)) { a = b; continue; }
            while (b < L && (get(c.bad, b) || get(c.drop, b))) { clr(c.knot, b); b = next_knot(&c, b); }   /* knots inside a dirty run go */
            searched += b - a;
            search(&c, a, b);
            a = b;
        }
    }
    /* values: slope of the piece each cell lies on */
    for (int a = 0; a < L;) {
        const int b = next_knot(&c, a);
        const double x = c0 + (height(&c, b) - height(&c, a)) / (b - a);
        for (int k = a; k < b; ++k) out[k] = x;
        a = b;
    }
    free(S); free(mask);
    return searched;
}
