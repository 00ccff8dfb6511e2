/*
 * oracle/gfl_port.c -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * Plain-C CPU restatement of the graph-fused-lasso ADMM solver that the
 * reference's fast-binless path calls:
 *
 *   graph_fused_lasso_weight_warm  /root/reference/binless/src/gfl_graph_fl.c:53-213
 *   update_beta_weight             gfl_graph_fl.c:241-264
 *   update_z                       gfl_graph_fl.c:267-304
 *   update_u                       gfl_graph_fl.c:306-327
 *   primal_resnorm / dual_resnorm  gfl_graph_fl.c:329-365
 *   tf_dp_weight (1-D DP)          /root/reference/binless/src/gfl_tf.c:184-321
 *
 * It is written so that every floating-point expression is evaluated with the
 * same operands in the same order as the reference, hence results are expected
 * to be bit-identical to the reference C code compiled with the same compiler
 * flags (tests/test_oracle_gfl.py checks exactly that against oracle/_ref,
 * which is the reference's own source compiled in place).
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline/reference
 * legs may load this. The product (binless_b200/) never links or calls it.
 */
#include <stdlib.h>
#include <string.h>
#include <math.h>
#include <stdint.h>

#define GFLP_PENALTY_PERIOD 50 /* VARYING_PENALTY_DELAY, gfl_graph_fl.h:18 */

/* ---- 1-D weighted fused lasso by dynamic programming (gfl_tf.c:184-321) ----
 * Minimises  sum_k w_k/2 (b_k - y_k)^2 + lam * sum_k |b_{k+1} - b_k|.
 * kx/ka/kb hold the knots of the clipped derivative of the running message as
 * a double-ended queue living in [lo_end, hi_end] of arrays of size 2n;
 * tneg/tpos are the back-pointers (where the derivative hits -lam / +lam).   */
void gflp_tf_dp_weight(int n, const double *y, const double *w, double lam,
                       double *out, double *kx, double *ka, double *kb,
                       double *tneg, double *tpos)
{
    int k, lo, hi, lo_end, hi_end;
    double a_left, b_left, a_right, b_right; /* fresh quadratic seen from either side */
    double sa, sb;                           /* running slope / intercept while scanning */
    double ta, tb;

    if (n == 0) return;
    if (n == 1 || lam == 0) { /* gfl_tf.c:210-215 */
        for (k = 0; k < n; k++) out[k] = y[k];
        return;
    }

    /* first point: derivative is w0*(b - y0), clipped at +-lam (gfl_tf.c:229-242) */
    tneg[0] = -lam / w[0] + y[0];
    tpos[0] = lam / w[0] + y[0];
    lo_end = n - 1;
    hi_end = n;
    kx[lo_end] = tneg[0];
    kx[hi_end] = tpos[0];
    ka[lo_end] = w[0];
    kb[lo_end] = -w[0] * y[0] + lam;
    ka[hi_end] = -w[0];
    kb[hi_end] = w[0] * y[0] + lam;
    a_left = w[1];
    b_left = -w[1] * y[1] - lam;
    a_right = -w[1];
    b_right = w[1] * y[1] - lam;

    for (k = 1; k < n - 1; k++) { /* gfl_tf.c:245-290 */
        /* from the left: pop knots until the derivative exceeds -lam */
        sa = a_left;
        sb = b_left;
        for (lo = lo_end; lo <= hi_end; lo++) {
            if (sa * kx[lo] + sb > -lam) break;
            sa += ka[lo];
            sb += kb[lo];
        }
        /* from the right: pop knots until the derivative is below +lam */
        ta = a_right;
        tb = b_right;
        for (hi = hi_end; hi >= lo; hi--) {
            if (-ta * kx[hi] - tb < lam) break;
            ta += ka[hi];
            tb += kb[hi];
        }
        tneg[k] = (-lam - sb) / sa;
        lo_end = lo - 1;
        kx[lo_end] = tneg[k];

        tpos[k] = (lam + tb) / (-ta);
        hi_end = hi + 1;
        kx[hi_end] = tpos[k];

        ka[lo_end] = sa;
        kb[lo_end] = sb + lam;
        ka[hi_end] = ta;
        kb[hi_end] = tb + lam;
        a_left = w[k + 1];
        b_left = -w[k + 1] * y[k + 1] - lam;
        a_right = -w[k + 1];
        b_right = w[k + 1] * y[k + 1] - lam;
    }

    /* last point: zero of the derivative (gfl_tf.c:294-302) */
    sa = a_left;
    sb = b_left;
    for (lo = lo_end; lo <= hi_end; lo++) {
        if (sa * kx[lo] + sb > 0) break;
        sa += ka[lo];
        sb += kb[lo];
    }
    out[n - 1] = -sb / sa;

    /* back-substitution through the stored knots (gfl_tf.c:306-311) */
    for (k = n - 2; k >= 0; k--) {
        if (out[k + 1] > tpos[k]) out[k] = tpos[k];
        else if (out[k + 1] < tneg[k]) out[k] = tneg[k];
        else out[k] = out[k + 1];
    }
}

/* ---- ADMM over trails with warm start (gfl_graph_fl.c:53-213) -------------
 * trails[0..nz): cell index of each trail slot; breakpoints[t]: end slot of
 * trail t. Returns step (= iterations performed + 1), like the reference.    */
int gflp_graph_fused_lasso_weight_warm(int n, const double *y, const double *w,
                                       int ntrails, const unsigned *trails,
                                       const unsigned *breakpoints, double lam,
                                       double *alpha, double inflate, int maxsteps,
                                       double converge, double *beta, double *z,
                                       double *u)
{
    const int64_t nz = breakpoints[ntrails - 1];
    int64_t s, t, first, last, len;
    int i, c, step, longest;
    double *zcur = z, *zprev, *swap;
    double *ybuf, *wbuf, *dp;
    int *cnt, *start, *fill, *slot_of;
    double rp, rd, cur, acc, bi, diff;

    zprev = (double *)malloc(nz * sizeof(double));
    cnt = (int *)calloc(n, sizeof(int));      /* nzmap */
    start = (int *)malloc((n + 1) * sizeof(int));
    fill = (int *)malloc(n * sizeof(int));
    slot_of = (int *)malloc(nz * sizeof(int)); /* zmap: slots of each cell, ascending */
    dp = (double *)malloc((nz * 8 - 2 * (int64_t)ntrails) * sizeof(double));
    memcpy(zprev, z, nz * sizeof(double));

    longest = breakpoints[0];
    for (i = 1; i < ntrails; i++)
        if ((int)(breakpoints[i] - breakpoints[i - 1]) > longest)
            longest = breakpoints[i] - breakpoints[i - 1];
    ybuf = (double *)malloc(longest * sizeof(double));
    wbuf = (double *)malloc(longest * sizeof(double));
    for (i = 0; i < longest; i++) wbuf[i] = (*alpha) / 2.0;

    for (s = 0; s < nz; s++) cnt[trails[s]]++;
    start[0] = 0;
    for (c = 0; c < n; c++) { start[c + 1] = start[c] + cnt[c]; fill[c] = 0; }
    for (s = 0; s < nz; s++) { c = trails[s]; slot_of[start[c] + fill[c]++] = (int)s; }

    step = 1;
    cur = converge + 1;
    while (step < maxsteps && cur > converge) {
        /* beta update (gfl_graph_fl.c:241-264) */
        for (c = 0; c < n; c++) {
            if (cnt[c] == 0) continue;
            acc = 0;
            for (i = start[c]; i < start[c + 1]; i++) acc += zcur[slot_of[i]] - u[slot_of[i]];
            beta[c] = (2.0 * y[c] * w[c] + (*alpha) * acc) / (2.0 * w[c] + (*alpha) * cnt[c]);
        }
        swap = zcur; zcur = zprev; zprev = swap;

        /* z update: one 1-D fused lasso per trail (gfl_graph_fl.c:267-304) */
        first = 0;
        for (t = 0; t < ntrails; t++) {
            last = breakpoints[t];
            len = last - first;
            for (s = first; s < last; s++) ybuf[s - first] = beta[trails[s]] + u[s];
            {
                double *kx = dp + 8 * first - 2 * t;
                gflp_tf_dp_weight((int)len, ybuf, wbuf, lam, zcur + first, kx, kx + 2 * len,
                                  kx + 4 * len, kx + 6 * len, kx + 7 * len - 1);
            }
            first = last;
        }

        /* u update (gfl_graph_fl.c:306-327) */
        for (c = 0; c < n; c++) {
            if (cnt[c] == 0) continue;
            bi = beta[c];
            for (i = start[c]; i < start[c + 1]; i++) u[slot_of[i]] += bi - zcur[slot_of[i]];
        }

        /* residual norms (gfl_graph_fl.c:329-365) */
        acc = 0;
        for (c = 0; c < n; c++) {
            if (cnt[c] == 0) continue;
            bi = beta[c];
            for (i = start[c]; i < start[c + 1]; i++)
                acc += (bi - zcur[slot_of[i]]) * (bi - zcur[slot_of[i]]);
        }
        rp = sqrt(acc / n);
        acc = 0;
        for (s = 0; s < (int)nz; s++) {
            diff = (*alpha) * (zcur[s] - zprev[s]);
            acc += diff * diff;
        }
        rd = sqrt(acc / (int)nz);
        cur = rp > rd ? rp : rd;

        /* varying penalty (gfl_graph_fl.c:168-179) */
        if (step % GFLP_PENALTY_PERIOD == 0 && rp > 10 * rd) {
            (*alpha) *= inflate;
            for (s = 0; s < nz; s++) u[s] /= inflate;
            for (i = 0; i < longest; i++) wbuf[i] = (*alpha) / 2.0;
        } else if (step % GFLP_PENALTY_PERIOD == 0 && rd > 10 * rp) {
            (*alpha) /= inflate;
            for (s = 0; s < nz; s++) u[s] *= inflate;
            for (i = 0; i < longest; i++) wbuf[i] = (*alpha) / 2.0;
        }
        step++;
    }
    if (zcur != z) memcpy(z, zcur, nz * sizeof(double));
    free(zcur == z ? zprev : zcur);
    free(cnt); free(start); free(fill); free(slot_of); free(dp); free(ybuf); free(wbuf);
    return step;
}
