"""CPU prototype of the CERTIFICATE-FIRST chain prox planned for the next round.

The 1-D fused-lasso prox  min_x 1/2 sum (x_k - v_k)^2 + lam sum |x_{k+1} - x_k|  is the taut string through the tube
S_k -+ lam around the prefix sums S of v: a piecewise-linear path from (0, 0) to (L, S_L) that stays inside the tube,
bends only where it touches the tube, upwards (slope increases) at a contact with the UPPER edge and downwards at a
contact with the LOWER edge. Those conditions are the KKT conditions of a strictly convex problem, so ANY path that
satisfies them is the solution. Hence:

  1. take the knots (positions and sides) of the previous ADMM iterate as a hypothesis;
  2. join consecutive knots by straight lines and test, for every cell independently, that the line stays in the tube,
     and for every knot that the bend has the right sign  -- a streaming, fully parallel pass;
  3. only the ranges whose certificate fails are searched exactly (between verified knots as fixed end points, which the
     CUDA kernel already supports), and the bends at their ends are re-tested; failing ends merge ranges.

This file proves the scheme on the CPU against the reference's own tf_dp_weight (oracle/_ref, compiled from
/root/reference) and measures, on the iterates of a real solve, how many cells still need the sequential search.
TEST / DESIGN AID ONLY: nothing here is on the product path.

    python tools/certificate_prox_prototype.py [nbins]
"""
import ctypes as C
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def tf_dp_reference():
    """tf_dp_weight of the reference (gfl_tf.c:184-321) through ctypes."""
    path = os.path.join(ROOT, "oracle", "_ref", "libgfl_ref.so")
    lib = C.CDLL(path)
    dp = C.POINTER(C.c_double)
    lib.tf_dp_weight.argtypes = [C.c_int, dp, dp, C.c_double, dp, dp, dp, dp, dp, dp]
    lib.tf_dp_weight.restype = None

    def solve(v, w, lam):
        n = v.size
        v = np.ascontiguousarray(v, dtype=np.float64); wv = np.full(n, float(w))
        out = np.empty(n); buf = np.empty(8 * n)
        p = lambda a: a.ctypes.data_as(dp)
        off = lambda k: C.cast(C.addressof(p(buf).contents) + 8 * k, dp)
        lib.tf_dp_weight(n, p(v), p(wv), float(lam), p(out), off(0), off(2 * n), off(4 * n), off(6 * n), off(7 * n - 1))
        return out
    return solve


def taut_between(S, lam, a, b, sa, sb):
    """Exact taut string on [a, b] through (a, S[a] + sa*lam) and (b, S[b] + sb*lam), tube S[k] -+ lam inside.
    Returns the interior knots [(k, side)], side +1 = upper edge. Plain sequential search (what the CUDA kernel does
    160 candidates at a time)."""
    knots = []
    k0, y0 = a, S[a] + sa * lam
    while True:
        hi, lo = np.inf, -np.inf
        khi = klo = k0
        found = None
        for k in range(k0 + 1, b + 1):
            if k == b:
                h = l = (S[b] + sb * lam - y0) / (k - k0)
            else:
                h = (S[k] + lam - y0) / (k - k0); l = (S[k] - lam - y0) / (k - k0)
            if l > hi:                      # must bend upwards: the segment ends where the upper bound was attained
                found = (khi, +1); break
            if h < lo:
                found = (klo, -1); break
            if h <= hi:
                hi, khi = h, k
            if l >= lo:
                lo, klo = l, k
        if found is None:
            return knots
        knots.append(found)
        k0, y0 = found[0], S[found[0]] + found[1] * lam


def solution_from_knots(S, lam, knots, L):
    pts = [(0, 0.0)] + [(k, S[k] + s * lam) for k, s in knots] + [(L, S[L])]
    x = np.empty(L)
    for (k1, y1), (k2, y2) in zip(pts[:-1], pts[1:]):
        x[k1:k2] = (y2 - y1) / (k2 - k1)
    return x


def certified_prox(v, lam, z_prev):
    """Returns (solution, cells searched sequentially, number of passes)."""
    L = v.size
    c = v[0]
    S = np.concatenate([[0.0], np.cumsum(v - c)])
    if lam == 0 or L == 1:
        return v.copy(), 0, 0
    # hypothesis: where the previous solution jumps, and in which direction
    d = np.diff(z_prev)
    knots = [(int(k) + 1, 1 if d[k] > 0 else -1) for k in np.flatnonzero(d != 0)]
    searched = 0
    passes = 0
    while True:
        passes += 1
        pts = [(0, 0)] + knots + [(L, 0)]
        ys = [S[k] + s * lam for k, s in pts]
        slopes = [(ys[i + 1] - ys[i]) / (pts[i + 1][0] - pts[i][0]) for i in range(len(pts) - 1)]
        bad_piece = np.zeros(len(slopes), dtype=bool)
        for i, ((k1, _), (k2, _)) in enumerate(zip(pts[:-1], pts[1:])):      # parallel over cells on the GPU
            if k2 - k1 > 1:
                kk = np.arange(k1 + 1, k2)
                line = ys[i] + slopes[i] * (kk - k1)
                tol = 1e-12 * (1.0 + np.abs(S[kk]))
                bad_piece[i] = bool(np.any(line > S[kk] + lam + tol) or np.any(line < S[kk] - lam - tol))
        bad_knot = np.zeros(len(pts), dtype=bool)
        for i in range(1, len(pts) - 1):                                       # parallel over knots
            bend = slopes[i] - slopes[i - 1]
            bad_knot[i] = not (pts[i][1] * bend > 0)
        if not bad_piece.any() and not bad_knot.any():
            return c + solution_from_knots(S, lam, knots, L), searched, passes
        # dirty ranges: failing pieces, plus the two pieces around a failing knot (the knot is dropped); merge overlaps
        dirty = []
        for i in np.flatnonzero(bad_piece):
            dirty.append([i, i + 1])
        for i in np.flatnonzero(bad_knot):
            dirty.append([i - 1, i + 1])
        dirty.sort()
        merged = [dirty[0]]
        for lo_, hi_ in dirty[1:]:
            if lo_ < merged[-1][1]:       # shares a piece with the previous range: the knot between them goes too
                merged[-1][1] = max(merged[-1][1], hi_)
            else:
                merged.append([lo_, hi_])
        new_knots = []
        last = 0
        for lo_, hi_ in merged:           # independent searches: one warp-pair each on the GPU
            new_knots += [p for p in pts[last + 1:lo_ + 1] if 0 < p[0] < L]
            (a, sa), (b, sb) = pts[lo_], pts[hi_]
            new_knots += taut_between(S, lam, a, b, sa, sb)
            searched += b - a
            last = hi_ - 1
        new_knots += [p for p in pts[last + 1:-1]]
        # drop duplicates keeping order
        seen, knots = set(), []
        for p in sorted(new_knots):
            if p[0] not in seen:
                seen.add(p[0]); knots.append(p)


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 300
    from oracle.loader import oracle_lib
    from binless_b200._abi import gfl_solve
    from binless_b200.synth import synth
    ref = tf_dp_reference()
    rng = np.random.default_rng(0)
    # 1. random chains, random (wrong) hypotheses: the result must equal tf_dp_weight whatever the hypothesis
    worst = 0.0
    for trial in range(200):
        L = int(rng.integers(2, 120)); lam = float(rng.choice([0.3, 1.0, 3.0]))
        v = rng.normal(size=L) + np.repeat(rng.normal(0, 1.5, L // 8 + 1), 8)[:L]
        want = ref(v, 1.0, lam)      # tf_dp_weight minimises 1/2 sum w (y - x)^2 + lam sum |dx|: tube radius lam / w
        hyp = ref(v + rng.normal(0, rng.choice([0.0, 0.05, 0.5]), L), 1.0, lam * rng.choice([0.7, 1.0, 1.4]))
        got, _, _ = certified_prox(v, lam, hyp)
        worst = max(worst, float(np.abs(got - want).max()))
    print(f"random chains with arbitrary hypotheses: max |certified - tf_dp_weight| = {worst:.2e}")
    assert worst < 1e-9
    # 2. the iterates of a real solve: hypothesis = previous iterate
    ora = oracle_lib("auto")
    mat = synth(n, 1, 5000, 2)
    with ora.create(mat, n) as h:
        out = h.run_signal(0.5, [2.5], 0.0)
    y, w = out["phihat"], out["weight"]
    nc = n * (n + 1) // 2
    alpha, lam2 = 5.0, 2.5
    lam = lam2 / (alpha / 2.0)
    prev = gfl_solve(ora, n, y, w, lam2, 1e-12, 2, 0)
    print("iter  cells   searched  share   passes(max)  max|err| vs reference iterate")
    for k in range(2, 14):
        cur = gfl_solve(ora, n, y, w, lam2, 1e-12, k + 1, 0)      # state after k iterations
        off = 0; tot = 0; srch = 0; pmax = 0; err = 0.0
        for i in range(n - 1):                                     # row chains (slots 0 .. nc-2, row-major)
            L = n - i
            z_k, u_k = cur["z"][off:off + L], cur["u"][off:off + L]
            v = z_k + u_k                                          # = beta_k + u_{k-1}: the prox input of iteration k
            z_prev = prev["z"][off:off + L]
            got, s, p = certified_prox(v, lam, z_prev)
            err = max(err, float(np.abs(got - z_k).max()))
            tot += L; srch += s; pmax = max(pmax, p)
            off += L
        print(f"{k:4d} {tot:7d} {srch:9d}  {srch / tot:5.2f}   {pmax:5d}        {err:.2e}")
        assert err < 1e-8
        prev = cur


if __name__ == "__main__":
    main()
