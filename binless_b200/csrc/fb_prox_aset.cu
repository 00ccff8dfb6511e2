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
#include "fb_prox_aset_passes.cuh"
#include "fb_ctl.cuh"
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstddef>

namespace fb {

namespace {
using namespace aset;
constexpr unsigned FULL = 0xffffffffu;
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

template <int CAP>
__global__ void __launch_bounds__(AS_NT, (CAP <= 2432 ? 4 : CAP <= 3072 ? 3 : 2))
k_prox_aset(GflCtl *__restrict__ ctl, const AsChain *__restrict__ chains, const AsBatch *__restrict__ batches,
            const AsWord *__restrict__ words, const int nbatches, const int maxpass,
            const double *__restrict__ beta, const double *__restrict__ betaT,
            double *__restrict__ zr, double *__restrict__ ur, double *__restrict__ zc, double *__restrict__ uc,
            double *__restrict__ partials, unsigned long long *__restrict__ stats, const GflFold fold) {
    if (ctl->done || !ctl->prox_sel) return;   // !prox_sel: this iteration belongs to the team kernel k_prox
    extern __shared__ __align__(16) unsigned char smem_raw[];
    SmT<CAP> &s = *reinterpret_cast<SmT<CAP> *>(smem_raw);
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const unsigned lt = (1u << lane) - 1u;
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
            s.Lc[c] = ch.L; s.Bc[c] = ch.base; s.Qc[c] = ch.q;
        }
        if (tid < nwu) s.WD[tid] = words[bt.w0 + tid];
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
                        const AsWord d = s.WD[g];
                        const long long o = d.goff + lane;
                        if (lane < d.nvalid) {
                            bq[x] = __ldcg(Bv + o); uq[x] = __ldcg(Uv + o); zq[x] = __ldcg(Zv + o);
                            if (lane == 0 && d.jw > 0) z0[x] = __ldcg(Zv + o - 1);
                        }
                    }
                }
#pragma unroll
                for (int x = 0; x < AS_UN; ++x) {
                    const int g = g0 + x * AS_NW;
                    if (g < nwu) {
                        const AsWord d = s.WD[g];
                        const bool in = lane < d.nvalid;
                        const double zp = zq[x];
                        double zl = __shfl_up_sync(FULL, zp, 1);
                        if (lane == 0) zl = z0[x];
                        const double v = bq[x] + uq[x];
                        if (in) s.S[(g << 5) + lane + 1] = v;
                        const bool head = (d.jw == 0) && lane == 0;       // the first cell of a chain
                        if (head) s.Cc[d.cidx] = v;
                        const bool knot = in && !head && zp != zl;
                        const unsigned kb = __ballot_sync(FULL, knot);
                        const unsigned ub = __ballot_sync(FULL, knot && zp > zl);
                        s.PM[(g << 5) + lane] = 0u;
                        if (lane == 0) {
                            const unsigned fx = fix_bits(d);
                            s.K[g] = kb | fx; s.UP[g] = ub; s.FIXM[g] = fx; s.CH[g] = 0u; s.ENDB[g] = (unsigned char)(d.endbit >= 0 ? d.endbit : 32);
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
                    double *S = s.S + s.Bc[c];
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
#ifndef FB_AS_NO_PREFETCH
            // the next batch of this CTA is known (queue head fetched at the top): ask L2 for its beta / u / z lines now,
            // so that the staging loads of the next round find them there instead of paying the HBM latency
            {
                const int bnext = s.ci[C_BATCH];
                if (bnext < nbatches) {
                    const AsBatch nb = batches[bnext];
                    if (tid < (nb.npts >> 5)) {
                        const AsWord d = words[nb.w0 + tid];
                        if (d.nvalid > 0) {
                            const double *pb = (nb.dir == 0 ? beta : betaT) + d.goff, *pu = (nb.dir == 0 ? ur : uc) + d.goff,
                                         *pz = (nb.dir == 0 ? zr : zc) + d.goff;
                            const int last = d.nvalid - 1;
#pragma unroll
                            for (int q = 0; q < 3; ++q) {
                                const int o = q == 0 ? 0 : (q == 1 ? min(16, last) : last);
                                asm volatile("prefetch.global.L2 [%0];" ::"l"(pb + o));
                                asm volatile("prefetch.global.L2 [%0];" ::"l"(pu + o));
                                asm volatile("prefetch.global.L2 [%0];" ::"l"(pz + o));
                            }
                        }
                    }
                }
            }
#endif
            // ---- P: passes until no knot changes ----
            if (tid < AS_TWN) {     // the first pass looks at every word
                const int left = nwu - (tid << 5);
                s.TW[0][tid] = left >= 32 ? 0xffffffffu : left > 0 ? (1u << left) - 1u : 0u;
                s.TW[1][tid] = 0u;
            }
            __syncthreads();
            bool first = true;
            int cur = 0;
            for (;;) {
                if (npass >= maxpass) { fell = 1; break; }
                ++npass;
                const bool act = tid < nwu && ((s.TW[cur][tid >> 5] >> (tid & 31)) & 1u);
                if (act) pass_slopes(s, tid, lam, first);
                __syncthreads();
                if (act) pass_bends(s, tid);
                for (int g = wid; g < nwu; g += AS_NW)
                    if ((s.TW[cur][g >> 5] >> (g & 31)) & 1u) pass_points(s, g, lam, lane, lt);
                __syncthreads();
                if (act) pass_apply(s, tid, cur ^ 1);
                __syncthreads();
                unsigned any = 0u;
#pragma unroll
                for (int k = 0; k < AS_TWN; ++k) any |= s.TW[cur ^ 1][k];
                if (tid < AS_TWN) s.TW[cur][tid] = 0u;
                if (!any) break;
                cur ^= 1;
                first = false;
            }
            if (fell) {     // not settled: the sequential taut string, one lane per chain; then every slope again
                if (tid < bt.nch) chain_sequential(s, s.Bc[tid], s.Lc[tid], lam);
                __syncthreads();
                if (tid < nwu) pass_slopes(s, tid, lam, true);
                __syncthreads();
            }
        }
        if (stats) tc3 = clock64();
        // ---- D: z = c + slope of the piece, u += beta - z, residual sums (gfl_graph_fl.c:306-365) ----
        // sums: per lane over the words a warp sees of one chain (chain word j goes to slot j mod 8 whatever the batch),
        // then over the lanes, then over the slots: one order per chain
        double *WP = s.S;            // S is free now: 16 doubles per chain
        {
            double accp = 0.0, accd = 0.0;
            int curc = -1, curs = 0;
#pragma unroll 1
            for (int g0 = wid; g0 < nwu; g0 += AS_NW * AS_UN) {
                double bq[AS_UN], uq[AS_UN], zq[AS_UN];
#pragma unroll
                for (int x = 0; x < AS_UN; ++x) {
                    const int g = g0 + x * AS_NW;
                    bq[x] = uq[x] = zq[x] = 0.0;
                    if (g < nwu) {
                        const AsWord d = s.WD[g];
                        const long long o = d.goff + lane;
                        if (lane < d.nvalid) { bq[x] = __ldcg(Bv + o); uq[x] = __ldcg(Uv + o); zq[x] = __ldcg(Zv + o); }
                    }
                }
#pragma unroll
                for (int x = 0; x < AS_UN; ++x) {
                    const int g = g0 + x * AS_NW;
                    if (g < nwu) {
                        const AsWord d = s.WD[g];
                        const bool in = lane < d.nvalid;
                        if (d.cidx != curc) {
                            if (curc >= 0) {
                                accp = warp_sum(accp); accd = warp_sum(accd);
                                if (lane == 0) { WP[16 * curc + 2 * curs] = accp; WP[16 * curc + 2 * curs + 1] = accd; }
                            }
                            curc = d.cidx; curs = d.jw & (AS_NW - 1); accp = 0.0; accd = 0.0;
                        }
                        double z = bq[x] + uq[x];
                        if (!copy) {
                            // cell i lies between the points p = 32 g + lane and p + 1: its piece starts at the last knot <= p
                            const unsigned lm = s.K[g] & (0xffffffffu >> (31 - lane));
                            const int pa = !in ? 0 : lm ? (g << 5) + 31 - __clz(lm) : (int)s.WPREV[g];
                            z = s.Cc[d.cidx] + s.SLP[pa];
                        }
                        if (in) {
                            const long long o = d.goff + lane;
                            const double r = bq[x] - z;
                            __stcg(Uv + o, uq[x] + r);
                            __stcg(Zv + o, z);
                            accp += r * r;
                            const double dd = alpha * (z - zq[x]);
                            accd += dd * dd;
                        }
                    }
                }
            }
            if (curc >= 0) {
                accp = warp_sum(accp); accd = warp_sum(accd);
                if (lane == 0) { WP[16 * curc + 2 * curs] = accp; WP[16 * curc + 2 * curs + 1] = accd; }
            }
        }
        __syncthreads();
        for (int c = tid; c < bt.nch; c += AS_NT) {
            const int nslot = min(AS_NW, (s.Lc[c] >> 5) + 1);
            double ap = 0.0, ad = 0.0;
            for (int k = 0; k < nslot; ++k) { ap += WP[16 * c + 2 * k]; ad += WP[16 * c + 2 * k + 1]; }
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
    // the last CTA to finish ends the iteration (residual norms, convergence, penalty): no k_ctl launch. S and SLP are
    // adjacent arrays of CAP >= 1024 doubles each: room for the 2 x 1024 sums of the reduction
    static_assert(offsetof(SmT<CAP>, SLP) == sizeof(double) * CAP && CAP >= CTL_VT, "S and SLP hold the reduction");
    ctl_fold(ctl, fold, partials, &s.ci[1], &s.S[0]);
}

} // namespace

// ==========================================================================
namespace {
constexpr int AS_CAPS[] = {1024, 2048, 2432, 3072, 4096, 5632};
size_t sizeof_smem(int cap) {
    switch (cap) {
        case 1024: return sizeof(SmT<1024>); case 2048: return sizeof(SmT<2048>); case 2432: return sizeof(SmT<2432>);
        case 3072: return sizeof(SmT<3072>); case 4096: return sizeof(SmT<4096>); case 5632: return sizeof(SmT<5632>);
    }
    return ~size_t(0);
}
const void *kernel_of(int cap) {
    switch (cap) {
        case 1024: return (const void *)k_prox_aset<1024>; case 2048: return (const void *)k_prox_aset<2048>;
        case 2432: return (const void *)k_prox_aset<2432>; case 3072: return (const void *)k_prox_aset<3072>;
        case 4096: return (const void *)k_prox_aset<4096>; case 5632: return (const void *)k_prox_aset<5632>;
    }
    return nullptr;
}
} // namespace

void aset_plan_build(AsetPlan &plan, const Grid &g, int r0, int r1, int c0, int c1, int n_sm) {
    plan.usable = false;
    int dev = 0, smem_max = 0;
    FB_CUDA(cudaGetDevice(&dev));
    FB_CUDA(cudaDeviceGetAttribute(&smem_max, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev));
    const int lmax = g.row_len(0);
    // capacity classes: the kernel is instantiated per class (compile-time shared-memory layout)
    const int need_cap = ((lmax + 1 + 31) / 32) * 32;
    int cap = 0;
    for (int c : AS_CAPS) if (c >= need_cap) { cap = c; break; }
    if (const char *e = std::getenv("FB_ASET_CAP")) { const int want = std::atoi(e); cap = 0; for (int c : AS_CAPS) if (c >= need_cap && c >= want) { cap = c; break; } }
    if (!cap || sizeof_smem(cap) > (size_t)smem_max) return;
    // measured on B200 (tools/prox_ab.py): faster than k_prox up to ~3600 bins, slower above (two CTAs per SM, one chain per batch)
    if (cap > 3072 && !std::getenv("FB_ASET_CAP") && !std::getenv("FB_PROX_ASET")) return;
    std::vector<AsChain> chains;
    std::vector<AsBatch> batches;
    std::vector<AsWord> words;
    std::vector<int64_t> cells;
    auto add_dir = [&](int dir, int lo, int hi) {
        AsBatch cur{(int)chains.size(), 0, dir, 0, (int)words.size()};
        int64_t cur_cells = 0;
        auto flush = [&]() {
            if (cur.nch) { batches.push_back(cur); cells.push_back(cur_cells); }
            cur = AsBatch{(int)chains.size(), 0, dir, 0, (int)words.size()}; cur_cells = 0;
        };
        for (int x = lo; x < hi; ++x) {
            AsChain ch;
            if (dir == 0) { ch.L = g.row_len(x); ch.g = g.rm_row_start(x); ch.q = 2 * x; }
            else { ch.L = g.col_len(x); ch.g = g.cm_col_start(x); ch.q = 2 * (g.n - 1 - x) + 1; }
            const int need = ((ch.L + 1 + 31) / 32) * 32;
            if (cur.npts + need > cap) flush();
            ch.base = cur.npts; ch.pad = 0;
            chains.push_back(ch);
            for (int j = 0; j < need / 32; ++j) {
                AsWord wd;
                wd.goff = ch.g + 32 * (int64_t)j;
                wd.nvalid = (short)std::max(0, std::min(32, ch.L - 32 * j));
                wd.cidx = (short)cur.nch;
                wd.endbit = (short)((ch.L >= 32 * j && ch.L < 32 * j + 32) ? ch.L - 32 * j : -1);
                wd.jw = (short)j;
                words.push_back(wd);
            }
            cur.npts += need; cur.nch += 1; cur_cells += ch.L;
        }
        flush();
    };
    add_dir(0, r0, std::min(r1, g.n - 1));      // rows 0..n-2 are chains
    add_dir(1, std::max(c0, 1), c1);            // columns 1..n-1
    plan.maxpass = 64;
    if (const char *e = std::getenv("FB_ASET_MAXPASS")) plan.maxpass = std::max(0, std::atoi(e));   // test aid: 0 = always the fallback
    if (batches.empty()) { plan.nbatches = 0; plan.usable = true; plan.cap = cap; plan.smem = sizeof_smem(cap); plan.grid = 1; return; }
    // most cells first: the dynamic queue then ends with the small batches
    std::vector<int> order(batches.size());
    for (size_t k = 0; k < order.size(); ++k) order[k] = (int)k;
    std::stable_sort(order.begin(), order.end(), [&](int a, int b) { return cells[a] > cells[b]; });
    std::vector<AsBatch> sorted(batches.size());
    for (size_t k = 0; k < order.size(); ++k) sorted[k] = batches[order[k]];
    plan.chains.alloc(chains.size()); plan.batches.alloc(sorted.size()); plan.words.alloc(words.size());
    FB_CUDA(cudaMemcpy(plan.words.p, words.data(), words.size() * sizeof(AsWord), cudaMemcpyHostToDevice));
    FB_CUDA(cudaMemcpy(plan.chains.p, chains.data(), chains.size() * sizeof(AsChain), cudaMemcpyHostToDevice));
    FB_CUDA(cudaMemcpy(plan.batches.p, sorted.data(), sorted.size() * sizeof(AsBatch), cudaMemcpyHostToDevice));
    plan.cap = cap; plan.nbatches = (int)sorted.size(); plan.smem = sizeof_smem(cap);
    const void *fn = kernel_of(cap);
    FB_CUDA(cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)plan.smem));
    int per_sm = 1;
    FB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, fn, AS_NT, plan.smem));
    plan.grid = std::max(1, std::min(plan.nbatches, n_sm * std::max(1, per_sm)));
    if (std::getenv("FB_ASET_STATS")) plan.stats.alloc(8);
    plan.usable = true;
}

void aset_prox_launch(const AsetPlan &plan, GflCtl *ctl, const double *beta, const double *betaT, double *zr, double *ur,
                      double *zc, double *uc, double *partials, const GflFold &fold, cudaStream_t st) {
    if (!plan.nbatches) return;
    if (plan.stats.p) FB_CUDA(cudaMemsetAsync(plan.stats.p, 0, 8 * sizeof(unsigned long long), st));
#define FB_ASET_LAUNCH(CAP_)                                                                                                   \
    case CAP_: k_prox_aset<CAP_><<<plan.grid, AS_NT, plan.smem, st>>>(ctl, plan.chains.p, plan.batches.p, plan.words.p, plan.nbatches, \
                                                                     plan.maxpass, beta, betaT, zr, ur, zc, uc, partials, plan.stats.p, fold); break;
    switch (plan.cap) {
        FB_ASET_LAUNCH(1024) FB_ASET_LAUNCH(2048) FB_ASET_LAUNCH(2432) FB_ASET_LAUNCH(3072) FB_ASET_LAUNCH(4096) FB_ASET_LAUNCH(5632)
        default: throw StateError("k_prox_aset: no kernel for this capacity");
    }
#undef FB_ASET_LAUNCH
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
