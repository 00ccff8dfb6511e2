// fb_ctl.cuh -- the end of one ADMM iteration (gfl_graph_fl.c:159-181): residual norms from the per-chain partial sums,
// convergence test, varying-penalty rule every 50 steps, bookkeeping for the next iteration. One device function used
// by the stand-alone kernel k_ctl (trail-block sharding: the partials are all-reduced first) and by the LAST CTA of the
// chain-prox kernels (everything else: no extra launch). The sums are formed in ONE order whatever the block size: 1024
// virtual threads, thread v adding the partials v, v + 1024, ... and a binary tree over the 1024 sums, so a sharded and
// an unsharded solve take bit-identical decisions.
#pragma once
#include "fb_gfl.cuh"

namespace fb {

constexpr int CTL_VT = 1024;   // virtual threads of the reduction (its shape)

// sm: 2 * CTL_VT doubles of shared memory; every thread of the calling block must call
static __device__ __noinline__ void ctl_tail(GflCtl *ctl, const double *partials, const int npart, const int64_t ncells, const int64_t nz,
                                         cudaGraphConditionalHandle cond, const bool graph, double *sm) {
    double *s_p = sm, *s_d = sm + CTL_VT;
    const int NT = (int)blockDim.x;
    // virtual thread v adds the partials v, v + 1024, ... in that order; a real thread serves four virtual ones at a time
    // and keeps their loads in flight together (the reduction is latency bound: ~10 rounds of L2 reads)
    constexpr int MAXV = 4;
    for (int vb = 0; vb < CTL_VT; vb += MAXV * NT) {
        double p[MAXV], d[MAXV];
#pragma unroll
        for (int i = 0; i < MAXV; ++i) { p[i] = 0.0; d[i] = 0.0; }
        for (int k0 = 0; k0 < npart; k0 += CTL_VT) {
            double a[MAXV], b[MAXV];
#pragma unroll
            for (int i = 0; i < MAXV; ++i) {
                const int v = vb + (int)threadIdx.x + i * NT, k = k0 + v;
                const bool ok = v < CTL_VT && k < npart;
                a[i] = ok ? __ldcg(partials + 2 * k) : 0.0;
                b[i] = ok ? __ldcg(partials + 2 * k + 1) : 0.0;
            }
#pragma unroll
            for (int i = 0; i < MAXV; ++i) {
                const int v = vb + (int)threadIdx.x + i * NT, k = k0 + v;
                if (v < CTL_VT && k < npart) { p[i] += a[i]; d[i] += b[i]; }
            }
        }
#pragma unroll
        for (int i = 0; i < MAXV; ++i) {
            const int v = vb + (int)threadIdx.x + i * NT;
            if (v < CTL_VT) { s_p[v] = p[i]; s_d[v] = d[i]; }
        }
    }
    __syncthreads();
    for (int o = CTL_VT / 2; o > 0; o >>= 1) {
        for (int v = threadIdx.x; v < o; v += NT) { s_p[v] += s_p[v + o]; s_d[v] += s_d[v + o]; }
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        const double rp = sqrt(s_p[0] / (double)ncells);
        const double rd = sqrt(s_d[0] / (double)nz);
        const double cur = rp > rd ? rp : rd;
        ctl->presnorm = rp; ctl->dresnorm = rd;
        int step = ctl->step;
        int rescale = 0;
        if (step % 50 == 0 && rp > 10 * rd) { ctl->alpha *= ctl->inflate; rescale = 1; }
        else if (step % 50 == 0 && rd > 10 * rp) { ctl->alpha /= ctl->inflate; rescale = 2; }
        ctl->rescale = rescale;
        ++step;
        ctl->step = step;
        ctl->next_chain = 0; // work queue of the next chain-prox launch
        ctl->ctas_done = 0;
        ctl->prox_sel = (ctl->aset_from >= 0 && step >= ctl->aset_from) ? 1 : 0;
        const int done = !(step < ctl->maxsteps && cur > ctl->converge);
        ctl->done = done;
        if (graph) cudaGraphSetConditional(cond, done ? 0u : 1u);   // the WHILE node runs the next iteration or stops
    }
}

// Called by every thread of every CTA of a chain-prox kernel when its work is done: the last CTA to arrive runs the tail.
// s_flag: one int of shared memory. The partials of this CTA must have been written before the call.
__device__ __forceinline__ void ctl_fold(GflCtl *ctl, const GflFold &f, const double *partials, int *s_flag, double *sm) {
    if (!f.on) return;
    __threadfence();                       // this CTA's partial sums are visible before its ticket
    __syncthreads();
    if (threadIdx.x == 0) *s_flag = (atomicAdd(&ctl->ctas_done, 1) == (int)gridDim.x - 1) ? 1 : 0;
    __syncthreads();
    if (*s_flag) {
        __threadfence();
        ctl_tail(ctl, partials, f.npart, f.ncells, f.nz, f.cond, f.graph != 0, sm);
    }
}

} // namespace fb
