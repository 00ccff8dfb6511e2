// fb_bin.cu -- bin_data() on the device: cut sites and read pairs -> dense packed triangles
// of observed / nobs per dataset, without the dense matrix ever existing on the host.
// Follows /root/reference/binless/R/preprocess.R:114-146 (CSnorm branch) and R/common.R:115-132. Integer work:
// results are bit-exact (int32 atomics are order independent).
#include "fb_model.cuh"

namespace fb {

namespace {

struct BinGeom {
    long long b0, top, res;   // borders b0 + k * res, k = 0..n; top = b0 + n * res
    int n;
    // cut(pos, borders, right = FALSE, include.lowest = TRUE); -1 = NA
    __host__ __device__ __forceinline__ int bin_of(int pos) const {
        if (pos < b0 || pos > top) return -1;
        if (pos == top) return n - 1;
        return (int)(((long long)pos - b0) / res);
    }
};

__global__ void k_cut_hist(BinGeom g, int D, int64_t ncuts, const int32_t *__restrict__ name, const int32_t *__restrict__ pos,
                           int *__restrict__ cuts, int *__restrict__ bad) {
    for (int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; k < ncuts; k += (int64_t)gridDim.x * blockDim.x) {
        const int d = name[k] - 1;
        if (d < 0 || d >= D) { *bad = 1; continue; }
        const int b = g.bin_of(pos[k]);
        if (b >= 0) atomicAdd(&cuts[(int64_t)d * g.n + b], 1);
    }
}

__global__ void k_read_hist(BinGeom g, int D, int64_t ncells, int64_t nreads, const int32_t *__restrict__ name,
                            const int32_t *__restrict__ p1, const int32_t *__restrict__ p2, const int32_t *__restrict__ c0,
                            const int32_t *__restrict__ c1, const int32_t *__restrict__ c2, const int32_t *__restrict__ c3,
                            int32_t *__restrict__ observed, int32_t *__restrict__ present, int *__restrict__ bad) {
    for (int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; k < nreads; k += (int64_t)gridDim.x * blockDim.x) {
        const int d = name[k] - 1;
        if (d < 0 || d >= D) { *bad = 1; continue; }
        const int i = g.bin_of(p1[k]), j = g.bin_of(p2[k]);
        if (i < 0 || j < 0 || i > j) continue;   // NA bins and pairs below the diagonal match no cell of the matrix
        const int v = c0[k] + c1[k] + c2[k] + c3[k];
        const int64_t cell = (int64_t)d * ncells + rm_start(i, g.n) + (j - i);
        if (v) atomicAdd(&observed[cell], v);
        present[cell] = 1;   // the cell has a row in poscounts even when its counts sum to zero (observed = 0, not NA)
    }
}

// nobs = 4 cuts(i) cuts(j), halved on the diagonal; cells without a read row (nobs holds the presence flag on entry):
// observed = nobs = 0; then the corner read
__global__ void k_fill_nobs(int n, int64_t ncells, const int *__restrict__ cuts, int32_t *__restrict__ observed, int32_t *__restrict__ nobs) {
    const int i = blockIdx.y, d = blockIdx.z;
    const int64_t base = (int64_t)d * ncells + rm_start(i, n);
    const long long ci = cuts[(int64_t)d * n + i];
    for (int j = i + blockIdx.x * blockDim.x + threadIdx.x; j < n; j += gridDim.x * blockDim.x) {
        const int64_t gidx = base + (j - i);
        long long no = 4 * ci * (long long)cuts[(int64_t)d * n + j];
        if (i == j) no /= 2;
        if (no > 2147483647LL) no = 2147483647LL;
        const int ob = observed[gidx];
        nobs[gidx] = nobs[gidx] == 0 ? 0 : (int)no;
        if (i == 0 && j == n - 1) observed[gidx] = ob + 1;   // counts[bin1==min(bin1)&bin2==max(bin2), observed := observed+1]
    }
}

} // namespace

void Model::ingest_reads(const fb_reads *r, int *nbins_out) {
    if (!r || !nbins_out) throw ArgError("null argument");
    if (r->ndatasets < 1 || r->resolution < 1 || r->ncuts < 1 || !r->cut_name || !r->cut_pos) throw ArgError("empty input");
    if (r->nreads > 0 && (!r->rd_name || !r->rd_pos1 || !r->rd_pos2 || !r->contact_close || !r->contact_far || !r->contact_up || !r->contact_down))
        throw ArgError("missing input column");
    int mn = r->cut_pos[0], mx = r->cut_pos[0];
    for (int64_t k = 0; k < r->ncuts; ++k) { mn = std::min(mn, r->cut_pos[k]); mx = std::max(mx, r->cut_pos[k]); }
    BinGeom g;
    g.res = r->resolution; g.b0 = (long long)mn - 1;
    const long long nb = ((long long)mx + 1 - g.b0) / g.res;     // seq(min-1, max+1, resolution), last incomplete bin discarded
    if (nb < 1) throw ArgError("no complete bin at this resolution");
    if (nb > 65535) throw ArgError("at most 65535 bins are supported");
    g.n = (int)nb; g.top = g.b0 + nb * g.res;
    const int n = g.n, D = r->ndatasets;
    const int64_t ncells = (int64_t)n * (n + 1) / 2, N = (int64_t)D * ncells;
    t_.n = n; t_.D = D; t_.ncells = ncells;
    obs_.alloc(N); nobs_.alloc(N);
    obs_.zero(st_); nobs_.zero(st_);
    DevBuf<int> cuts((size_t)D * n + 1);
    cuts.zero(st_);
    int *bad = cuts.p + (size_t)D * n;
    {
        DevBuf<int32_t> cn(r->ncuts), cp(r->ncuts);
        cn.upload(r->cut_name, r->ncuts, st_); cp.upload(r->cut_pos, r->ncuts, st_);
        k_cut_hist<<<(unsigned)std::min<int64_t>(1184, (r->ncuts + 255) / 256), 256, 0, st_>>>(g, D, r->ncuts, cn.p, cp.p, cuts.p, bad);
        FB_CUDA(cudaStreamSynchronize(st_));
    }
    // reads in slabs so that the staging buffers stay small next to the matrix
    const int64_t SLAB = int64_t(1) << 24;
    if (r->nreads > 0) {
        const int64_t cap = std::min<int64_t>(SLAB, r->nreads);
        DevBuf<int32_t> buf[7];
        for (auto &b : buf) b.alloc(cap);
        const int32_t *src[7] = {r->rd_name, r->rd_pos1, r->rd_pos2, r->contact_close, r->contact_far, r->contact_up, r->contact_down};
        for (int64_t o = 0; o < r->nreads; o += cap) {
            const int64_t cnt = std::min<int64_t>(cap, r->nreads - o);
            for (int q = 0; q < 7; ++q) buf[q].upload(src[q] + o, cnt, st_);
            k_read_hist<<<(unsigned)std::min<int64_t>(2368, (cnt + 255) / 256), 256, 0, st_>>>(g, D, ncells, cnt, buf[0].p, buf[1].p, buf[2].p,
                                                                                            buf[3].p, buf[4].p, buf[5].p, buf[6].p, obs_.p, nobs_.p, bad);
            FB_CUDA(cudaStreamSynchronize(st_));   // the host arrays are pageable: the next slab reuses the staging buffers
        }
    }
    dim3 grid((unsigned)std::max(1, std::min(8, (n + 255) / 256)), n, D);
    k_fill_nobs<<<grid, 256, 0, st_>>>(n, ncells, cuts.p, obs_.p, nobs_.p);
    int hbad = 0;
    FB_CUDA(cudaMemcpyAsync(&hbad, bad, sizeof(int), cudaMemcpyDeviceToHost, st_));
    FB_CUDA(cudaStreamSynchronize(st_));
    FB_LAUNCH_CHECK();
    kernel_launches += 3;
    if (hbad) throw ArgError("Name column values should range from 1 to the number of datasets");
    pos_.resize(n); dist_.resize(n);
    for (int k = 0; k < n; ++k) {
        const long long beg = g.b0 + (long long)k * g.res;
        pos_[k] = (double)((beg + beg + g.res) / 2);                      // as.integer((end + begin) / 2)
        dist_[k] = (double)((long long)k * g.res) + (double)g.res / 2.0;  // begin2 - begin1 + resolution / 2
    }
    alloc_work();
    *nbins_out = n;
}

void Model::get_mat(int32_t *name, int32_t *bin1, int32_t *pos1, int32_t *bin2, int32_t *pos2, double *distance,
                    int32_t *observed, int32_t *nobs) {
    const int n = t_.n, D = t_.D;
    const int64_t N = (int64_t)D * t_.ncells;
    if (observed) FB_CUDA(cudaMemcpyAsync(observed, obs_.p, N * sizeof(int32_t), cudaMemcpyDeviceToHost, st_));
    if (nobs) FB_CUDA(cudaMemcpyAsync(nobs, nobs_.p, N * sizeof(int32_t), cudaMemcpyDeviceToHost, st_));
    int64_t g = 0;
    for (int d = 0; d < D; ++d)
        for (int i = 0; i < n; ++i)
            for (int j = i; j < n; ++j, ++g) {
                if (name) name[g] = d + 1;
                if (bin1) bin1[g] = i + 1;
                if (bin2) bin2[g] = j + 1;
                if (pos1) pos1[g] = (int32_t)pos_[i];
                if (pos2) pos2[g] = (int32_t)pos_[j];
                if (distance) distance[g] = dist_[j - i];
            }
    FB_CUDA(cudaStreamSynchronize(st_));
}

} // namespace fb
