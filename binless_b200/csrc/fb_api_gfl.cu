// fb_api_gfl.cu -- C-ABI entry point for the stand-alone fused-lasso solve (include/fastbinless.h: fb_gfl_solve).
#include "../../include/fastbinless.h"
#include "fb_gfl.cuh"
#include <string>
#include <vector>

namespace fb { std::string g_last_error; }

extern "C" FB_API int fb_gfl_solve(int nbins, unsigned maxdiag, const double *y, const double *w, double lam2,
                                   double converge, int maxsteps, int warm, fb_gfl_state *st, double *beta_full) {
    try {
        int ndev = 0;
        if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) throw fb::CudaError("no CUDA device available (there is no CPU fallback)");
        if (!y || !w || !st) throw fb::ArgError("null argument");
        const size_t nc = (size_t)nbins * (nbins + 1) / 2;
        fb::GflSolver s(nbins, maxdiag, 0);
        s.ninner_max = maxsteps;
        fb::DevBuf<double> dy(nc), dw(nc), dout(nc);
        dy.upload(y, nc); dw.upload(w, nc);
        if (warm) { s.alpha = st->alpha; s.import_state(st->beta, st->z, st->u); }
        s.optimize(dy.p, dw.p, lam2, converge);
        st->alpha = s.alpha; st->steps = s.last_steps;
        s.export_state(st->beta, st->z, st->u);
        if (beta_full) { s.get_beta_full(dout.p); FB_CUDA(cudaDeviceSynchronize()); dout.download(beta_full, nc); FB_CUDA(cudaDeviceSynchronize()); }
        return FB_OK;
    } catch (const fb::ArgError &e) { fb::g_last_error = e.what(); return FB_ERR_ARG;
    } catch (const fb::CudaError &e) { fb::g_last_error = e.what(); return FB_ERR_CUDA;
    } catch (const std::exception &e) { fb::g_last_error = e.what(); return FB_ERR_NUMERIC; }
}
