// fb_perf.cuh -- optimized-path performance iteration (wgfl_signal_BIC / wgfl_diff_BIC) on the device; see fb_perf.cu
#pragma once
#include "../../include/fastbinless.h"

namespace fb {
void wgfl_signal_bic(int nbins, const fb_cts *cts, const fb_perf_meta *meta, const fb_perf_params *p, const double *beta_init,
                     int warm, fb_gfl_state *gfl, fb_perf_signal_out *out);
void wgfl_diff_bic(int nbins, const fb_cts *cts, const fb_cts *ref, const fb_perf_meta *meta, const fb_perf_params *p,
                   const double *phi_ref_init, const double *beta_init, int warm, fb_gfl_state *gfl, fb_perf_diff_out *out);
} // namespace fb
