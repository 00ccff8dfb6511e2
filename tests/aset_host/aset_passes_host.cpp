// aset_passes_host.cpp -- TEST INFRASTRUCTURE. Runs the device passes of k_prox_aset (the very source the kernel
// compiles: binless_b200/csrc/fb_prox_aset_passes.cuh) on the host, one "thread" after the other in the order of the
// kernel's phases, for a batch of chains laid out as the plan builder lays them out. Built with -fsanitize=address by
// tests/test_aset_passes_host.py, so an index that leaves the shared-memory arrays is reported.
#include "shim.h"
#include "../../binless_b200/csrc/fb_prox_aset_passes.cuh"
#include <cstdlib>
#include <initializer_list>
#include <vector>

using namespace fb;
using namespace fb::aset;

// nch chains, chain c = v[off[c] .. off[c+1]); z_prev likewise; out likewise. Returns passes (<0: fell back).
extern "C" int aset_passes_host(int nch, const int *off, const double *v, const double *zprev, double lam, double *out,
                                int maxpass, int cap_min) {
    // layout
    std::vector<int> base(nch), L(nch);
    int npts = 0;
    std::vector<AsWord> words;
    for (int c = 0; c < nch; ++c) {
        L[c] = off[c + 1] - off[c]; base[c] = npts;
        const int need = ((L[c] + 1 + 31) / 32) * 32;
        for (int j = 0; j < need / 32; ++j) {
            AsWord wd;
            wd.goff = off[c] + 32 * (int64_t)j;
            wd.nvalid = (short)std::max(0, std::min(32, L[c] - 32 * j));
            wd.cidx = (short)c;
            wd.endbit = (short)((L[c] >= 32 * j && L[c] < 32 * j + 32) ? L[c] - 32 * j : -1);
            wd.jw = (short)j;
            words.push_back(wd);
        }
        npts += need;
    }
    const int cap = std::max(cap_min, npts), nwu = npts / 32;
    // every array on the heap, exactly as long as the kernel's (ASan guards the ends of each)
    const int nwd = cap / 32;
    struct SmH {
        std::vector<double> S, SLP, R, Cc, WOFF; std::vector<AsWord> WD;
        std::vector<unsigned> PM, K, UP, FIXM, DIRTY, REM, CH; std::vector<int> Lc, Bc, Qc; std::vector<unsigned short> WPREV, WFL; std::vector<unsigned char> ENDB;
        unsigned TW[2][AS_TWN];
    } s;
    s.S.assign(cap, 1e300); s.SLP.assign(cap, 1e300); s.R.assign(AS_RT, 0.0); s.Cc.assign(nwd, 0.0); s.WOFF.assign(nwd, 1e300);
    s.WD.resize(nwd); s.PM.assign(cap, 0xCDCDCDCDu);
    for (auto *m : {&s.K, &s.UP, &s.FIXM, &s.DIRTY, &s.REM, &s.CH}) m->assign(nwd, 0xCDCDCDCDu);
    s.Lc.assign(nwd, 0); s.Bc.assign(nwd, 0); s.Qc.assign(nwd, 0); s.WPREV.assign(nwd, 0xCDCD); s.WFL.assign(nwd, 0xCDCD); s.ENDB.assign(nwd, 0xCD);
    for (int m = 0; m < AS_RT; ++m) s.R[m] = m ? 1.0 / (double)m : 0.0;
    for (int c = 0; c < nch; ++c) { s.Lc[c] = L[c]; s.Bc[c] = base[c]; s.Qc[c] = c; }
    for (int g = 0; g < nwu; ++g) s.WD[g] = words[g];
    // A1 + A2
    for (int g = 0; g < nwu; ++g) {
        const AsWord d = s.WD[g];
        unsigned kb = 0, ub = 0;
        for (int lane = 0; lane < 32; ++lane) {
            s.PM[(g << 5) + lane] = 0u;
            if (lane >= d.nvalid) continue;
            const int64_t o = d.goff + lane;
            s.S[(g << 5) + lane + 1] = v[o];
            const bool head = d.jw == 0 && lane == 0;
            if (head) s.Cc[d.cidx] = v[o];
            if (!head && zprev[o] != zprev[o - 1]) { kb |= 1u << lane; if (zprev[o] > zprev[o - 1]) ub |= 1u << lane; }
        }
        s.K[g] = kb | fix_bits(d); s.UP[g] = ub; s.FIXM[g] = fix_bits(d); s.CH[g] = 0u; s.ENDB[g] = (unsigned char)(d.endbit >= 0 ? d.endbit : 32);
    }
    for (int c = 0; c < nch; ++c) {
        double *S = s.S.data() + base[c]; const double cc = s.Cc[c];
        double run = 0.0;
        for (int k = 0; k < L[c]; ++k) { run += S[k + 1] - cc; S[k + 1] = run; }
        S[0] = 0.0;
    }
    int npass = 0, fell = 0;
    if (lam != 0.0) {
        for (int k = 0; k < AS_TWN; ++k) {
            const int left = nwu - (k << 5);
            s.TW[0][k] = left >= 32 ? 0xffffffffu : left > 0 ? (1u << left) - 1u : 0u;
            s.TW[1][k] = 0u;
        }
        bool first = true;
        int cur = 0;
        auto active = [&](int w) { return ((s.TW[cur][w >> 5] >> (w & 31)) & 1u) != 0u; };
        for (;;) {
            if (npass >= maxpass) { fell = 1; break; }
            ++npass;
            for (int w = 0; w < nwu; ++w) if (active(w)) pass_slopes(s, w, lam, first);
            for (int w = 0; w < nwu; ++w) if (active(w)) pass_bends(s, w);
            for (int g = 0; g < nwu; ++g) if (active(g)) for (int lane = 0; lane < 32; ++lane) pass_points(s, g, lam, lane, (1u << lane) - 1u);
            for (int w = nwu - 1; w >= 0; --w) if (active(w)) pass_apply(s, w, cur ^ 1);      // any order must do
            unsigned any = 0u;
            for (int k = 0; k < AS_TWN; ++k) { any |= s.TW[cur ^ 1][k]; s.TW[cur][k] = 0u; }
            if (!any) break;
            cur ^= 1;
            first = false;
        }
        if (fell) {
            for (int c = 0; c < nch; ++c) chain_sequential(s, base[c], L[c], lam);
            for (int w = 0; w < nwu; ++w) pass_slopes(s, w, lam, true);
        }
    }
    for (int g = 0; g < nwu; ++g) {
        const AsWord d = s.WD[g];
        for (int lane = 0; lane < d.nvalid; ++lane) {
            double z = v[d.goff + lane];
            if (lam != 0.0) {
                const unsigned lm = s.K[g] & (0xffffffffu >> (31 - lane));
                const int pa = lm ? (g << 5) + 31 - __clz(lm) : (int)s.WPREV[g];
                z = s.Cc[d.cidx] + s.SLP[pa];
            }
            out[d.goff + lane] = z;
        }
    }
    return fell ? -npass - 1 : npass;
}
