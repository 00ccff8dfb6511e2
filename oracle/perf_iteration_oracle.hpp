/* perf_iteration_oracle.hpp -- TEST INFRASTRUCTURE (CPU oracle), included by fast_binless_oracle.cpp.
 *
 * Restatement of the optimized path's performance iteration on the triangle-grid fused lasso,
 * paths relative to /root/reference/binless/src:
 *   wgfl_signal_BIC            perf_iteration_signal.cpp:22-157
 *   wgfl_diff_BIC              perf_iteration_diff.cpp:20-117
 *   cts_to_mat<Signal/Diff>    cts_to_mat.cpp:14-124   (remove_outliers :14-32)
 *   cts_to_signal_mat_core     cts_core.c:5-47         (the reference's own object when built with ORACLE_USE_REFERENCE_GFL,
 *                                                        else the restatement below; tests require them equal bit for bit)
 *   IRLSEstimator::optimize    IRLSEstimator.hpp:28-60
 *   WeightsUpdater             WeightsUpdater.hpp:17-92
 *   compute_phi_ref            util.cpp:20-41
 *   EstimatedSparsity          EstimatedSparsity.ipp:11-30;  FixedSparsity.hpp:35-45;  BoundsRectifier.hpp
 *   candidates                 candidate_generation.cpp:8-44, candidate_generation.hpp:27-40, CandidatesValues.hpp,
 *                              minimum_UB.cpp:6-40, minimum_UB.hpp:25-33
 *   BoundsOptimizer            BoundsOptimizer.ipp:5-52
 *   ScorePreparator<BIC>       ScorePreparator.ipp:53-64 (cold-start re-solve)
 *   Likelihood                 Likelihoods.hpp:25-95;  DOFComputer.hpp:16-24;  ScoreAssembler<BIC> ScoreAssembler.hpp:41-55
 *   ScoreOptimizer<BIC>        ScoreOptimizer.hpp:13-21 (k = 0: the first minimum)
 * Parity: cts_core.c pinned by the reference object; the C++ around it needs Rcpp sugar and is restated (unpinned).
 */
#pragma once

#ifdef ORACLE_USE_REFERENCE_GFL
extern "C" void cts_to_signal_mat_core(int N, int *cts_bin1, int *cts_bin2, double *count, double *lmu_nosig, double *nobs,
                                       double *log_decay, int nbins, double dispersion, double *phi, double *phihat,
                                       double *phihat_var, double *phihat_var_nodecay, double *ncounts, int *bin1, int *bin2);
#define ORACLE_CTS_CORE cts_to_signal_mat_core
#else
#define ORACLE_CTS_CORE ctsp_signal_mat_core
#endif

namespace {

/* cts_core.c:5-47, statement by statement */
void ctsp_signal_mat_core(int N, int *cb1, int *cb2, double *count, double *lmu_nosig, double *nobs, double *log_decay,
                          int nbins, double dispersion, double *phi, double *phihat, double *phihat_var,
                          double *phihat_var_nodecay, double *ncounts, int *bin1, int *bin2) {
    const int nbetas = nbins * (nbins + 1) / 2;
    for (int i = 0; i < N; ++i) {
        const int b1 = cb1[i], b2 = cb2[i];
        const int pos = (b1 - 1) * (nbins + 1) - (b1 - 1) * b1 / 2 + b2 - b1;
        const double mu = exp(lmu_nosig[i] + phi[pos]);
        const double mu_nodecay = exp(lmu_nosig[i] - log_decay[i] + phi[pos]);
        const double z = count[i] / mu - 1;
        const double var = 1 / mu + 1 / dispersion;
        const double var_nodecay = 1 / mu_nodecay + 1 / dispersion;
        const double wt = nobs[i] / var;
        const double wt_nodecay = nobs[i] / var_nodecay;
        ncounts[pos] += nobs[i];
        phihat_var[pos] += wt;
        phihat_var_nodecay[pos] += wt_nodecay;
        phihat[pos] += (z + phi[pos]) * wt;
    }
    int b1 = 1, b2 = 1;
    for (int i = 0; i < nbetas; ++i) {
        phihat_var[i] = 1 / phihat_var[i];
        phihat_var_nodecay[i] = 1 / phihat_var_nodecay[i];
        if (ncounts[i] > 0) phihat[i] = phihat[i] * phihat_var[i];
        else phihat[i] = 0;
        bin1[i] = b1;
        bin2[i] = b2++;
        if (b2 > nbins) b2 = ++b1;
    }
}

struct PerfCts { ivec bin1, bin2; vec count, lmu_nosig, nobs, log_decay; };
PerfCts perf_copy_cts(const fb_cts *c) {
    if (!c || c->N < 0 || (c->N > 0 && (!c->bin1 || !c->bin2 || !c->count || !c->lmu_nosig || !c->nobs || !c->log_decay)))
        throw ArgError("missing cts column");
    PerfCts r; const int64_t n = c->N;
    r.bin1.assign(c->bin1, c->bin1 + n); r.bin2.assign(c->bin2, c->bin2 + n);
    r.count.assign(c->count, c->count + n); r.lmu_nosig.assign(c->lmu_nosig, c->lmu_nosig + n);
    r.nobs.assign(c->nobs, c->nobs + n); r.log_decay.assign(c->log_decay, c->log_decay + n);
    return r;
}

/* BinnedData<Signal> / <Difference> (BinnedData.hpp) */
struct PerfBinned {
    ivec bin1, bin2, diag_idx, diag_grp, patchno;
    vec beta, betahat, weight, weight_nodecay, nobs;
    vec weight_ref, weight_nodecay_ref, phihat_ref, phi_ref;   /* Difference */
};

struct PerfMeta { std::set<int> bad_rows, bad_diags; ivec diag_grp; };

/* cts_to_mat<Signal> (cts_to_mat.cpp:34-102) with phi = binned.beta */
void perf_cts_to_mat_signal(int nbins, double dispersion, PerfCts &cts, const PerfMeta &meta, PerfBinned &binned) {
    const int nbetas = nbins * (nbins + 1) / 2;
    vec phi = binned.beta;
    vec phihat(nbetas, 0), phihat_var(nbetas, 0), phihat_var_nodecay(nbetas, 0), ncounts(nbetas, 0);
    ivec bin1(nbetas, 0), bin2(nbetas, 0);
    ORACLE_CTS_CORE((int)cts.bin1.size(), cts.bin1.data(), cts.bin2.data(), cts.count.data(), cts.lmu_nosig.data(),
                    cts.nobs.data(), cts.log_decay.data(), nbins, dispersion, phi.data(), phihat.data(), phihat_var.data(),
                    phihat_var_nodecay.data(), ncounts.data(), bin1.data(), bin2.data());
    /* remove_outliers (:14-32) */
    for (int i = 0; i < nbetas; ++i)
        if (meta.bad_rows.count(bin1[i]) || (bin2[i] != bin1[i] && meta.bad_rows.count(bin2[i]))) phihat_var[i] = INFINITY;
    for (int i = 0; i < nbetas; ++i)
        if (meta.bad_diags.count(bin2[i] - bin1[i])) phihat_var[i] = INFINITY;
    binned.bin1 = bin1; binned.bin2 = bin2;
    binned.betahat = phihat;
    binned.weight.resize(nbetas); binned.weight_nodecay.resize(nbetas);
    binned.diag_idx.resize(nbetas); binned.diag_grp.resize(nbetas);
    for (int i = 0; i < nbetas; ++i) {
        binned.weight[i] = 1 / phihat_var[i];
        binned.weight_nodecay[i] = 1 / phihat_var_nodecay[i];
        binned.diag_idx[i] = bin2[i] - bin1[i];
        binned.diag_grp[i] = meta.diag_grp[binned.diag_idx[i]];
    }
    binned.nobs = ncounts;
}

/* compute_phi_ref (util.cpp:20-41); phihat = phi_ref + deltahat (BinnedData.hpp:76) */
vec perf_compute_phi_ref(const PerfBinned &b, const vec &delta) {
    const size_t N = delta.size();
    vec r(N);
    for (size_t i = 0; i < N; ++i) {
        const double phihat = b.phi_ref[i] + b.betahat[i];
        const double var = 1 / b.weight[i], var_ref = 1 / b.weight_ref[i];
        double val;
        if (var_ref == INFINITY && var == INFINITY) val = (b.phihat_ref[i] + phihat) / 2;
        else val = (b.phihat_ref[i] / var_ref + (phihat - delta[i]) / var) / (1 / var_ref + 1 / var);
        r[i] = std::max(val, 0.);
    }
    return r;
}

/* cts_to_mat<Difference> (cts_to_mat.cpp:104-124) */
void perf_cts_to_mat_diff(int nbins, double dispersion, PerfCts &cts, PerfCts &ref, const PerfMeta &meta, PerfBinned &binned) {
    PerfBinned sref, soth;
    sref.beta = binned.phi_ref;
    perf_cts_to_mat_signal(nbins, dispersion, ref, meta, sref);
    soth.beta.resize(binned.phi_ref.size());
    for (size_t i = 0; i < soth.beta.size(); ++i) soth.beta[i] = binned.phi_ref[i] + binned.beta[i];
    perf_cts_to_mat_signal(nbins, dispersion, cts, meta, soth);
    binned.bin1 = sref.bin1; binned.bin2 = sref.bin2;
    binned.weight = soth.weight; binned.weight_nodecay = soth.weight_nodecay;
    binned.betahat.resize(soth.betahat.size());
    binned.nobs.resize(soth.betahat.size());
    for (size_t i = 0; i < soth.betahat.size(); ++i) {
        binned.betahat[i] = soth.betahat[i] - binned.phi_ref[i];
        binned.nobs[i] = soth.nobs[i] + sref.nobs[i];
    }
    binned.diag_idx = sref.diag_idx; binned.diag_grp = sref.diag_grp;
    binned.weight_ref = sref.weight; binned.weight_nodecay_ref = sref.weight_nodecay; binned.phihat_ref = sref.betahat;
}

vec perf_soft(const vec &beta, double eCprime, double lam1) {   /* util.cpp:8-18 */
    vec phi(beta.size());
    for (size_t i = 0; i < beta.size(); ++i) {
        const double val = beta[i] - eCprime;
        phi[i] = (val > 0) ? std::max(0., val - lam1) : std::min(0., val + lam1);
    }
    return phi;
}

struct PerfScore { double eCprime = 0, lambda1 = 0, BIC = 0, BIC_sd = -1, UB = 0, LB = 0; int dof = 0; };

/* the four policy tags of EstimatedSparsity / FixedSparsity */
struct PerfPolicy { bool zero_offset, positive, forbid_degeneracy, difference; };

struct PerfSparsity {
    int nbins; double tol_val; const PerfBinned &binned; PerfPolicy pol;
    vec beta_cold;      /* ScorePreparator<BIC>::beta_ */
    double lsnc;        /* ScoreAssembler<BIC>: log(sum(nobs)) */
    int ncandidates = 0;

    PerfSparsity(int nb, double tol, const PerfBinned &b, PerfPolicy p, GflSolver &flo, double lam2, int64_t &admm)
        : nbins(nb), tol_val(tol), binned(b), pol(p) {
        flo.reset();                                   /* ScorePreparator.ipp:61-63; reset() keeps alpha (GFLLibraryTriangleImpl.cpp:52-58) */
        flo.optimize(b.betahat, b.weight, lam2, tol / 20.);
        admm += flo.last_steps - 1;
        beta_cold = flo.get();
        double s = 0; for (double v : b.nobs) s += v;
        lsnc = std::log(s);
    }
    vec values() const {                               /* CandidatesValues.hpp */
        vec v = binned.beta;
        if (pol.zero_offset) for (auto &x : v) x = std::fabs(x);
        return v;
    }
    double minimum_UB() const {                        /* minimum_UB.hpp:25-33, minimum_UB.cpp:6-40 */
        const vec v = values();
        const double minbeta = *std::min_element(binned.beta.begin(), binned.beta.end());
        double minUBo;
        if (pol.zero_offset) minUBo = pol.positive ? std::max(-minbeta, 0.) : 0.;
        else minUBo = minbeta;
        double minUBd;
        if (!pol.forbid_degeneracy) minUBd = *std::min_element(v.begin(), v.end()) - 1;
        else {
            const int ndiags = *std::max_element(binned.diag_grp.begin(), binned.diag_grp.end()) + 1;
            std::vector<char> touch(ndiags, 0);
            vec diagvals(ndiags, *std::max_element(v.begin(), v.end()));
            for (size_t i = 0; i < binned.diag_grp.size(); ++i) {
                diagvals[binned.diag_grp[i]] = std::min(diagvals[binned.diag_grp[i]], v[i]);
                touch[binned.diag_grp[i]] = 1;
            }
            double m = -INFINITY;
            for (int g = 0; g < ndiags; ++g) if (touch[g]) m = std::max(m, diagvals[g]);
            minUBd = m;
        }
        return std::max(minUBd, minUBo);
    }
    vec candidates() const {                           /* candidate_generation.hpp:27-40 */
        const double minUB = minimum_UB();
        const vec v = values();
        std::vector<int32_t> patchno(v.size());
        patch_numbers(nbins, tol_val, v.data(), patchno.data());
        const int npatches = *std::max_element(patchno.begin(), patchno.end()) + 1;
        vec borders(npatches);
        for (size_t i = 0; i < v.size(); ++i) borders[patchno[i]] = v[i];
        std::sort(borders.begin(), borders.end());
        vec ret; ret.push_back(minUB);                 /* cleanup_borders */
        bool any = false; double last = 0;
        for (double b : borders) if (b >= minUB) { ret.push_back(b); any = true; last = b; }
        ret.push_back(any ? last + 1 : minUB + 1);
        vec cand;                                      /* candidates_from_borders */
        for (size_t i = 0; i + 1 < ret.size(); ++i) cand.push_back((ret[i] + ret[i + 1]) / 2);
        return cand;
    }
    std::pair<double, double> optimize_bounds(double val) const {   /* BoundsOptimizer.ipp */
        const vec &beta = binned.beta, &w = binned.weight;
        if (!pol.zero_offset) {
            const double xmin = *std::min_element(beta.begin(), beta.end()), xmax = *std::max_element(beta.begin(), beta.end());
            double s1 = 0, s2 = 0, mx1 = -INFINITY; size_t n1 = 0, n2 = 0;
            for (size_t i = 0; i < beta.size(); ++i) {
                if (beta[i] <= val) { s1 += w[i]; ++n1; mx1 = std::max(mx1, beta[i]); } else { s2 += w[i]; ++n2; }
            }
            const bool grp2empty = n2 == 0 || s2 == 0, grp1empty = n1 == 0 || s1 == 0;
            double UB, LB;
            if (!grp2empty && !grp1empty) { UB = mx1; LB = xmin; }
            else if (grp2empty) { UB = xmax; LB = xmin; }
            else { UB = xmin; LB = xmin; }
            if (UB < LB) throw NumError("Found LB > UB!");
            return {LB, UB};
        }
        double mn = INFINITY, mx = -INFINITY, xk = -INFINITY;
        for (double b : beta) { const double a = std::fabs(b); mn = std::min(mn, a); mx = std::max(mx, a); if (a <= val) xk = std::max(xk, a); }
        if (val > mx) xk = mx; else if (val < mn) xk = 0;
        return {-xk, xk};
    }
    PerfScore evaluate(double LB, double UB) const {    /* ScoreComputer.ipp:24-41 */
        const double lambda1 = (UB - LB) / 2, eCprime = (UB + LB) / 2.;
        const size_t N = binned.beta.size();
        const double logpri = lambda1 * lambda1;
        double sum = 0;
        if (!pol.difference) {                         /* Likelihoods.hpp:25-58 */
            const vec phi = perf_soft(beta_cold, eCprime, lambda1);
            for (size_t i = 0; i < N; ++i) {
                const double d = binned.betahat[i] - (phi[i] + eCprime);
                sum += binned.weight_nodecay[i] * (d * d) + logpri;
            }
        } else {                                       /* Likelihoods.hpp:61-95 */
            const vec delta = perf_soft(beta_cold, eCprime, lambda1);
            const vec phi_ref = perf_compute_phi_ref(binned, delta);
            for (size_t i = 0; i < N; ++i) {
                const double phihat = binned.phi_ref[i] + binned.betahat[i];
                const double d = phihat - (delta[i] + eCprime + phi_ref[i]), dr = binned.phihat_ref[i] - phi_ref[i];
                sum += binned.weight_nodecay[i] * (d * d) + binned.weight_nodecay_ref[i] * (dr * dr) + logpri;
            }
        }
        const vec soft = perf_soft(binned.beta, eCprime, lambda1);     /* DOFComputer.hpp:16-24 */
        std::set<int> sel;
        for (size_t i = 0; i < N; ++i) if (std::fabs(soft[i]) > tol_val / 2) sel.insert(binned.patchno[i]);
        PerfScore s;
        s.eCprime = eCprime; s.lambda1 = lambda1; s.dof = (int)sel.size();
        s.BIC = sum + lsnc * s.dof; s.BIC_sd = -1; s.UB = UB; s.LB = LB;
        return s;
    }
    PerfScore estimated() {                            /* EstimatedSparsity.ipp:11-30 */
        const vec cand = candidates();
        ncandidates = (int)cand.size();
        std::vector<PerfScore> values;
        for (double c : cand) {
            auto b = optimize_bounds(c);
            const double lambda1 = (b.second - b.first) / 2.;
            if (lambda1 < tol_val && c < cand.back() - 1e-12) continue;
            values.push_back(evaluate(b.first, b.second));
        }
        if (values.empty()) throw NumError("Found 0 values!");
        size_t best = 0;
        for (size_t k = 1; k < values.size(); ++k) if (values[k].BIC < values[best].BIC) best = k;
        return values[best];
    }
    PerfScore fixed(double lambda1_candidate) {        /* FixedSparsity.hpp:35-45, BoundsRectifier.hpp */
        const double minUB = minimum_UB();
        ncandidates = 1;
        if (!pol.zero_offset) {
            const double LB = *std::min_element(binned.beta.begin(), binned.beta.end());
            double UB = LB + 2 * lambda1_candidate;
            if (minUB > UB) UB = minUB;
            return evaluate(LB, UB);
        }
        double UB = lambda1_candidate;
        if (minUB > lambda1_candidate) UB = minUB;
        return evaluate(-UB, UB);
    }
};

PerfMeta perf_meta(int nbins, const fb_perf_meta *m) {
    if (!m || !m->diag_grp) throw ArgError("metadata$diag.grp is required");
    PerfMeta r;
    for (int k = 0; k < m->n_bad_rows; ++k) r.bad_rows.insert(m->bad_rows[k]);
    for (int k = 0; k < m->n_bad_diags; ++k) r.bad_diags.insert(m->bad_diags[k]);
    r.diag_grp.assign(m->diag_grp, m->diag_grp + nbins);
    return r;
}

void perf_import_state(GflSolver &flo, int warm, const fb_gfl_state *st) {      /* set_state, GFLLibraryTriangleImpl.cpp:78-87 */
    if (warm && st && st->z && st->u && st->beta) {
        flo.z.assign(st->z, st->z + flo.tsz); flo.u.assign(st->u, st->u + flo.tsz);
        flo.beta_impl.assign(st->beta, st->beta + flo.N);
        flo.alpha = st->alpha;
    }
}
void perf_export_state(const GflSolver &flo, fb_gfl_state *st) {
    if (!st) return;
    st->alpha = flo.alpha; st->steps = flo.last_steps;
    if (st->beta) std::memcpy(st->beta, flo.beta_impl.data(), flo.N * sizeof(double));
    if (st->z) std::memcpy(st->z, flo.z.data(), flo.tsz * sizeof(double));
    if (st->u) std::memcpy(st->u, flo.u.data(), flo.tsz * sizeof(double));
}

/* IRLSEstimator::optimize (IRLSEstimator.hpp:28-60); update(beta) refreshes (y, w) from the counts */
template <class Update>
bool perf_irls(unsigned nouter, vec &beta, double lam2, double converge, GflSolver &flo, PerfBinned &binned, Update &&update,
               int &counter, int64_t &admm) {
    counter = 0;
    bool has_converged = false;
    vec beta_old = beta;
    do {
        update(beta);
        flo.optimize(binned.betahat, binned.weight, lam2, converge);
        admm += flo.last_steps - 1;
        beta = flo.get();
        double maxval = std::fabs(beta[0] - beta_old[0]);
        for (size_t i = 1; i < beta.size(); ++i) maxval = std::max(std::fabs(beta[i] - beta_old[i]), maxval);
        has_converged = maxval < converge;
        ++counter;
        beta_old = beta;
    } while ((unsigned)counter < nouter && !has_converged);
    return has_converged;
}

void perf_wgfl_signal(int nbins, const fb_cts *ctsp, const fb_perf_meta *metap, const fb_perf_params *p, const double *beta_init,
                      int warm, fb_gfl_state *gfl, fb_perf_signal_out *out) {
    if (!p || !out || !beta_init || nbins < 2) throw ArgError("null argument");
    PerfCts cts = perf_copy_cts(ctsp);
    const PerfMeta meta = perf_meta(nbins, metap);
    const int nbetas = nbins * (nbins + 1) / 2;
    PerfBinned binned;
    const double converge = p->tol_val / 20.;
    GflSolver flo(nbins, 0);
    perf_import_state(flo, warm, gfl);
    vec beta(beta_init, beta_init + nbetas);
    int64_t admm = 0;
    auto update = [&](const vec &b) { binned.beta = b; perf_cts_to_mat_signal(nbins, p->dispersion, cts, meta, binned); };
    unsigned nwarm, ncold;
    if (p->nouter == 1) nwarm = 1; else nwarm = p->nouter / 2;
    ncold = p->nouter - nwarm;
    int counter = 0;
    bool converged = perf_irls(nwarm, beta, p->lam2, converge, flo, binned, update, counter, admm);
    binned.beta = beta;
    if (ncold > 0 && !converged) {
        std::fill(beta.begin(), beta.end(), 0);
        flo.reset();
        converged = perf_irls(ncold, beta, p->lam2, converge, flo, binned, update, counter, admm);
        binned.beta = beta;
    }
    perf_export_state(flo, gfl);
    binned.patchno.resize(nbetas);
    patch_numbers(nbins, p->tol_val, binned.beta.data(), binned.patchno.data());
    const bool zero_offset = p->fixed != 0, positive = p->fixed ? p->positive != 0 : true;
    PerfSparsity sp(nbins, p->tol_val, binned, PerfPolicy{zero_offset, positive, p->constrained != 0, false}, flo, p->lam2, admm);
    const PerfScore opt = p->lambda1_fixed ? sp.fixed(p->lambda1_fix_value) : sp.estimated();
    const vec phi = perf_soft(binned.beta, opt.eCprime, opt.lambda1);
    copy_out(out->phi, phi); copy_out(out->beta, binned.beta); copy_out(out->phihat, binned.betahat);
    copy_out(out->nobs, binned.nobs); copy_out(out->weight, binned.weight);
    if (out->patchno) std::copy(binned.patchno.begin(), binned.patchno.end(), out->patchno);
    out->dof = opt.dof; out->converged = converged; out->nouter_done = counter; out->ncandidates = sp.ncandidates;
    out->BIC = opt.BIC; out->BIC_sd = opt.BIC_sd; out->eCprime = opt.eCprime; out->lambda1 = opt.lambda1;
    out->admm_iters = admm;
}

void perf_wgfl_diff(int nbins, const fb_cts *ctsp, const fb_cts *refp, const fb_perf_meta *metap, const fb_perf_params *p,
                    const double *phi_ref_init, const double *beta_init, int warm, fb_gfl_state *gfl, fb_perf_diff_out *out) {
    if (!p || !out || !beta_init || !phi_ref_init || nbins < 2) throw ArgError("null argument");
    PerfCts cts = perf_copy_cts(ctsp), ref = perf_copy_cts(refp);
    const PerfMeta meta = perf_meta(nbins, metap);
    const int nbetas = nbins * (nbins + 1) / 2;
    PerfBinned binned;
    const double converge = p->tol_val / 20.;
    GflSolver flo(nbins, 0);
    perf_import_state(flo, warm, gfl);
    vec beta(beta_init, beta_init + nbetas);
    int64_t admm = 0;
    /* WeightsUpdater<Difference>::setUp (WeightsUpdater.hpp:57-62) */
    binned.beta = beta; binned.phi_ref.assign(phi_ref_init, phi_ref_init + nbetas);
    perf_cts_to_mat_diff(nbins, p->dispersion, cts, ref, meta, binned);
    auto set_beta_delta = [&](const vec &b) { binned.beta = b; binned.phi_ref = perf_compute_phi_ref(binned, binned.beta); };
    auto update = [&](const vec &b) { set_beta_delta(b); perf_cts_to_mat_diff(nbins, p->dispersion, cts, ref, meta, binned); };
    int counter = 0;
    bool converged = perf_irls(p->nouter, beta, p->lam2, converge, flo, binned, update, counter, admm);
    set_beta_delta(beta);
    if (!converged) {
        std::fill(beta.begin(), beta.end(), 0);
        flo.reset();
        converged = perf_irls(p->nouter, beta, p->lam2, converge, flo, binned, update, counter, admm);
        set_beta_delta(beta);
    }
    perf_export_state(flo, gfl);
    binned.patchno.resize(nbetas);
    patch_numbers(nbins, p->tol_val, binned.beta.data(), binned.patchno.data());
    PerfSparsity sp(nbins, p->tol_val, binned, PerfPolicy{true, false, p->constrained != 0, true}, flo, p->lam2, admm);
    const PerfScore opt = p->lambda1_fixed ? sp.fixed(p->lambda1_fix_value) : sp.estimated();
    const vec delta = perf_soft(binned.beta, 0, opt.lambda1);
    const vec phi_ref_r = perf_compute_phi_ref(binned, delta);
    vec phihat(nbetas), var(nbetas), var_ref(nbetas);
    for (int i = 0; i < nbetas; ++i) {
        phihat[i] = binned.phi_ref[i] + binned.betahat[i]; var[i] = 1 / binned.weight[i]; var_ref[i] = 1 / binned.weight_ref[i];
    }
    copy_out(out->phi_ref, phi_ref_r); copy_out(out->delta, delta); copy_out(out->beta, binned.beta);
    copy_out(out->phihat, phihat); copy_out(out->phihat_var, var); copy_out(out->phihat_ref, binned.phihat_ref);
    copy_out(out->phihat_var_ref, var_ref); copy_out(out->nobs, binned.nobs); copy_out(out->deltahat, binned.betahat);
    copy_out(out->weight, binned.weight);
    if (out->patchno) std::copy(binned.patchno.begin(), binned.patchno.end(), out->patchno);
    out->dof = opt.dof; out->converged = converged; out->nouter_done = counter; out->ncandidates = sp.ncandidates;
    out->BIC = opt.BIC; out->BIC_sd = opt.BIC_sd; out->lambda1 = opt.lambda1;
    out->admm_iters = admm;
}

} // namespace
