// The priors and the state of an RVmodel as plain data for the device-side sampler
// (include/kima_sampler.h): what kc_sampler_create / kc_sampler_upload take.
//
// Included AFTER the reference's headers have been included with the access keywords opened
// (host/overlay/kima.h, oracle/ref_sampler_harness.cpp), so that the private members of the
// distributions (DNest4/code/Distributions/*.h, src/distributions/Fixed.h), of RVmodel
// (src/RVmodel.h:36-104), RJObject (RJObject.h:27-58) and RVConditionalPrior
// (src/RVConditionalPrior.h:13-47) can be read.
#pragma once
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>
#include "kima_sampler.h"

extern const bool GP, MA, hyperpriors, trend, multi_instrument, known_object, studentt;
extern const int degree, n_known_object;

namespace kima_b200 {

// The reference's own constructors derive k[] here (host libm), through kc_prior_make when the
// library is linked, or through the caller's `make` (the oracle harness has no libkima_cuda).
using PriorMake = int (*)(int type, const double* p, int n_p, kc_prior* out);

// One ContinuousDistribution -> kc_prior; false when the class has no device counterpart
// (Empirical: a table of samples, and its log_pdf throws).  Jeffreys is an alias of LogUniform.
inline bool extract_prior(const DNest4::ContinuousDistribution* d, kc_prior& out, PriorMake make)
{
    using namespace DNest4;
    out = kc_prior{};
    out.type = KC_PRIOR_NONE;
    if (!d) return true;
    auto mk = [&](int type, std::initializer_list<double> p) {
        std::vector<double> v(p);
        return make(type, v.data(), (int)v.size(), &out) == 0;
    };
    if (auto* q = dynamic_cast<const Fixed*>(d)) return mk(KC_PRIOR_FIXED, {q->val});
    if (auto* q = dynamic_cast<const Uniform*>(d)) return mk(KC_PRIOR_UNIFORM, {q->lower, q->upper});
    if (auto* q = dynamic_cast<const LogUniform*>(d)) return mk(KC_PRIOR_LOGUNIFORM, {q->lower, q->upper});
    if (auto* q = dynamic_cast<const ModifiedLogUniform*>(d)) return mk(KC_PRIOR_MODLOGUNIFORM, {q->knee, q->upper});
    if (auto* q = dynamic_cast<const Gaussian*>(d)) return mk(KC_PRIOR_GAUSSIAN, {q->center, q->width});
    if (auto* q = dynamic_cast<const TruncatedGaussian*>(d))
        return mk(KC_PRIOR_TRUNC_GAUSSIAN, {q->center, q->width, q->lower, q->upper});
    if (auto* q = dynamic_cast<const Cauchy*>(d)) return mk(KC_PRIOR_CAUCHY, {q->center, q->width});
    if (auto* q = dynamic_cast<const TruncatedCauchy*>(d))
        return mk(KC_PRIOR_TRUNC_CAUCHY, {q->center, q->width, q->lower, q->upper});
    if (auto* q = dynamic_cast<const Exponential*>(d)) return mk(KC_PRIOR_EXPONENTIAL, {q->scale});
    if (auto* q = dynamic_cast<const TruncatedExponential*>(d))
        return mk(KC_PRIOR_TRUNC_EXPONENTIAL, {q->scale, q->lower, q->upper});
    if (auto* q = dynamic_cast<const Laplace*>(d)) return mk(KC_PRIOR_LAPLACE, {q->center, q->width});
    if (auto* q = dynamic_cast<const Kumaraswamy*>(d)) return mk(KC_PRIOR_KUMARASWAMY, {q->a, q->b});
    if (auto* q = dynamic_cast<const Rayleigh*>(d)) return mk(KC_PRIOR_RAYLEIGH, {q->scale});
    if (auto* q = dynamic_cast<const TruncatedRayleigh*>(d))
        return mk(KC_PRIOR_TRUNC_RAYLEIGH, {q->scale, q->lower, q->upper});
    if (auto* q = dynamic_cast<const Triangular*>(d)) return mk(KC_PRIOR_TRIANGULAR, {q->lower, q->centre, q->upper});
    if (auto* q = dynamic_cast<const Pareto*>(d)) return mk(KC_PRIOR_PARETO, {q->min, q->alpha});
    // kima's own classes (src/distributions/): the Gaussian for phi built from a transit time is the
    // Gaussian(m, s) it wraps (every virtual forwards to it; its perturb is not virtual and never reached
    // through the base pointer); the Gaussian + log-uniform mixture has its own table form
    if (auto* q = dynamic_cast<const Gaussian_from_Tc*>(d)) return mk(KC_PRIOR_GAUSSIAN, {q->m, q->s});
    if (auto* q = dynamic_cast<const mixGaussianLogUniform*>(d))
        return mk(KC_PRIOR_MIX_GAUSS_LOGUNIFORM, {q->mean, q->sigma, q->lower, q->upper});
    return false;
}

// Every prior of a model on which setPriors() has run (from_prior calls it, src/RVmodel.cpp:91).
// Returns the name of the first prior without a device counterpart, or "" when all map.
inline std::string extract_prior_table(const RVmodel& m, kc_prior_table& T, PriorMake make)
{
    T = kc_prior_table{};
    std::string bad;
    auto get = [&](const std::shared_ptr<DNest4::ContinuousDistribution>& p, kc_prior& out, const char* name) {
        if (!extract_prior(p.get(), out, make) && bad.empty()) bad = name;
    };
    get(m.Cprior, T.Cprior, "Cprior");
    get(m.Jprior, T.Jprior, "Jprior");
    get(m.slope_prior, T.slope_prior, "slope_prior");
    get(m.quadr_prior, T.quadr_prior, "quadr_prior");
    get(m.cubic_prior, T.cubic_prior, "cubic_prior");
    get(m.offsets_prior, T.offsets_prior, "offsets_prior");
    get(m.betaprior, T.betaprior, "betaprior");
    get(m.sigmaMA_prior, T.sigmaMA_prior, "sigmaMA_prior");
    get(m.tauMA_prior, T.tauMA_prior, "tauMA_prior");
    get(m.log_eta1_prior, T.log_eta1_prior, "log_eta1_prior");
    get(m.eta2_prior, T.eta2_prior, "eta2_prior");
    get(m.eta3_prior, T.eta3_prior, "eta3_prior");
    get(m.log_eta4_prior, T.log_eta4_prior, "log_eta4_prior");
    get(m.nu_prior, T.nu_prior, "nu_prior");
    for (int i = 0; i < KC_MAX_KO; i++) {
        const bool have = known_object && i < n_known_object;
        std::shared_ptr<DNest4::ContinuousDistribution> none;
        get(have ? m.KO_Pprior[i] : none, T.KO_Pprior[i], "KO_Pprior");
        get(have ? m.KO_Kprior[i] : none, T.KO_Kprior[i], "KO_Kprior");
        get(have ? m.KO_eprior[i] : none, T.KO_eprior[i], "KO_eprior");
        get(have ? m.KO_phiprior[i] : none, T.KO_phiprior[i], "KO_phiprior");
        get(have ? m.KO_wprior[i] : none, T.KO_wprior[i], "KO_wprior");
    }
    const RVConditionalPrior& c = m.planets.conditional_prior;
    if (hyperpriors) {
        // Laplace / Exponential re-parameterised from the particle's hyperparameters on every
        // use (src/RVConditionalPrior.cpp:99-100, 122-123, 139-140)
        T.Pprior.type = KC_PRIOR_LAPLACE;
        T.Kprior.type = KC_PRIOR_EXPONENTIAL;
    } else {
        get(c.Pprior, T.Pprior, "Pprior");
        get(c.Kprior, T.Kprior, "Kprior");
    }
    get(c.eprior, T.eprior, "eprior");
    get(c.phiprior, T.phiprior, "phiprior");
    get(c.wprior, T.wprior, "wprior");
    get(c.log_muP_prior, T.log_muP_prior, "log_muP_prior");
    get(c.wP_prior, T.wP_prior, "wP_prior");
    get(c.log_muK_prior, T.log_muK_prior, "log_muK_prior");
    T.fix = m.fix ? 1 : 0;
    if (m.planets.prior_type != DNest4::PriorType::uniform && bad.empty()) bad = "RJObject prior_type";
    return bad;
}

// ---- a particle <-> one state column (kc_sampler_layout) -------------------------------------
// col(row) is the particle's value of `row`.
template <class ColRef>
inline void model_to_column(const RVmodel& m, const kc_layout& L, const kc_sampler_layout& SL, ColRef col,
                            int32_t& np)
{
    const auto& pl = m.planets;
    np = pl.num_components;
    for (int j = 0; j < L.npmax; j++)
        for (int d = 0; d < 5; d++) {
            col(L.row_P + d * L.npmax + j) = j < np ? pl.components[j][d] : 0.0;
            col(SL.row_u + d * L.npmax + j) = j < np ? pl.u_components[j][d] : 0.0;
        }
    col(SL.row_hyper + 0) = pl.conditional_prior.center;
    col(SL.row_hyper + 1) = pl.conditional_prior.width;
    col(SL.row_hyper + 2) = pl.conditional_prior.muK;
    col(L.row_vsys) = m.background;
    if (multi_instrument) {
        for (int i = 0; i < L.n_jit; i++) col(L.row_jit + i) = m.jitters[i];
        for (int i = 0; i < L.n_off; i++) col(L.row_off + i) = m.offsets[i];
    } else {
        col(L.row_jit) = m.extra_sigma;
    }
    if (trend) {
        col(L.row_trend + 0) = m.slope;
        col(L.row_trend + 1) = m.quadr;
        col(L.row_trend + 2) = m.cubic;
    }
    for (int i = 0; i < L.n_beta; i++) col(L.row_beta + i) = m.betas[i];
    if (GP) {
        col(L.row_eta + 0) = m.eta1;
        col(L.row_eta + 1) = m.eta2;
        col(L.row_eta + 2) = m.eta3;
        col(L.row_eta + 3) = m.eta4;
    }
    if (studentt) col(L.row_nu) = m.nu;
    if (MA) {
        col(L.row_MA + 0) = m.sigmaMA;
        col(L.row_MA + 1) = m.tauMA;
    }
    for (int i = 0; i < L.n_ko; i++) {
        col(L.row_KO_P + i) = m.KO_P[i];
        col(L.row_KO_K + i) = m.KO_K[i];
        col(L.row_KO_e + i) = m.KO_e[i];
        col(L.row_KO_phi + i) = m.KO_phi[i];
        col(L.row_KO_w + i) = m.KO_w[i];
    }
}

template <class ColRef>
inline void column_to_model(RVmodel& m, const kc_layout& L, const kc_sampler_layout& SL, ColRef col, int32_t np)
{
    auto& pl = m.planets;
    pl.num_components = np;
    pl.components.assign(np, std::vector<double>(5));
    pl.u_components.assign(np, std::vector<double>(5));
    for (int j = 0; j < np; j++)
        for (int d = 0; d < 5; d++) {
            pl.components[j][d] = col(L.row_P + d * L.npmax + j);
            pl.u_components[j][d] = col(SL.row_u + d * L.npmax + j);
        }
    pl.added = pl.components;  // same size as components: calculate_mu's from-scratch branch (:296)
    pl.removed.clear();
    m.staleness = 0;
    pl.conditional_prior.center = col(SL.row_hyper + 0);
    pl.conditional_prior.width = col(SL.row_hyper + 1);
    pl.conditional_prior.muK = col(SL.row_hyper + 2);
    m.background = col(L.row_vsys);
    if (multi_instrument) {
        for (int i = 0; i < L.n_jit; i++) m.jitters[i] = col(L.row_jit + i);
        for (int i = 0; i < L.n_off; i++) m.offsets[i] = col(L.row_off + i);
    } else {
        m.extra_sigma = col(L.row_jit);
    }
    if (trend) {
        m.slope = col(L.row_trend + 0);
        m.quadr = col(L.row_trend + 1);
        m.cubic = col(L.row_trend + 2);
    }
    for (int i = 0; i < L.n_beta; i++) m.betas[i] = col(L.row_beta + i);
    if (GP) {
        m.eta1 = col(L.row_eta + 0);
        m.eta2 = col(L.row_eta + 1);
        m.eta3 = col(L.row_eta + 2);
        m.eta4 = col(L.row_eta + 3);
    }
    if (studentt) m.nu = col(L.row_nu);
    if (MA) {
        m.sigmaMA = col(L.row_MA + 0);
        m.tauMA = col(L.row_MA + 1);
    }
    if (L.n_ko) {
        for (auto* v : {&m.KO_P, &m.KO_K, &m.KO_e, &m.KO_phi, &m.KO_w}) v->resize(L.n_ko);
        for (int i = 0; i < L.n_ko; i++) {
            m.KO_P[i] = col(L.row_KO_P + i);
            m.KO_K[i] = col(L.row_KO_K + i);
            m.KO_e[i] = col(L.row_KO_e + i);
            m.KO_phi[i] = col(L.row_KO_phi + i);
            m.KO_w[i] = col(L.row_KO_w + i);
        }
    }
}

inline kc_sampler_layout sampler_layout(const kc_layout& L, int n_particles)
{
    kc_sampler_layout SL{};
    SL.row_u = L.n_rows;
    SL.row_hyper = L.n_rows + 5 * L.npmax;
    SL.n_state = SL.row_hyper + 3;
    SL.n_particles = n_particles;
    return SL;
}

}  // namespace kima_b200
