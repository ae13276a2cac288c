// A particle as a flat array of doubles, for the exchange between ranks (host/BatchedSampler.h::exchange):
// everything of an RVmodel that a later perturb(), print() or stage() reads
// (src/RVmodel.h:44-76, RJObject.h:34-58, RVConditionalPrior.h:17-20) -- not mu and C, which the GPU path
// never keeps on the host.  Shared by the evaluators (host/RVmodelDevice.h, tests/cpp/RefEvaluator.h).
//
// Include after the reference's headers (RVmodel.h with its private section opened).
#pragma once
#include <cstddef>
#include <vector>

extern const bool known_object;
extern const int n_known_object;

namespace kima_b200 {

inline size_t particle_state_doubles()
{
    const Data& d = Data::get_instance();
    RVmodel probe;
    return 8 + 20 * (size_t)probe.npmax + 3 + 24 + 2 * (size_t)d.number_instruments + d.number_indicators +
           5 * (size_t)n_known_object;
}
inline void particle_pack(const RVmodel& m, double* out)
{
    size_t o = 0;
    const auto& pl = m.planets;
    auto put_vv = [&](const std::vector<std::vector<double>>& vv) {
        out[o++] = (double)vv.size();
        for (const auto& v : vv)
            for (double x : v) out[o++] = x;
    };
    out[o++] = (double)pl.num_components;
    out[o++] = pl.components_changed_flag ? 1.0 : 0.0;
    put_vv(pl.components);
    put_vv(pl.u_components);
    put_vv(pl.added);
    put_vv(pl.removed);
    out[o++] = pl.conditional_prior.center;
    out[o++] = pl.conditional_prior.width;
    out[o++] = pl.conditional_prior.muK;
    const double sc[] = {m.background, m.slope, m.quadr, m.cubic, m.sigmaMA, m.tauMA, m.extra_sigma, m.nu,
                         m.eta1, m.eta2, m.eta3, m.eta4, m.eta5, m.log_eta1, m.log_eta2, m.log_eta3,
                         m.log_eta4, m.log_eta5, m.a, m.b, m.c, m.P, (double)m.staleness, 0.0};
    for (double x : sc) out[o++] = x;
    for (double x : m.offsets) out[o++] = x;
    for (double x : m.jitters) out[o++] = x;
    for (double x : m.betas) out[o++] = x;
    for (const auto* v : {&m.KO_P, &m.KO_K, &m.KO_e, &m.KO_phi, &m.KO_w})
        for (int i = 0; i < n_known_object; i++) out[o++] = i < (int)v->size() ? (*v)[i] : 0.0;
}
inline void particle_unpack(RVmodel& m, const double* in)
{
    size_t o = 0;
    auto& pl = m.planets;
    auto get_vv = [&](std::vector<std::vector<double>>& vv) {
        const size_t n = (size_t)in[o++];
        vv.assign(n, std::vector<double>(5));
        for (auto& v : vv)
            for (double& x : v) x = in[o++];
    };
    pl.num_components = (int)in[o++];
    pl.components_changed_flag = in[o++] != 0.0;
    get_vv(pl.components);
    get_vv(pl.u_components);
    get_vv(pl.added);
    get_vv(pl.removed);
    pl.conditional_prior.center = in[o++];
    pl.conditional_prior.width = in[o++];
    pl.conditional_prior.muK = in[o++];
    double* sc[] = {&m.background, &m.slope, &m.quadr, &m.cubic, &m.sigmaMA, &m.tauMA, &m.extra_sigma, &m.nu,
                    &m.eta1, &m.eta2, &m.eta3, &m.eta4, &m.eta5, &m.log_eta1, &m.log_eta2, &m.log_eta3,
                    &m.log_eta4, &m.log_eta5, &m.a, &m.b, &m.c, &m.P};
    for (double* x : sc) *x = in[o++];
    m.staleness = (unsigned int)in[o++];
    o++;
    for (double& x : m.offsets) x = in[o++];
    for (double& x : m.jitters) x = in[o++];
    for (double& x : m.betas) x = in[o++];
    for (auto* v : {&m.KO_P, &m.KO_K, &m.KO_e, &m.KO_phi, &m.KO_w}) {
        if (known_object) v->resize(n_known_object);
        for (int i = 0; i < n_known_object; i++) {
            if (i < (int)v->size()) (*v)[i] = in[o];
            o++;
        }
    }
}

}  // namespace kima_b200
