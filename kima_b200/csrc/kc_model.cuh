// Per-particle setup and the model RV at one epoch — the device form of
// RVmodel::calculate_mu's from-scratch branch
// (/root/reference/src/RVmodel.cpp:306-375, add_known_object :416-432).
// Used by both kernel families: the white-noise likelihood reduces it on the
// fly, the GP kernel turns it into the residual row of its factorisation.
#pragma once
#include "kc_internal.h"
#include "kc_kepler.cuh"

namespace kc {

// Scalars of one particle, staged in shared memory by its owning warp.
struct ParticleScalars {
    double vsys, slope, quadr, cubic, nu;
    double eta1, eta2, eta3, eta4;
    double jit2[KC_MAX_INST];  // jitter^2 per instrument (index 0 = extra_sigma^2)
    double off[KC_MAX_INST];   // offsets, 0 for the last instrument
    double beta[KC_MAX_IND];
    int npl;                   // known objects + planets actually present
    int pad;
};

__device__ __forceinline__ double ldp(const KParams& p, int row, int b)
{
    return __ldg(p.params + (size_t)row * p.ld + b);
}

// One warp prepares particle b.  Known objects come first, as in the reference
// (add_known_object runs before the planet loop, :340-342 then :355).
__device__ __forceinline__ void particle_setup(const KParams& p, int b, PlanetInv* inv,
                                               ParticleScalars* sc, int lane)
{
    const kc_layout& L = p.L;
    const int n_ko = L.n_ko;
    int np = __ldg(p.np + b);
    np = np < 0 ? 0 : (np > L.npmax ? L.npmax : np);
    const int npl = n_ko + np;
    for (int j = lane; j < npl; j += 32) {
        double P, K, phi, e, w;
        if (j < n_ko) {
            P = ldp(p, L.row_KO_P + j, b);
            K = ldp(p, L.row_KO_K + j, b);
            phi = ldp(p, L.row_KO_phi + j, b);
            e = ldp(p, L.row_KO_e + j, b);
            w = ldp(p, L.row_KO_w + j, b);
        } else {
            int k = j - n_ko;
            P = ldp(p, L.row_P + k, b);
            if (p.F.hyperpriors) P = exp_dd(P);  // P = exp(components[j][0]) (:357-358)
            K = ldp(p, L.row_K + k, b);
            phi = ldp(p, L.row_phi + k, b);
            e = ldp(p, L.row_ecc + k, b);
            w = ldp(p, L.row_w + k, b);
        }
        inv[j] = planet_invariants(P, K, phi, e, w, p.M0_epoch);
    }
    if (lane == 0) {
        sc->npl = npl;
        sc->vsys = ldp(p, L.row_vsys, b);
        sc->slope = p.F.trend ? ldp(p, L.row_trend, b) : 0.0;
        sc->quadr = p.F.trend ? ldp(p, L.row_trend + 1, b) : 0.0;
        sc->cubic = p.F.trend ? ldp(p, L.row_trend + 2, b) : 0.0;
        sc->nu = p.F.studentt ? ldp(p, L.row_nu, b) : 0.0;
        if (p.F.GP) {
            sc->eta1 = ldp(p, L.row_eta, b);
            sc->eta2 = ldp(p, L.row_eta + 1, b);
            sc->eta3 = ldp(p, L.row_eta + 2, b);
            sc->eta4 = ldp(p, L.row_eta + 3, b);
        }
    }
    if (lane < KC_MAX_INST) {
        double j2 = 0.0, of = 0.0;
        if (lane < L.n_jit) {
            double jt = ldp(p, L.row_jit + lane, b);
            j2 = kmul(jt, jt);
        }
        if (lane < L.n_off) of = ldp(p, L.row_off + lane, b);
        sc->jit2[lane] = j2;
        sc->off[lane] = of;
    }
    if (lane < KC_MAX_IND) sc->beta[lane] = lane < L.n_beta ? ldp(p, L.row_beta + lane, b) : 0.0;
    __syncwarp();
}

// Everything of mu[i] except the Keplerians (:309-338).
__device__ __forceinline__ double model_background(const KParams& p, const ParticleScalars* sc,
                                                   int i, double ti, int inst)
{
    double mu = sc->vsys;
    if (p.F.trend) {
        double tm = kadd(ti, -p.t_middle);
        double tm2 = kmul(tm, tm);
        // slope*tm + quadr*tm^2 + cubic*tm^3, summed before it is added to mu (:316)
        double tr = kadd(kadd(kmul(sc->slope, tm), kmul(sc->quadr, tm2)),
                         kmul(sc->cubic, kmul(tm2, tm)));
        mu = kadd(mu, tr);
    }
    if (p.F.multi_instrument) mu = kadd(mu, sc->off[inst]);  // off[] is 0 past n_off (:326)
    for (int j = 0; j < p.n_ind; j++)
        mu = kadd(mu, kmul(sc->beta[j], __ldg(p.actind + (size_t)j * p.N + i)));
    return mu;
}

// mu at two epochs at once: two independent Kepler chains per planet give the
// FP64 pipe the instruction-level parallelism a single 8-cycle-latency chain lacks.
template <int MODE>
__device__ __forceinline__ void model_mu2(const KParams& p, const ParticleScalars* sc,
                                          const PlanetInv* inv, int i0, int i1, double t0,
                                          double t1, int inst0, int inst1, double& mu0,
                                          double& mu1)
{
    mu0 = model_background(p, sc, i0, t0, inst0);
    mu1 = model_background(p, sc, i1, t1, inst1);
    const int npl = sc->npl;
    for (int j = 0; j < npl; j++) {
        const PlanetInv pl = inv[j];
        mu0 += kepler_term<MODE>(pl, t0);
        mu1 += kepler_term<MODE>(pl, t1);
    }
}

}  // namespace kc
