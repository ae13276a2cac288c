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
    double sigmaMA, tauMA;
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
        if (p.F.MA) {
            sc->sigmaMA = ldp(p, L.row_MA, b);
            sc->tauMA = ldp(p, L.row_MA + 1, b);
        }
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

// ---- MA(1) term of calculate_mu (:378-386) --------------------------------------------------
// The reference walks the epochs in order:  mu[i] += sigmaMA exp(-|t[i-1]-t[i]|/tauMA) (y[i-1] - mu[i-1]),
// mu[i-1] already updated.  In terms of the residual r = y - mu this is the first-order linear
// recurrence  r[i] = b[i] + c[i] r[i-1]  with  b = y - mu0  (mu0: the model without the MA
// term),  c[i] = -sigmaMA exp(-|t[i-1]-t[i]|/tauMA),  c[0] = 0.

// c[i] for one epoch.
__device__ __forceinline__ double ma_coeff(const KParams& p, const ParticleScalars* sc, int i)
{
    if (i == 0) return 0.0;
    const double dt = fabs(kadd(__ldg(p.t + i - 1), -__ldg(p.t + i)));
    return -kmul(sc->sigmaMA, exp(-kdiv(dt, sc->tauMA)));
}

// One warp turns b[0..N) into r[0..N) in place (shared memory): each lane composes the affine
// maps of a contiguous chunk, the 32 chunk maps are combined with a shuffle scan
// ((C2,B2) o (C1,B1) = (C2 C1, B2 + C2 B1)), then each lane replays its chunk from the
// residual that enters it.
__device__ __forceinline__ void ma_scan_warp(double* b, const double* c, int N, int lane)
{
    const int chunk = (N + 31) / 32;
    const int lo = min(N, lane * chunk), hi = min(N, lo + chunk);
    double C = 1.0, Bc = 0.0;
    for (int i = lo; i < hi; i++) {
        Bc = kfma(c[i], Bc, b[i]);
        C = c[i] * C;
    }
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const double Cp = __shfl_up_sync(0xffffffffu, C, o);
        const double Bp = __shfl_up_sync(0xffffffffu, Bc, o);
        if (lane >= o) {
            Bc = kfma(C, Bp, Bc);
            C = C * Cp;
        }
    }
    double r = __shfl_up_sync(0xffffffffu, Bc, 1);
    if (lane == 0) r = 0.0;
    for (int i = lo; i < hi; i++) {
        r = kfma(c[i], r, b[i]);
        b[i] = r;
    }
}

// All threads of a CTA: b = y - mu0 and c into shared memory, the scan, r left in rbuf.
template <int MODE>
__device__ __forceinline__ void ma_residuals(const KParams& p, const ParticleScalars* sc, const PlanetInv* inv,
                                             double* rbuf, double* cbuf, int tid, int nthreads)
{
    const int N = p.N;
    const bool multi = p.F.multi_instrument != 0;
    for (int i0 = tid; i0 < N; i0 += 2 * nthreads) {
        int i1 = i0 + nthreads;
        const bool has1 = i1 < N;
        if (!has1) i1 = i0;
        double t0 = __ldg(p.t + i0), t1 = __ldg(p.t + i1);
        int in0 = multi ? __ldg(p.obsi + i0) - 1 : 0;
        int in1 = multi ? __ldg(p.obsi + i1) - 1 : 0;
        double mu0, mu1;
        model_mu2<MODE>(p, sc, inv, i0, i1, t0, t1, in0, in1, mu0, mu1);
        rbuf[i0] = __ldg(p.y + i0) - mu0;
        cbuf[i0] = ma_coeff(p, sc, i0);
        if (has1) {
            rbuf[i1] = __ldg(p.y + i1) - mu1;
            cbuf[i1] = ma_coeff(p, sc, i1);
        }
    }
    __syncthreads();
    if (tid < 32) ma_scan_warp(rbuf, cbuf, N, tid);
    __syncthreads();
}

}  // namespace kc
