// Pieces shared by the two GP kernel files: the quasi-periodic kernel with the
// reference's roundings (/root/reference/src/RVmodel.cpp:193-194), the DMMA wrapper
// and the per-lane state of the blocked path.
#pragma once
#include "kc_model.cuh"

namespace kc {

// sin(x) alone, |x| < ~1e5: one polynomial, chosen by quadrant.
__device__ __forceinline__ double sin_core(double x);
__device__ __forceinline__ double sin_fast(double x)
{
    if (fabs(x) > 1.0e5) return sin(x);
    return sin_core(x);
}

// The same without the large-argument guard (callers check |x| <= 1e5 themselves).
__device__ __forceinline__ double sin_core(double x)
{
    const double S1 = -1.66666666666666324348e-01, S2 = 8.33333333332248946124e-03,
                 S3 = -1.98412698298579493134e-04, S4 = 2.75573137070700676789e-06,
                 S5 = -2.50507602534068634195e-08, S6 = 1.58969099521155010221e-10;
    const double C1 = 4.16666666666666019037e-02, C2 = -1.38888888888741095749e-03,
                 C3 = 2.48015872894767294178e-05, C4 = -2.75573143513906633035e-07,
                 C5 = 2.08757232129817482790e-09, C6 = -1.13596475577881948265e-11;
    int k;
    double kd = rint_magic_i(x * TWO_OVER_PI, k);
    double r = kfma(-kd, PIO2_HI, x);
    r = kfma(-kd, PIO2_MID, r);
    r = kfma(-kd, PIO2_LO, r);
    const bool odd = k & 1;
    double z = r * r;
    double p = kfma(z, odd ? C6 : S6, odd ? C5 : S5);
    p = kfma(z, p, odd ? C4 : S4);
    p = kfma(z, p, odd ? C3 : S3);
    p = kfma(z, p, odd ? C2 : S2);
    p = kfma(z, p, odd ? C1 : S1);
    double base = odd ? kfma(-0.5, z, 1.0) : r;
    double mult = odd ? z * z : r * z;
    double v = kfma(mult, p, base);
    return (k & 2) ? -v : v;
}

struct GpConsts {
    double eta1sq;            // eta1*eta1
    double eta2, r_eta2;      // eta and its correctly rounded reciprocal
    double eta3, r_eta3;
    double eta4, r_eta4;
};

// Per-lane state of the blocked path, living beside the factor storage in HBM.
struct __align__(16) GpLane {
    double dinv[2][64];       // 1 / diag(L_JJ), double-buffered by block-column parity (cp.async source)
    GpConsts c;
    double jit2[KC_MAX_INST];
    double logdet, quad;
    int fail, pad;
    double pad2;
};
static_assert(sizeof(GpLane) % 16 == 0, "lanes[] elements stay 16-byte aligned");

__device__ __forceinline__ GpConsts gp_consts(const ParticleScalars& sc)
{
    GpConsts g;
    g.eta1sq = kmul(sc.eta1, sc.eta1);
    g.eta2 = sc.eta2; g.r_eta2 = kdiv(1.0, sc.eta2);
    g.eta3 = sc.eta3; g.r_eta3 = kdiv(1.0, sc.eta3);
    g.eta4 = sc.eta4; g.r_eta4 = kdiv(1.0, sc.eta4);
    return g;
}

// C(i,j) off the diagonal, with the reference's roundings (:193-194):
// eta1*eta1*exp(-0.5*pow(tau/eta2,2) - 2.0*pow(sin(M_PI*tau/eta3)/eta4,2)).
__device__ __forceinline__ double qp_kernel(const GpConsts& g, double tau)
{
    double u = div_by(tau, g.eta2, g.r_eta2);
    double a1 = -0.5 * kmul(u, u);
    double s = sin_fast(div_by(kmul(PI, tau), g.eta3, g.r_eta3));
    double v = div_by(s, g.eta4, g.r_eta4);
    double a2 = 2.0 * kmul(v, v);
    return kmul(g.eta1sq, exp_nonpos(kadd(a1, -a2)));
}

// Two entries at once: two independent, branch-free dependency chains for the FP64 pipe
// (one chain per thread leaves it half idle at 4 warps per scheduler).
__device__ __forceinline__ void qp_kernel2(const GpConsts& g, double tau0, double tau1,
                                           double& c0, double& c1)
{
    double u0 = div_by(tau0, g.eta2, g.r_eta2), u1 = div_by(tau1, g.eta2, g.r_eta2);
    double x0 = div_by(kmul(PI, tau0), g.eta3, g.r_eta3), x1 = div_by(kmul(PI, tau1), g.eta3, g.r_eta3);
    double s0 = sin_core(x0), s1 = sin_core(x1);
    if (!(fabs(x0) <= 1.0e5) || !(fabs(x1) <= 1.0e5)) {  // never taken with any sane eta3
        s0 = sin(x0);
        s1 = sin(x1);
    }
    double v0 = div_by(s0, g.eta4, g.r_eta4), v1 = div_by(s1, g.eta4, g.r_eta4);
    double e0 = kadd(-0.5 * kmul(u0, u0), -2.0 * kmul(v0, v0));
    double e1 = kadd(-0.5 * kmul(u1, u1), -2.0 * kmul(v1, v1));
    c0 = kmul(g.eta1sq, exp_nonpos(e0));
    c1 = kmul(g.eta1sq, exp_nonpos(e1));
}

__device__ __forceinline__ void dmma(double& c0, double& c1, double a, double b)
{
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                 : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}

}  // namespace kc
