// Device-side proposals and accept logic: the per-lane part of
// one DNest4 MCMC step, restated from the reference operation for operation and random draw
// for random draw, so that the unmodified reference sampler driven by the same stream
// (the RNG stand-in of the test harness) takes the same decisions.
//
//   Sampler::mcmc_thread / update_particle / update_level_assignment
//                                   DNest4/code/SamplerImpl.h:134-265
//   RVmodel::perturb                src/RVmodel.cpp:435-820 (the parameter moves; the incremental
//                                   updates of mu / C between them have no counterpart here: the
//                                   likelihood kernels recompute from the parameters)
//   RJObject::perturb & friends     DNest4/code/RJObject/RJObjectImpl.h:62-231
//   ConditionalPrior::perturb1/2    DNest4/code/RJObject/ConditionalPriors/ConditionalPrior.cpp:17-55
//   RVConditionalPrior              src/RVConditionalPrior.cpp:57-158
//   ContinuousDistribution::perturb DNest4/code/Distributions/ContinuousDistribution.cpp:12-20
//   the distributions               DNest4/code/Distributions/*.cpp, src/distributions/Fixed.cpp
//   DNest4::mod / wrap / normal_cdf / normal_inverse_cdf   DNest4/code/Utils.cpp:6-26, 74-159
//   LikelihoodType                  DNest4/code/LikelihoodType.cpp:17-43
//
// The header also compiles for the host (g++, kc_host_check): tests/ run it on the CPU against
// the reference harness without a GPU.  The product only ever runs the device instantiation.
#pragma once
#include <cmath>
#include <cstddef>
#include <cstdint>
#include "../../include/kima_sampler.h"
#include "../../include/kc_philox.h"

// KS_FN functions are real calls on the device (__noinline__): the step is branchy scalar code
// whose call tree (16 distributions x cdf / cdf_inverse / log_pdf under every move of
// perturb) would otherwise be inlined into each of the kernels that reach it.
#if defined(__CUDACC__)
#define KS_HD __host__ __device__ __forceinline__
#define KS_FN __host__ __device__ __noinline__
#else
#define KS_HD inline
#define KS_FN inline
#endif

namespace kc {
namespace smp {

constexpr double KS_PI = 3.14159265358979323846;  // M_PI

// ---- the lane's random stream (include/kc_philox.h) ---------------------------------------
struct Rng {
    uint64_t key, ctr;
    KS_FN double rand() { return kc_philox_uniform(key, ctr++); }
    KS_FN double randn()
    {
        const double u1 = rand();
        const double u2 = rand();
        return sqrt(-2.0 * log(1.0 - u1)) * cos(2.0 * KS_PI * u2);
    }
    KS_HD double randt2()  // RNG.cpp:38-42
    {
        const double n = randn();
        const double u = rand();
        return n / sqrt(-log(u));
    }
    KS_FN double randh()  // RNG.cpp:44-49 (randh_is_randh2 is false everywhere in kima)
    {
        const double t = randt2();
        const double n = randn();
        return pow(10.0, 1.5 - 3 * fabs(t)) * n;
    }
    KS_HD int rand_int(int N) { return (int)floor(N * rand()); }  // RNG.cpp:57-60
};

// ---- DNest4/code/Utils.cpp ------------------------------------------------------------------
KS_HD double mod_d(double y, double x) { return (y / x - floor(y / x)) * x; }   // :6-10
KS_HD void wrap(double& x, double lo, double hi) { x = mod_d(x - lo, hi - lo) + lo; }  // :12-15
KS_FN int mod_i(int y, int x)  // :17-24
{
    if (y >= 0) return y - (y / x) * x;
    const int z = -y - 1;
    return (x - 1) - (z - (z / x) * x);
}
KS_HD double normal_cdf(double x) { return 0.5 * (1. + erf(x / sqrt(2.))); }  // :74-77

KS_HD double poly8(const double* a, double x)  // r8poly_value, Utils.cpp:163-170
{
    double v = 0.0;
    for (int i = 7; i >= 0; i--) v = v * x + a[i];
    return v;
}

// Wichura's algorithm AS 241 (PPND16), Applied Statistics 37 (1988) 477-484: the published
// rational approximations; Utils.cpp:84-159 is the same algorithm.
KS_FN double normal_inverse_cdf(double p)
{
    const double A[8] = {3.3871328727963666080, 1.3314166789178437745e+2, 1.9715909503065514427e+3,
                         1.3731693765509461125e+4, 4.5921953931549871457e+4, 6.7265770927008700853e+4,
                         3.3430575583588128105e+4, 2.5090809287301226727e+3};
    const double B[8] = {1.0, 4.2313330701600911252e+1, 6.8718700749205790830e+2, 5.3941960214247511077e+3,
                         2.1213794301586595867e+4, 3.9307895800092710610e+4, 2.8729085735721942674e+4,
                         5.2264952788528545610e+3};
    const double C[8] = {1.42343711074968357734, 4.63033784615654529590, 5.76949722146069140550,
                         3.64784832476320460504, 1.27045825245236838258, 2.41780725177450611770e-1,
                         2.27238449892691845833e-2, 7.74545014278341407640e-4};
    const double D[8] = {1.0, 2.05319162663775882187, 1.67638483018380384940, 6.89767334985100004550e-1,
                         1.48103976427480074590e-1, 1.51986665636164571966e-2, 5.47593808499534494600e-4,
                         1.05075007164441684324e-9};
    const double E[8] = {6.65790464350110377720, 5.46378491116411436990, 1.78482653991729133580,
                         2.96560571828504891230e-1, 2.65321895265761230930e-2, 1.24266094738807843860e-3,
                         2.71155556874348757815e-5, 2.01033439929228813265e-7};
    const double F[8] = {1.0, 5.99832206555887937690e-1, 1.36929880922735805310e-1, 1.48753612908506148525e-2,
                         7.86869131145613259100e-4, 1.84631831751005468180e-5, 1.42151175831644588870e-7,
                         2.04426310338993978564e-15};
    const double q = p - 0.5;
    if (fabs(q) <= 0.425) {
        const double r = 0.180625 - q * q;
        return q * poly8(A, r) / poly8(B, r);
    }
    double r = (q < 0.0) ? p : 1.0 - p;
    if (r <= 0.0) return 0.;
    r = sqrt(-log(r));
    double value;
    if (r <= 5.0) {
        r = r - 1.6;
        value = poly8(C, r) / poly8(D, r);
    } else {
        r = r - 5.0;
        value = poly8(E, r) / poly8(F, r);
    }
    return (q < 0.0) ? -value : value;
}

// ---- the distributions ------------------------------------------------------------------------
KS_HD double neg_inf() { return -HUGE_VAL; }
KS_HD int sign_i(double x) { return x == 0.0 ? 0 : (x > 0.0 ? 1 : -1); }  // Laplace::sign
KS_HD double cauchy_cdf(double c, double w, double x) { return atan((x - c) / w) / KS_PI + 0.5; }
KS_HD double cauchy_icdf(double c, double w, double x) { return c + w * tan(KS_PI * (x - 0.5)); }
KS_HD double gauss_cdf(double c, double w, double x) { return normal_cdf((x - c) / w); }
KS_HD double gauss_icdf(double c, double w, double x) { return c + w * normal_inverse_cdf(x); }
KS_HD double expo_cdf(double s, double x) { return 1 - exp(-x / s); }
KS_HD double expo_icdf(double s, double p) { return -s * log(1. - p); }
KS_HD double clamp_lu(double x, double lo, double hi) { return fmax(fmin(x, hi), lo); }

// mixGaussianLogUniform::cdf (src/distributions/mixGaussianLogUniform.cpp:19-22)
KS_FN double mix_glu_cdf(const kc_prior& q, double x)
{
    const double a = q.p[0], b = q.p[1], c = q.p[2], d = q.p[3];
    double lu = (log(x) - log(c)) / q.k[0];
    if (x < c) lu = 0.0;
    if (x > d) lu = 1.0;
    return 0.5 * gauss_cdf(a, b, x) + 0.5 * lu;
}

KS_FN double prior_cdf(const kc_prior& q, double x)
{
    const double a = q.p[0], b = q.p[1], c = q.p[2], d = q.p[3];
    switch (q.type) {
    case KC_PRIOR_FIXED: return x == a ? 1.0 : 0.0;
    case KC_PRIOR_UNIFORM:
        if (x < a) return 0.;
        if (x >= b) return 1.;
        return (x - a) / (b - a);
    case KC_PRIOR_LOGUNIFORM:
        if (x < a) return 0.0;
        if (x > b) return 1.0;
        return (log(x) - log(a)) / q.k[0];
    case KC_PRIOR_MODLOGUNIFORM: return q.k[0] * log(1.0 + x / a);
    case KC_PRIOR_GAUSSIAN: return gauss_cdf(a, b, x);
    case KC_PRIOR_TRUNC_GAUSSIAN:
        return (gauss_cdf(a, b, clamp_lu(x, c, d)) - gauss_cdf(a, b, c)) / q.k[1];
    case KC_PRIOR_CAUCHY: return cauchy_cdf(a, b, x);
    case KC_PRIOR_TRUNC_CAUCHY:
        return (cauchy_cdf(a, b, clamp_lu(x, c, d)) - cauchy_cdf(a, b, c)) / q.k[1];
    case KC_PRIOR_EXPONENTIAL: return expo_cdf(a, x);
    case KC_PRIOR_TRUNC_EXPONENTIAL:
        return (expo_cdf(a, clamp_lu(x, b, c)) - expo_cdf(a, b)) / q.k[1];
    case KC_PRIOR_LAPLACE: return 0.5 + 0.5 * sign_i(x - a) * (1.0 - exp(-fabs(x - a) / b));
    case KC_PRIOR_KUMARASWAMY: return 1 - pow(1 - pow(x, a), b);
    case KC_PRIOR_RAYLEIGH: return 1 - exp(-(x * x) / (2. * a * a));
    case KC_PRIOR_TRUNC_RAYLEIGH: return ((1 - exp(-(x * x) / (2. * a * a))) - q.k[0]) / q.k[1];
    case KC_PRIOR_TRIANGULAR:
        if (x <= a) return 0.;
        if (x <= b) return pow(x - a, 2) / ((c - a) * (b - a));
        if (x < c) return 1. - pow(c - x, 2) / ((c - a) * (c - b));
        return 1.;
    case KC_PRIOR_MIX_GAUSS_LOGUNIFORM: return mix_glu_cdf(q, x);
    case KC_PRIOR_PARETO:
        if (x <= a) return 0.0;
        return 1.0 - pow(x / a, b);
    }
    return NAN;
}

KS_FN double prior_cdf_inverse(const kc_prior& q, double x)
{
    const double a = q.p[0], b = q.p[1], c = q.p[2];
    switch (q.type) {
    case KC_PRIOR_FIXED: return a;
    case KC_PRIOR_UNIFORM: return a + x * (b - a);
    case KC_PRIOR_LOGUNIFORM: return a * pow(b / a, x);
    case KC_PRIOR_MODLOGUNIFORM: return (exp(x / q.k[0]) - 1.0) * a;
    case KC_PRIOR_GAUSSIAN: return gauss_icdf(a, b, x);
    case KC_PRIOR_TRUNC_GAUSSIAN: return gauss_icdf(a, b, gauss_cdf(a, b, c) + x * q.k[1]);
    case KC_PRIOR_CAUCHY: return cauchy_icdf(a, b, x);
    case KC_PRIOR_TRUNC_CAUCHY: return cauchy_icdf(a, b, cauchy_cdf(a, b, c) + x * q.k[1]);
    case KC_PRIOR_EXPONENTIAL: return expo_icdf(a, x);
    case KC_PRIOR_TRUNC_EXPONENTIAL: return expo_icdf(a, expo_cdf(a, b) + x * q.k[1]);
    case KC_PRIOR_LAPLACE: return a - b * sign_i(x - 0.5) * log(1.0 - 2 * fabs(x - 0.5));
    case KC_PRIOR_KUMARASWAMY: return pow(1 - pow(1 - x, 1 / b), 1 / a);
    case KC_PRIOR_RAYLEIGH: return a * sqrt(-2. * log(1 - x));
    case KC_PRIOR_TRUNC_RAYLEIGH: return a * sqrt(-2. * log(1 - (q.k[0] + x * q.k[1])));
    case KC_PRIOR_TRIANGULAR: {
        const double centre_ccdf = (b - a) / (c - a);
        if (x <= centre_ccdf) return a + sqrt((c - a) * (b - a) * x);
        return c - sqrt((1. - x) * (c - a) * (c - b));
    }
    case KC_PRIOR_PARETO: return a * pow(1.0 - x, 1.0 / b);
    case KC_PRIOR_MIX_GAUSS_LOGUNIFORM: {
        // the reference's bisection on [lower, upper], tolerance 2e-12 (mixGaussianLogUniform.cpp:25-46)
        const double tol = 2e-12;
        double lo = c, hi = q.p[3], mid = lo;
        while ((hi - lo) >= tol) {
            mid = (lo + hi) / 2;
            const double fm = mix_glu_cdf(q, mid) - x;
            if (fm == 0.0) break;
            else if (fm * (mix_glu_cdf(q, lo) - x) < 0) hi = mid;
            else lo = mid;
        }
        return mid;
    }
    }
    return NAN;
}

KS_FN double prior_log_pdf(const kc_prior& q, double x)
{
    const double a = q.p[0], b = q.p[1], c = q.p[2], d = q.p[3];
    switch (q.type) {
    case KC_PRIOR_FIXED: return x == a ? 0.0 : neg_inf();
    case KC_PRIOR_UNIFORM:
        if (x < a || x > b) return neg_inf();
        return -log(b - a);
    case KC_PRIOR_LOGUNIFORM:
        if (x < a || x > b) return neg_inf();
        return -log(x) - log(q.k[0]);
    case KC_PRIOR_MODLOGUNIFORM: return log(q.k[0]) - log(1.0 + x / a);
    case KC_PRIOR_GAUSSIAN: {
        const double r = (x - a) / b;
        return -0.5 * r * r - log(sqrt(2 * KS_PI));
    }
    case KC_PRIOR_TRUNC_GAUSSIAN: {
        if (x < c || x > d) return neg_inf();
        const double r = (x - a) / b;
        return (-0.5 * r * r - log(sqrt(2 * KS_PI))) - log(q.k[1]);
    }
    case KC_PRIOR_CAUCHY: return -log(KS_PI * b) - log(1.0 + pow((x - a) / b, 2));
    case KC_PRIOR_TRUNC_CAUCHY:
        if (x < c || x > d) return neg_inf();
        return (-log(KS_PI * b) - log(1.0 + pow((x - a) / b, 2))) - log(q.k[1]);
    case KC_PRIOR_EXPONENTIAL:
        if (x < 0.0) return neg_inf();
        return -log(a) - x / a;
    case KC_PRIOR_TRUNC_EXPONENTIAL:
        if (x < b || x > c) return neg_inf();
        if (x < 0.0) return neg_inf();
        return (-log(a) - x / a) - log(q.k[1]);
    case KC_PRIOR_LAPLACE: return -log(2.0 * b) - fabs(x - a) / b;
    case KC_PRIOR_KUMARASWAMY: return log(a) + log(b) + (a - 1) * log(x) + (b - 1) * log(1 - pow(x, a));
    case KC_PRIOR_RAYLEIGH: {
        const double r = x / a;
        return log(r) - 0.5 * r * r;
    }
    case KC_PRIOR_TRUNC_RAYLEIGH: {
        if (x < b || x >= c) return neg_inf();
        const double r = x / a;
        return log(r) - 0.5 * r * r - q.k[2];
    }
    case KC_PRIOR_TRIANGULAR:
        if (x < a || x > c) return neg_inf();
        if (x > a && x <= b) return log(2.0 * (x - a)) - log((c - a) * (b - a));
        return log((c - a)) - log((c - a) * (c - b));  // the reference's third branch, as written
    case KC_PRIOR_MIX_GAUSS_LOGUNIFORM: {
        const double r = (x - a) / b;
        const double lg = -0.5 * r * r - log(sqrt(2 * KS_PI));                              // Gaussian::log_pdf
        const double ll = (x < c || x > d) ? neg_inf() : -log(x) - log(q.k[0]);             // LogUniform::log_pdf
        return log(0.5 * exp(lg) + 0.5 * exp(ll));
    }
    case KC_PRIOR_PARETO:
        if (x <= a) return -1E300;
        return log(b) + b * log(a) - (b + 1.0) * log(x);
    }
    return NAN;
}

// ContinuousDistribution::perturb (ContinuousDistribution.cpp:12-20).  It is not virtual
// (ContinuousDistribution.h:19), so Fixed::perturb (src/distributions/Fixed.h:30-31) is never
// reached through the shared_ptr<ContinuousDistribution> members: a Fixed prior draws its
// randh like every other and lands on val again through cdf_inverse.
KS_FN void prior_perturb(const kc_prior& q, double& x, Rng& rng)
{
    x = prior_cdf(q, x);
    x += rng.randh();
    wrap(x, 0.0, 1.0);
    x = prior_cdf_inverse(q, x);
}

// ---- everything a step reads, passed by value ----------------------------------------------
struct SamplerParams {
    kc_model_flags F;
    kc_layout L;
    kc_sampler_layout SL;
    int n_ind;
    const kc_prior_table* pri;
    // particles: state S [n_state][ldp], one column per particle
    double* S;
    size_t ldp;
    int32_t* np;
    double* logl;
    double* tb;
    int32_t* level;
    // proposals: Q [n_state][ldq], one column per lane (rows [0, n_rows) are what the likelihood reads)
    double* Q;
    size_t ldq;
    int32_t* npq;
    const double* logl_q;  // the likelihood kernels' output for the proposals
    // lanes
    uint64_t* ctr;
    uint64_t key0;     // lane k draws with key key0 + k
    int32_t* which;    // particle chosen by the step in flight
    int32_t* flags;    // bit 0 = particle_first (the coin of mcmc_thread), bit 1 = scored
    double* above;     // [lanes][max_steps][2]
    int32_t* n_above;
    int max_steps;
    // levels
    int n_levels;
    const double* lv_logl;
    const double* lv_tb;
    const double* lv_logX;
    const uint64_t* lv_tries;
    uint32_t* cnt;     // [4][Lcap][ldl]: accepts, tries, visits, exceeds increments per lane
    size_t ldl;
    int Lcap;
    int enough, at_max;
    double lambda, beta;
    int lanes, P;
};

struct Col {
    double* base;
    size_t ld;
    KS_HD double& operator()(int row) const { return base[(size_t)row * ld]; }
};

KS_HD bool lt(double v1, double t1, double v2, double t2)  // LikelihoodType::operator<
{
    if (v1 < v2) return true;
    if (v1 == v2 && t1 < t2) return true;
    return false;
}

KS_HD uint32_t& counter(const SamplerParams& sp, int c, int j, int lane)
{
    return sp.cnt[((size_t)c * sp.Lcap + j) * sp.ldl + lane];
}

KS_HD double log_push(const SamplerParams& sp, int which_level)  // SamplerImpl.h:442-452
{
    if (sp.enough) return 0.0;
    const int i = which_level - (sp.n_levels - 1);
    return (double)i / sp.lambda;
}

// Sampler::update_level_assignment, SamplerImpl.h:224-265
KS_FN void update_level_assignment(const SamplerParams& sp, Rng& rng, int lane, int which)
{
    const int cur = sp.level[which];
    // `pow(10., 2.*rng.rand())*rng.randn()`: g++ evaluates the left operand first
    const double u = rng.rand();
    const double n = rng.randn();
    int proposal = cur + (int)(pow(10., 2. * u) * n);
    if (proposal == cur) proposal = (rng.rand() < 0.5) ? (proposal - 1) : (proposal + 1);
    proposal = mod_i(proposal, sp.n_levels);
    double log_A = -sp.lv_logX[proposal] + sp.lv_logX[cur];
    log_A += log_push(sp, proposal) - log_push(sp, cur);
    if (sp.at_max) {
        const uint64_t tc = sp.lv_tries[cur] + counter(sp, 1, cur, lane);
        const uint64_t tp = sp.lv_tries[proposal] + counter(sp, 1, proposal, lane);
        // Level::get_tries() returns unsigned int (Level.h:50-51): 32-bit arithmetic, as there
        log_A += sp.beta * log((double)((uint32_t)tc + 1u) / (double)((uint32_t)tp + 1u));
    }
    if (log_A > 0.) log_A = 0.;
    const double r = rng.rand();
    if (r <= exp(log_A) && lt(sp.lv_logl[proposal], sp.lv_tb[proposal], sp.logl[which], sp.tb[which]))
        sp.level[which] = proposal;
}

// ---- RVConditionalPrior (src/RVConditionalPrior.cpp) -------------------------------------------
struct CondPrior {
    kc_prior P, K;           // Laplace(center, width) / Exponential(muK) under hyperpriors
    const kc_prior *e, *phi, *w;
    bool hyper;
};

KS_FN CondPrior cond_prior(const SamplerParams& sp, const Col& q)
{
    CondPrior cp;
    cp.hyper = sp.F.hyperpriors != 0;
    cp.e = &sp.pri->eprior;
    cp.phi = &sp.pri->phiprior;
    cp.w = &sp.pri->wprior;
    if (cp.hyper) {  // Pprior->setpars(center, width); Kprior->setpars(muK)
        cp.P.type = KC_PRIOR_LAPLACE;
        cp.P.p[0] = q(sp.SL.row_hyper + 0);
        cp.P.p[1] = q(sp.SL.row_hyper + 1);
        cp.K.type = KC_PRIOR_EXPONENTIAL;
        cp.K.p[0] = q(sp.SL.row_hyper + 2);
    } else {
        cp.P = sp.pri->Pprior;
        cp.K = sp.pri->Kprior;
    }
    return cp;
}

// component j of a column: rows row_P + d*npmax + j, d = 0..4 (P, K, phi, ecc, w)
KS_HD int crow(const SamplerParams& sp, int d, int j) { return sp.L.row_P + d * sp.L.npmax + j; }
KS_HD int urow(const SamplerParams& sp, int d, int j) { return sp.SL.row_u + d * sp.L.npmax + j; }

KS_FN void from_uniform(const CondPrior& cp, double v[5])  // :118-133 (vec order P, K, phi, e, w)
{
    v[0] = prior_cdf_inverse(cp.P, v[0]);
    v[1] = prior_cdf_inverse(cp.K, v[1]);
    v[2] = prior_cdf_inverse(*cp.phi, v[2]);
    v[3] = prior_cdf_inverse(*cp.e, v[3]);
    v[4] = prior_cdf_inverse(*cp.w, v[4]);
}
KS_FN void to_uniform(const CondPrior& cp, double v[5])  // :135-150
{
    v[0] = prior_cdf(cp.P, v[0]);
    v[1] = prior_cdf(cp.K, v[1]);
    v[2] = prior_cdf(*cp.phi, v[2]);
    v[3] = prior_cdf(*cp.e, v[3]);
    v[4] = prior_cdf(*cp.w, v[4]);
}
KS_FN double cond_log_pdf(const CondPrior& cp, const double v[5])  // :87-116
{
    if (cp.hyper) {
        if (v[2] < 0. || v[2] > 2. * KS_PI || v[3] < 0. || v[3] >= 1.0 || v[4] < 0. || v[4] > 2. * KS_PI)
            return -1E300;
    } else {
        if (v[0] < 1. || v[0] > 1E4 || v[1] < 0. || v[2] < 0. || v[2] > 2. * KS_PI || v[3] < 0. ||
            v[3] >= 1.0 || v[4] < 0. || v[4] > 2. * KS_PI)
            return -1E300;
    }
    return prior_log_pdf(cp.P, v[0]) + prior_log_pdf(cp.K, v[1]) + prior_log_pdf(*cp.phi, v[2]) +
           prior_log_pdf(*cp.e, v[3]) + prior_log_pdf(*cp.w, v[4]);
}

// The kc_layout rows hold (P, K, phi, ecc, w); a component vector of the reference is
// (P, K, phi, e, w) as well (src/RVmodel.cpp:355-366 reads components[j][0..4] in that order).
KS_HD void load_comp(const SamplerParams& sp, const Col& q, int j, double v[5])
{
    for (int d = 0; d < 5; d++) v[d] = q(crow(sp, d, j));
}
KS_HD void store_comp(const SamplerParams& sp, const Col& q, int j, const double v[5])
{
    for (int d = 0; d < 5; d++) q(crow(sp, d, j)) = v[d];
}

// RVConditionalPrior::perturb_hyperparameters, :66-85
KS_FN double perturb_hyperparameters(const SamplerParams& sp, const Col& q, Rng& rng)
{
    if (sp.F.hyperpriors) {
        const int which = rng.rand_int(3);
        if (which == 0) prior_perturb(sp.pri->log_muP_prior, q(sp.SL.row_hyper + 0), rng);
        else if (which == 1) prior_perturb(sp.pri->wP_prior, q(sp.SL.row_hyper + 1), rng);
        else {
            double muK = log(q(sp.SL.row_hyper + 2));
            prior_perturb(sp.pri->log_muK_prior, muK, rng);
            q(sp.SL.row_hyper + 2) = exp(muK);
        }
    }
    return 0.;
}

// ---- RJObject (DNest4/code/RJObject/RJObjectImpl.h) ----------------------------------------------
KS_FN void add_component(const SamplerParams& sp, const Col& q, int& np, Rng& rng)  // :85-108
{
    if (np >= sp.L.npmax) return;
    const int j = np++;
    double v[5];
    for (int d = 0; d < 5; d++) {
        v[d] = rng.rand();
        q(urow(sp, d, j)) = v[d];
    }
    const CondPrior cp = cond_prior(sp, q);
    from_uniform(cp, v);
    store_comp(sp, q, j, v);
}

KS_FN void remove_component(const SamplerParams& sp, const Col& q, int& np, Rng& rng)  // :214-233
{
    if (np <= 0) return;
    const int i = rng.rand_int(np);
    for (int j = i; j + 1 < np; j++)
        for (int d = 0; d < 5; d++) {
            q(crow(sp, d, j)) = q(crow(sp, d, j + 1));
            q(urow(sp, d, j)) = q(urow(sp, d, j + 1));
        }
    np--;
}

KS_FN double perturb_num_components(const SamplerParams& sp, const Col& q, int& np, Rng& rng)  // :110-141
{
    const int maxn = sp.L.npmax;
    const double delta = maxn * rng.randh();
    int difference = (int)delta;
    if (difference == 0) difference = (rng.rand() <= 0.5) ? -1 : 1;
    const int new_num = mod_i(np + difference, maxn + 1);
    difference = new_num - np;
    if (difference > 0)
        for (int i = 0; i < difference; i++) add_component(sp, q, np, rng);
    else
        for (int i = 0; i < -difference; i++) remove_component(sp, q, np, rng);
    return 0.;
}

KS_FN double perturb_components(const SamplerParams& sp, const Col& q, int np, Rng& rng)  // :44-83
{
    if (np == 0) return 0.;
    const double chance = pow(10.0, -3 * fabs(rng.randt2()));
    uint32_t change = 0;
    int count = 0;
    for (int i = 0; i < np; i++)
        if (rng.rand() <= chance) {
            change |= 1u << i;
            count++;
        }
    if (count == 0) change |= 1u << rng.rand_int(np);
    const CondPrior cp = cond_prior(sp, q);
    for (int i = 0; i < np; i++)
        if (change & (1u << i)) {
            double v[5];
            for (int d = 0; d < 5; d++) {
                double u = q(urow(sp, d, i));
                u += rng.randh();
                u = mod_d(u, 1.0);
                q(urow(sp, d, i)) = u;
                v[d] = u;
            }
            from_uniform(cp, v);
            store_comp(sp, q, i, v);
        }
    return 0.;
}

// ConditionalPrior::perturb1, ConditionalPrior.cpp:17-40
KS_FN double cond_perturb1(const SamplerParams& sp, const Col& q, int np, Rng& rng)
{
    double logH = 0.;
    double v[5];
    {
        const CondPrior cp = cond_prior(sp, q);
        for (int i = 0; i < np; i++) {
            load_comp(sp, q, i, v);
            logH -= cond_log_pdf(cp, v);
        }
    }
    logH += perturb_hyperparameters(sp, q, rng);
    const CondPrior cp = cond_prior(sp, q);
    for (int i = 0; i < np; i++) {
        load_comp(sp, q, i, v);
        logH += cond_log_pdf(cp, v);
        to_uniform(cp, v);
        for (int d = 0; d < 5; d++) {
            q(urow(sp, d, i)) = v[d];
            if (v[d] <= 0.0 || v[d] >= 1.0) logH = -1E300;
        }
    }
    return logH;
}

// ConditionalPrior::perturb2, ConditionalPrior.cpp:42-55
KS_FN double cond_perturb2(const SamplerParams& sp, const Col& q, int np, Rng& rng)
{
    double logH = 0.;
    logH += perturb_hyperparameters(sp, q, rng);
    const CondPrior cp = cond_prior(sp, q);
    for (int i = 0; i < np; i++) {
        double v[5];
        for (int d = 0; d < 5; d++) v[d] = q(urow(sp, d, i));
        from_uniform(cp, v);
        store_comp(sp, q, i, v);
    }
    return logH;
}

// RJObject::perturb, :143-193 (prior_type uniform, hyperparameters = true as kima calls it)
// The move a proposal makes is decided by the first draws of RVmodel::perturb and RJObject::perturb; the
// kernels sort the lanes of a CTA by it before doing the work, so that the threads of a warp run the
// same branch.  move_choose consumes exactly those selection draws, in the reference's order; move_apply
// continues on the same stream.
enum Move : int {
    MV_NONE = 0,        // npmax == 0: RJObject::perturb returns at once
    MV_NUM_COMPONENTS,  // RJObjectImpl.h:180-181 (birth / death)
    MV_COND1,           // :185-186 conditional prior's perturb1
    MV_COND2,           // :187-188 perturb2
    MV_COMPONENTS,      // :191
    MV_ETA1, MV_ETA2, MV_ETA3, MV_ETA4,
    MV_SIGMA_MA, MV_TAU_MA,
    MV_JITTER,          // jitters (+ nu, known objects)
    MV_BACKGROUND,
    MV_COUNT
};
KS_FN int planets_choose(const SamplerParams& sp, Rng& rng)
{
    if (sp.L.npmax == 0) return MV_NONE;
    const int which = sp.pri->fix ? (1 + rng.rand_int(2)) : rng.rand_int(3);
    if (which == 0) return MV_NUM_COMPONENTS;
    if (which == 1) return rng.rand() <= 0.5 ? MV_COND1 : MV_COND2;
    return MV_COMPONENTS;
}

// ---- RVmodel::perturb, src/RVmodel.cpp:435-820 ----------------------------------------------------
KS_FN void perturb_jitters(const SamplerParams& sp, const Col& q, Rng& rng)
{
    // multi_instrument: every jitters[i]; else extra_sigma (the single jitter row)
    for (int i = 0; i < sp.L.n_jit; i++) prior_perturb(sp.pri->Jprior, q(sp.L.row_jit + i), rng);
}
KS_FN void perturb_known_objects(const SamplerParams& sp, const Col& q, Rng& rng)
{
    for (int i = 0; i < sp.L.n_ko; i++) {
        prior_perturb(sp.pri->KO_Pprior[i], q(sp.L.row_KO_P + i), rng);
        prior_perturb(sp.pri->KO_Kprior[i], q(sp.L.row_KO_K + i), rng);
        prior_perturb(sp.pri->KO_eprior[i], q(sp.L.row_KO_e + i), rng);
        prior_perturb(sp.pri->KO_phiprior[i], q(sp.L.row_KO_phi + i), rng);
        prior_perturb(sp.pri->KO_wprior[i], q(sp.L.row_KO_w + i), rng);
    }
}
// vsys, offsets, trend, indicator slopes; the GP and MA branches move the slope only
// (:568-586, :708-721), the white-noise branch every coefficient up to `degree` (:782-800)
KS_FN void perturb_background(const SamplerParams& sp, const Col& q, Rng& rng, bool slope_only)
{
    prior_perturb(sp.pri->Cprior, q(sp.L.row_vsys), rng);
    for (int j = 0; j < sp.L.n_off; j++) prior_perturb(sp.pri->offsets_prior, q(sp.L.row_off + j), rng);
    if (sp.F.trend) {
        if (slope_only) prior_perturb(sp.pri->slope_prior, q(sp.L.row_trend + 0), rng);
        else {
            if (sp.F.degree >= 1) prior_perturb(sp.pri->slope_prior, q(sp.L.row_trend + 0), rng);
            if (sp.F.degree >= 2) prior_perturb(sp.pri->quadr_prior, q(sp.L.row_trend + 1), rng);
            if (sp.F.degree == 3) prior_perturb(sp.pri->cubic_prior, q(sp.L.row_trend + 2), rng);
        }
    }
    for (int j = 0; j < sp.L.n_beta; j++) prior_perturb(sp.pri->betaprior, q(sp.L.row_beta + j), rng);
}
KS_FN void perturb_log_of(const kc_prior& pr, double& x, Rng& rng)  // log_eta = log(eta); perturb; eta = exp
{
    double lx = log(x);
    prior_perturb(pr, lx, rng);
    x = exp(lx);
}

KS_FN int move_choose(const SamplerParams& sp, Rng& rng)
{
    if (sp.F.GP) {
        if (rng.rand() <= 0.5) return planets_choose(sp, rng);
        if (rng.rand() <= 0.5) {
            if (sp.F.kernel == 0) {  // standard, :458-487
                if (rng.rand() <= 0.25) return MV_ETA1;
                if (rng.rand() <= 0.33330) return MV_ETA2;
                if (rng.rand() <= 0.5) return MV_ETA3;
                return MV_ETA4;
            }
            // celerite, :489-509
            if (rng.rand() <= 0.33330) return MV_ETA1;
            if (rng.rand() <= 0.5) return MV_ETA2;
            return MV_ETA3;
        }
        return rng.rand() <= 0.5 ? MV_JITTER : MV_BACKGROUND;
    }
    if (sp.F.MA) {
        if (rng.rand() <= 0.5) return planets_choose(sp, rng);
        if (rng.rand() <= 0.5) return rng.rand() <= 0.5 ? MV_SIGMA_MA : MV_TAU_MA;
        return rng.rand() <= 0.5 ? MV_JITTER : MV_BACKGROUND;
    }
    if (rng.rand() <= 0.75) return planets_choose(sp, rng);
    return rng.rand() <= 0.5 ? MV_JITTER : MV_BACKGROUND;
}
KS_FN double move_apply(const SamplerParams& sp, int move, const Col& q, int& np, Rng& rng)
{
    const int e = sp.L.row_eta;
    switch (move) {
    case MV_NUM_COMPONENTS: return perturb_num_components(sp, q, np, rng);
    case MV_COND1: return cond_perturb1(sp, q, np, rng);
    case MV_COND2: return cond_perturb2(sp, q, np, rng);
    case MV_COMPONENTS: return perturb_components(sp, q, np, rng);
    case MV_ETA1: perturb_log_of(sp.pri->log_eta1_prior, q(e + 0), rng); break;
    case MV_ETA2: prior_perturb(sp.pri->eta2_prior, q(e + 1), rng); break;
    case MV_ETA3: prior_perturb(sp.pri->eta3_prior, q(e + 2), rng); break;
    case MV_ETA4: perturb_log_of(sp.pri->log_eta4_prior, q(e + 3), rng); break;
    case MV_SIGMA_MA: prior_perturb(sp.pri->sigmaMA_prior, q(sp.L.row_MA + 0), rng); break;
    case MV_TAU_MA: prior_perturb(sp.pri->tauMA_prior, q(sp.L.row_MA + 1), rng); break;
    case MV_JITTER:
        perturb_jitters(sp, q, rng);
        // the white-noise branch moves nu and the known objects with the jitters, the GP branch the known
        // objects only, the MA branch neither (the three branches of RVmodel::perturb)
        if (!sp.F.GP && !sp.F.MA && sp.F.studentt) prior_perturb(sp.pri->nu_prior, q(sp.L.row_nu), rng);
        if (!sp.F.MA && sp.F.known_object) perturb_known_objects(sp, q, rng);
        break;
    case MV_BACKGROUND: perturb_background(sp, q, rng, sp.F.GP || sp.F.MA); break;
    default: break;
    }
    return 0.;
}
KS_FN double perturb_model(const SamplerParams& sp, const Col& q, int& np, Rng& rng)
{
    const int move = move_choose(sp, rng);
    return move_apply(sp, move, q, np, rng);
}

// ---- one MCMC step of a lane, split around the likelihood ---------------------------------------
// mcmc_thread :146-158 + update_particle :168-190 up to the likelihood call
KS_FN int step_propose_begin(const SamplerParams& sp, int lane)
{
    Rng rng{sp.key0 + (uint64_t)lane, sp.ctr[lane]};
    const int which = lane * sp.P + rng.rand_int(sp.P);
    const bool particle_first = rng.rand() <= 0.5;
    if (!particle_first) update_level_assignment(sp, rng, lane, which);
    const Col q{sp.Q + lane, sp.ldq};
    const double* s = sp.S + which;
    for (int r = 0; r < sp.SL.n_state; r++) q(r) = s[(size_t)r * sp.ldp];
    sp.npq[lane] = sp.np[which];
    const int move = move_choose(sp, rng);
    sp.which[lane] = which;
    sp.flags[lane] = particle_first ? 1 : 0;
    sp.ctr[lane] = rng.ctr;
    return move;
}
// the move itself, the prior-ratio test (:186-190); any thread of the CTA may finish a lane
KS_FN void step_propose_finish(const SamplerParams& sp, int lane, int move)
{
    Rng rng{sp.key0 + (uint64_t)lane, sp.ctr[lane]};
    const Col q{sp.Q + lane, sp.ldq};
    int np = sp.npq[lane];
    double log_H = move_apply(sp, move, q, np, rng);
    sp.npq[lane] = np;
    if (log_H > 0.0) log_H = 0.0;
    const bool scored = rng.rand() <= exp(log_H);
    sp.flags[lane] |= scored ? 2 : 0;
    sp.ctr[lane] = rng.ctr;
}
KS_FN void step_propose(const SamplerParams& sp, int lane)
{
    const int move = step_propose_begin(sp, lane);
    step_propose_finish(sp, lane, move);
}

// update_particle :191-221, then the level assignment and the `above` list (mcmc_thread :153-163)
KS_FN void step_accept(const SamplerParams& sp, int lane)
{
    Rng rng{sp.key0 + (uint64_t)lane, sp.ctr[lane]};
    const int which = sp.which[lane];
    const int fl = sp.flags[lane];
    const int lvl = sp.level[which];
    if (fl & 2) {
        double value = sp.logl_q[lane];
        if (!(fabs(value) <= 1.7976931348623157e308)) value = -1E300;  // inf or NaN, LikelihoodType.cpp:20-21
        double tbq = sp.tb[which];
        tbq += rng.randh();  // LikelihoodType::perturb
        wrap(tbq, 0., 1.);
        if (lt(sp.lv_logl[lvl], sp.lv_tb[lvl], value, tbq)) {
            const Col q{sp.Q + lane, sp.ldq};
            double* s = sp.S + which;
            for (int r = 0; r < sp.SL.n_state; r++) s[(size_t)r * sp.ldp] = q(r);
            sp.np[which] = sp.npq[lane];
            sp.logl[which] = value;
            sp.tb[which] = tbq;
            counter(sp, 0, lvl, lane) += 1;
        }
    }
    counter(sp, 1, lvl, lane) += 1;
    const double lv = sp.logl[which], lt_ = sp.tb[which];
    for (int j = lvl; j < sp.n_levels - 1; ++j) {
        counter(sp, 2, j, lane) += 1;
        if (lt(sp.lv_logl[j + 1], sp.lv_tb[j + 1], lv, lt_)) counter(sp, 3, j, lane) += 1;
        else break;
    }
    if (fl & 1) update_level_assignment(sp, rng, lane, which);
    const int top = sp.n_levels - 1;
    if (!sp.enough && lt(sp.lv_logl[top], sp.lv_tb[top], lv, lt_)) {
        const int n = sp.n_above[lane];
        if (n < sp.max_steps) {
            double* a = sp.above + ((size_t)lane * sp.max_steps + n) * 2;
            a[0] = lv;
            a[1] = lt_;
            sp.n_above[lane] = n + 1;
        }
    }
    sp.ctr[lane] = rng.ctr;
}

// Host-side twin of the reference's constructors: the derived constants of a prior
// (LogUniform.cpp:17,62; Gaussian.cpp:49-50; Cauchy.cpp:50-51; Exponential.cpp:57-58;
// Rayleigh.cpp:49-52).
inline int prior_make(int type, const double* p, int n_p, kc_prior* out)
{
    static const int nargs[KC_PRIOR_N_TYPES] = {1, 2, 2, 2, 2, 4, 2, 4, 1, 3, 2, 2, 1, 3, 3, 2, 4};
    if (type < 0 || type >= KC_PRIOR_N_TYPES || !out || n_p != nargs[type]) return 1;
    kc_prior q{};
    q.type = type;
    for (int i = 0; i < n_p; i++) q.p[i] = p[i];
    const double a = q.p[0], b = q.p[1], c = q.p[2], d = q.p[3];
    switch (type) {
    case KC_PRIOR_UNIFORM: if (b <= a) return 1; break;
    case KC_PRIOR_LOGUNIFORM: if (a <= 0.0 || b <= a) return 1; q.k[0] = log(b / a); break;
    case KC_PRIOR_MODLOGUNIFORM: if (a <= 0.0 || b <= a) return 1; q.k[0] = 1.0 / log(1.0 + b / a); break;
    case KC_PRIOR_GAUSSIAN: case KC_PRIOR_CAUCHY: case KC_PRIOR_LAPLACE: if (b <= 0.0) return 1; break;
    case KC_PRIOR_TRUNC_GAUSSIAN:
        if (b <= 0.0 || c >= d) return 1;
        q.k[0] = gauss_cdf(a, b, c);
        q.k[1] = gauss_cdf(a, b, d) - gauss_cdf(a, b, c);
        break;
    case KC_PRIOR_TRUNC_CAUCHY:
        if (b <= 0.0 || c >= d) return 1;
        q.k[0] = cauchy_cdf(a, b, c);
        q.k[1] = cauchy_cdf(a, b, d) - cauchy_cdf(a, b, c);
        break;
    case KC_PRIOR_EXPONENTIAL: case KC_PRIOR_RAYLEIGH: if (a <= 0.0) return 1; break;
    case KC_PRIOR_TRUNC_EXPONENTIAL:
        if (a <= 0.0 || b >= c) return 1;
        q.k[0] = expo_cdf(a, b);
        q.k[1] = expo_cdf(a, c) - expo_cdf(a, b);
        break;
    case KC_PRIOR_KUMARASWAMY: if (a <= 0.0 || b <= 0.0) return 1; break;
    case KC_PRIOR_TRUNC_RAYLEIGH:
        if (a <= 0.0 || b >= c) return 1;
        q.k[0] = 1. - exp(-(b * b) / (2. * a * a));
        q.k[1] = (1. - exp(-(c * c) / (2. * a * a))) - q.k[0];
        q.k[2] = log(q.k[1]);
        break;
    case KC_PRIOR_TRIANGULAR: if (c < b || a > b || c <= a) return 1; break;
    case KC_PRIOR_PARETO: if (a <= 0.0 || b <= 0.0) return 1; break;
    case KC_PRIOR_MIX_GAUSS_LOGUNIFORM:
        if (b <= 0.0 || c <= 0.0 || d <= c) return 1;  // Gaussian / LogUniform constructors
        q.k[0] = log(d / c);
        break;
    default: break;
    }
    *out = q;
    return 0;
}

}  // namespace smp
}  // namespace kc
