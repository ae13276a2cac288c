// Kepler's equation and the Keplerian RV term, per (particle, planet, epoch).
//
// Replaces kepler::kepler / kepler::true_anomaly and the inner loop of
// RVmodel::calculate_mu (/root/reference/src/RVmodel.cpp:367-374, 1380-1475).
// The solver is the reference's: Markley (1995) cubic starter + one high-order
// refinement, no iteration.  What differs is how it is evaluated:
//   * everything that depends only on the planet is hoisted into PlanetInv
//     (one lane computes it once per particle, with IEEE div/sqrt so that
//     n = 2pi/P and Tp match the reference bit for bit);
//   * fmod, sqrt, x^(2/3), sin/cos and the divisions use kc_math.cuh;
//   * mode 0 ("algebraic") never forms the true anomaly:
//       cos(f+w) = [ (cosE - e) cos w - sqrt(1-e^2) sinE sin w ] / (1 - e cosE)
//     which is the same function as acos((cosE-e)/(1-e cosE)) followed by
//     cos(f+w), without the acos (ill-conditioned at f = 0, pi) and two cosines;
//   * mode 1 ("chain") keeps the reference's operation chain, acos included.
#pragma once
#include "kc_math.cuh"

namespace kc {

// dd arithmetic for the one place that needs more than double: P = exp(log P)
// under hyperpriors (src/RVmodel.cpp:357-360).  A 1-ulp difference in P against
// glibc's exp would shift the mean anomaly by M * 1e-16, i.e. up to 1e-10 rad
// for short periods, so exp is evaluated in double-double and rounded once.
struct dd { double hi, lo; };
KC_HD dd two_sum(double a, double b)
{
    double s = kadd(a, b);
    double bb = kadd(s, -a);
    double e = kadd(kadd(a, -kadd(s, -bb)), kadd(b, -bb));
    return {s, e};
}
KC_HD dd quick_two_sum(double a, double b)
{
    double s = kadd(a, b);
    return {s, kadd(b, -kadd(s, -a))};
}
KC_HD dd dd_add(dd a, dd b)
{
    dd s = two_sum(a.hi, b.hi);
    double e = kadd(kadd(s.lo, a.lo), b.lo);
    return quick_two_sum(s.hi, e);
}
KC_HD dd dd_mul(dd a, dd b)
{
    double p = kmul(a.hi, b.hi);
    double e = kfma(a.hi, b.hi, -p);
    e = kfma(a.hi, b.lo, e);
    e = kfma(a.lo, b.hi, e);
    return quick_two_sum(p, e);
}
KC_HD dd dd_mul_d(dd a, double b)
{
    double p = kmul(a.hi, b);
    double e = kfma(a.hi, b, -p);
    e = kfma(a.lo, b, e);
    return quick_two_sum(p, e);
}

// exp(x) rounded from a ~1e-30 double-double value.
KC_HD double exp_dd(double x)
{
    if (!(x > -700.0 && x < 700.0)) return std::exp(x);  // out of any prior's range
    const double LN2_HI = 0.6931471804855391;   // 33 bits: k*LN2_HI exact
    const double LN2_MID = 7.440617110012397e-11;
    const double LN2_LO = -1.3124698417785255e-27;  // ln2 - HI - MID
    const double INV_LN2 = 1.44269504088896338700e+00;
    double kd = rint_magic(x * INV_LN2);
    // r = x - k ln2 in dd
    double r0 = kfma(-kd, LN2_HI, x);  // exact
    dd t = {kmul(-kd, LN2_MID), 0.0};
    t.lo = kfma(-kd, LN2_MID, -t.hi);
    dd r = dd_add({r0, 0.0}, t);
    r = dd_add(r, {kmul(-kd, LN2_LO), 0.0});
    // scale down by 2^8, expm1 by Taylor, square back up
    r.hi *= 1.0 / 256.0; r.lo *= 1.0 / 256.0;
    // e = r + r^2/2 + ... + r^10/10!   (|r| < 1.4e-3)
    const double inv_fact[11] = {1.0, 1.0, 0.5, 1.0 / 6, 1.0 / 24, 1.0 / 120, 1.0 / 720,
                                 1.0 / 5040, 1.0 / 40320, 1.0 / 362880, 1.0 / 3628800};
    // low parts of the non-dyadic reciprocals that matter at 1e-30
    const double inv_fact_lo[11] = {0, 0, 0, 9.25185853854297e-18, 2.3129646346357427e-18,
                                    1.1564823173178714e-19, -5.300543954373577e-20,
                                    1.7209558293420705e-22, 2.1511947866775882e-23,
                                    -1.858393274046472e-22, 2.3767714622250297e-23};
    dd p = {inv_fact[10], 0.0};
    for (int i = 9; i >= 1; i--) {
        p = dd_mul(p, r);
        p = dd_add(p, {inv_fact[i], inv_fact_lo[i]});
    }
    dd e = dd_mul(p, r);
    for (int i = 0; i < 8; i++) {  // (1+e)^2 - 1 = 2e + e^2
        dd e2 = dd_mul(e, e);
        e = dd_add({2.0 * e.hi, 2.0 * e.lo}, e2);
    }
    dd y = dd_add({1.0, 0.0}, e);
    int k = (int)kd;
    // 2^k by exponent arithmetic (|k| <= 1010 here)
    double scale = from_bits((uint64_t)(1023 + k) << 52);
    return kmul(kadd(y.hi, y.lo), scale);
}

// Per-planet invariants.  Layout matters: the kernels keep an array of these in
// shared memory and every lane reads the same planet at the same time.
struct PlanetInv {
    double n;        // 2*pi/P                      (true_anomaly :1459)
    double Tp;       // M0_epoch - (P*phi)/(2*pi)   (calculate_mu :370)
    double e;        // eccentricity
    double ome;      // 1 - e                       (kepler :1435)
    double c1;       // FACTOR2 / (1 + e)           (starter :1387)
    double A;        // K cos w
    double Bs;       // K sqrt(1-e^2) sin w
    double eA;       // K e cos w                   (:372 second term)
    double K, w;     // chain mode only
};

constexpr double MARKLEY_F1 = 3 * PI / (PI - 6 / PI);
constexpr double MARKLEY_F2 = 1.6 / (PI - 6 / PI);

KC_HD PlanetInv planet_invariants(double P, double K, double phi, double e, double w,
                                  double M0_epoch)
{
    PlanetInv v;
    v.n = kdiv(kmul(2.0, PI), P);
    v.Tp = kadd(M0_epoch, -kdiv(kmul(P, phi), kmul(2.0, PI)));
    v.e = e;
    v.ome = kadd(1.0, -e);
    v.c1 = kdiv(MARKLEY_F2, kadd(1.0, e));
    double sw, cw;
    sincos_fast(w, sw, cw);
    v.A = K * cw;
    v.Bs = K * std::sqrt(kfma(-e, e, 1.0)) * sw;
    v.eA = e * v.A;
    v.K = K;
    v.w = w;
    return v;
}

// The reference's E = kepler::kepler(M, ecc) on the folded anomaly.
// In: M in [0, pi] (already wrapped and folded).  Out: E in [0, pi], plus sinE
// and cosE of that E (from the refinement's own sincos and a Taylor step, so no
// second full sincos is needed).
KC_HD double kepler_folded(double M, double e, double ome, double c1, double& sinE,
                           double& cosE)
{
    // Markley starter (:1380-1395)
    double M2 = M * M;
    double alpha = kfma(c1, PI - M, MARKLEY_F1);
    double d = kfma(alpha, e, 3.0 * ome);
    double alphad = alpha * d;
    double r = kfma(3.0 * alphad, d - ome, M2) * M;
    double q = kfma(2.0 * ome, alphad, -M2);
    double q2 = q * q;
    double s = kfma(q2, q, r * r);
    double w = pow23_fast(std::fabs(r) + sqrt_fast(s));
    double den = kfma(w, w, kfma(w, q, q2));
    // E0 = (2 r w / den + M) / d  as one quotient
    double E0 = kfma(M, den, 2.0 * r * w) * rcp_nr1(den * d);

    // refine_estimate (:1397-1413)
    double s0, c0;
    sincos_small(E0, s0, c0);
    double sE = E0 - s0;
    double cE = 1.0 - c0;
    double f_0 = kfma(e, sE, kfma(E0, ome, -M));
    double f_1 = kfma(e, cE, ome);
    double f_2 = e * s0;
    double f_3 = 1.0 - f_1;
    double f3_6 = f_3 * (1.0 / 6.0);
    // d_3 = -f_0 / (f_1 - 0.5 f_0 f_2 / f_1) = -f_0 f_1 / (f_1^2 - 0.5 f_0 f_2)
    double d_3 = -f_0 * f_1 * rcp_nr1(kfma(-0.5 * f_0, f_2, f_1 * f_1));
    double d_4 = -f_0 * rcp_nr1(kfma(d_3 * d_3, f3_6, kfma(0.5 * d_3, f_2, f_1)));
    double d_42 = d_4 * d_4;
    double den5 = kfma(-d_42 * d_4, f_2 * (1.0 / 24.0),
                       kfma(d_42, f3_6, kfma(0.5 * d_4, f_2, f_1)));
    double dE = -f_0 * rcp_nr2(den5);
    double E = E0 + dE;

    double sd, cd;
    sincos_tiny(dE, sd, cd);
    cosE = kfma(c0, cd, -s0 * sd);
    sinE = kfma(s0, cd, c0 * sd);
    return E;
}

// E = kepler::kepler(M, ecc) for any M (test hook and chain mode).
KC_HD double kepler_E(double M, double e)
{
    double Mw = mod_two_pi(M);
    bool high = Mw > PI;
    if (high) Mw = TWO_PI - Mw;
    double ome = kadd(1.0, -e);
    double c1 = kdiv(MARKLEY_F2, kadd(1.0, e));
    double sE, cE;
    double E = kepler_folded(Mw, e, ome, c1, sE, cE);
    if (high) E = TWO_PI - E;
    return E;
}

// One Keplerian term v = K (cos(f + w) + e cos w) at time t (:367-373).
template <int MODE>
KC_HD double kepler_term(const PlanetInv& p, double t)
{
    double M = kmul(p.n, kadd(t, -p.Tp));  // same two roundings as the reference
    double Mw = mod_two_pi(M);
    bool high = Mw > PI;
    if (high) Mw = TWO_PI - Mw;
    double sinE, cosE;
    double E = kepler_folded(Mw, p.e, p.ome, p.c1, sinE, cosE);
    if (MODE == 0) {
        if (high) sinE = -sinE;
        double rden = rcp_nr2(kfma(-p.e, cosE, 1.0));
        double num = kfma(p.A, cosE - p.e, -p.Bs * sinE);
        return kfma(num, rden, p.eA);
    } else {
        // the reference's chain: cos(E), acos, unfold, cos(f+w)   (:1466-1471, :372)
        if (high) E = TWO_PI - E;
        double sE2, cE2;
        sincos_fast(E, sE2, cE2);
        double arg = div_fast(cE2 - p.e, kfma(-p.e, cE2, 1.0));
        double f = std::acos(arg);
        if (E > PI) f = TWO_PI - f;
        double sfw, cfw;
        sincos_fast(f + p.w, sfw, cfw);
        double sw, cw;
        sincos_fast(p.w, sw, cw);
        return p.K * (cfw + p.e * cw);
    }
}

// f = kepler::true_anomaly(t, P, ecc, Tp) (:1457-1475), test hook.
KC_HD double true_anomaly(double t, double P, double e, double Tp)
{
    double n = kdiv(kmul(2.0, PI), P);
    double M = kmul(n, kadd(t, -Tp));
    double E = kepler_E(M, e);
    double s, c;
    sincos_fast(E, s, c);
    double f = std::acos(div_fast(c - e, kfma(-e, c, 1.0)));
    if (E > PI) f = TWO_PI - f;
    return f;
}

}  // namespace kc
