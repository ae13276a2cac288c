// fp64 building blocks of the Kepler / GP kernels.
//
// Everything here is written against the FP64 pipe budget measured on B200
// (profiles/FP64_PEAKS.json): CUDA's generic fmod / pow / log / sin cost 100 /
// 126 / 49 / 33 DFMA slots per call, which would make the Kepler term ~580
// slots.  The versions below use MUFU seeds + Newton steps and argument ranges
// that the Kepler chain guarantees, for ~135 slots per term.
//
// The header also compiles for the host (g++), with the MUFU seeds emulated,
// so tests/ can check the arithmetic against the reference without a GPU; the
// product only ever runs the device instantiation.
#pragma once
#include <cmath>
#include <cstdint>
#include <cstring>

#if defined(__CUDACC__)
#define KC_HD __host__ __device__ __forceinline__
#else
#define KC_HD inline
#endif

namespace kc {

constexpr double PI = 3.14159265358979323846;       // M_PI
constexpr double TWO_PI = 6.28318530717958647692;   // 2*M_PI (exact doubling)
constexpr double INV_TWO_PI = 0.15915494309189533577;
constexpr double TWO_OVER_PI = 0.63661977236758134308;
constexpr double MAGIC = 6755399441055744.0;        // 1.5 * 2^52
// pi/2 split for Cody-Waite reduction (the usual three-term split)
constexpr double PIO2_HI = 1.57079632679489655800e+00;
constexpr double PIO2_MID = 6.12323399573676603587e-17;
constexpr double PIO2_LO = -1.49738490485916983e-33;

// ---- exact-rounding primitives ------------------------------------------------
KC_HD double kfma(double a, double b, double c)
{
#if defined(__CUDA_ARCH__)
    return __fma_rn(a, b, c);
#else
    return __builtin_fma(a, b, c);
#endif
}
KC_HD double kadd(double a, double b)
{
#if defined(__CUDA_ARCH__)
    return __dadd_rn(a, b);  // never contracted into an fma
#else
    volatile double r = a + b;
    return r;
#endif
}
KC_HD double kmul(double a, double b)
{
#if defined(__CUDA_ARCH__)
    return __dmul_rn(a, b);
#else
    volatile double r = a * b;
    return r;
#endif
}
KC_HD double kdiv(double a, double b)
{
#if defined(__CUDA_ARCH__)
    return __ddiv_rn(a, b);
#else
    volatile double r = a / b;
    return r;
#endif
}

KC_HD double from_bits(uint64_t u)
{
#if defined(__CUDA_ARCH__)
    return __longlong_as_double((long long)u);
#else
    double d; std::memcpy(&d, &u, 8); return d;
#endif
}
KC_HD uint64_t to_bits(double d)
{
#if defined(__CUDA_ARCH__)
    return (uint64_t)__double_as_longlong(d);
#else
    uint64_t u; std::memcpy(&u, &d, 8); return u;
#endif
}

// Round to nearest integer (ties to even), valid for |x| < 2^51.
KC_HD double rint_magic(double x) { return kadd(kadd(x, MAGIC), -MAGIC); }
// Same, and also the integer (low 32 bits) read straight from the sum.
KC_HD double rint_magic_i(double x, int& k)
{
    double t = kadd(x, MAGIC);
    k = (int)(uint32_t)to_bits(t);
    return kadd(t, -MAGIC);
}

// ---- MUFU seeds -----------------------------------------------------------------
// rcp.approx.ftz.f64 / rsqrt.approx.ftz.f64: ~2^-20 relative, low word zero.
KC_HD double rcp_seed(double x)
{
#if defined(__CUDA_ARCH__)
    double r;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(x));
    return r;
#else
    return from_bits(to_bits(1.0 / x) & 0xFFFFFFFF00000000ull);
#endif
}
KC_HD double rsqrt_seed(double x)
{
#if defined(__CUDA_ARCH__)
    double r;
    asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(x));
    return r;
#else
    return from_bits(to_bits(1.0 / std::sqrt(x)) & 0xFFFFFFFF00000000ull);
#endif
}

// 1/x with one Newton step (~2^-40) or two (~1 ulp).  No special-case handling:
// callers guarantee a finite, normal, non-zero x (or accept NaN out).
KC_HD double rcp_nr1(double x)
{
    double r = rcp_seed(x);
    double e = kfma(-x, r, 1.0);
    return kfma(r, e, r);
}
KC_HD double rcp_nr2(double x)
{
    double r = rcp_seed(x);
    double e = kfma(-x, r, 1.0);
    r = kfma(r, e, r);
    e = kfma(-x, r, 1.0);
    return kfma(r, e, r);
}
// a/b to <= 1 ulp (correctly rounded except in rare boundary cases).
KC_HD double div_fast(double a, double b)
{
    double r = rcp_nr2(b);
    double q = a * r;
    double rem = kfma(-q, b, a);
    return kfma(rem, r, q);
}
// x/d given rd = correctly rounded 1/d: Markstein's final step, correctly
// rounded for all but pathological mantissas of d.  3 slots instead of ~15.
KC_HD double div_by(double x, double d, double rd)
{
    double q = x * rd;
    double rem = kfma(-q, d, x);
    return kfma(rem, rd, q);
}

// sqrt(s), s > 0 finite: Goldschmidt from the rsqrt seed, then one correction.
KC_HD double sqrt_fast(double s)
{
    double y = rsqrt_seed(s);
    double g = s * y, h = 0.5 * y;
    double e = kfma(-h, g, 0.5);
    g = kfma(g, e, g);
    h = kfma(h, e, h);
    e = kfma(-h, g, 0.5);
    g = kfma(g, e, g);
    h = kfma(h, e, h);
    double rem = kfma(-g, g, s);
    return kfma(rem, h, g);
}

// x^(2/3) for x in the normal float range (the Markley starter's argument is
// within [1e-30, 1e5]).  fp32 MUFU seed of x^(-1/3), one cubic correction.
KC_HD double pow23_fast(double x)
{
#if defined(__CUDA_ARCH__)
    float zf = exp2f(-0.333333333f * __log2f((float)x));
#else
    float zf = 1.0f / cbrtf((float)x);
#endif
    double z = (double)zf;
    double z2 = z * z;
    double z3 = z2 * z;
    double e = kfma(-x, z3, 1.0);
    double p = kfma(e, 2.0 / 9.0, 1.0 / 3.0);
    double ze = z * e;
    z = kfma(ze, p, z);
    return x * z;
}

// ---- fmod(M, 2*pi) followed by numpy's sign fix-up -----------------------------
// kepler::npy_mod(M, two_pi), /root/reference/src/RVmodel.cpp:1353-1378, bit for
// bit for |M| < 2^51: the IEEE remainder r = M - rint(M/b)*b is exactly
// representable and the fma delivers it without rounding; fmod's result is r or
// r +/- b (also exact), and npy_mod's "+= b" for negative results is the same
// floating-point add performed here.
KC_HD double mod_two_pi(double M)
{
    double q = rint_magic(M * INV_TWO_PI);
    double r = kfma(-q, TWO_PI, M);
    return (r < 0.0) ? kadd(r, TWO_PI) : r;
}

// ---- exp(x) for x <= 0 (the GP kernel's argument), branch-free ---------------------
// k = rint(x log2 e), r = x - k ln2 (two-term Cody-Waite, exact for |k| < 2^11 with the
// 42-bit high part), exp(r) by its Taylor series to r^13 (|r| <= 0.3466: truncation
// 4e-18), scaled by 2^k in two exact-until-the-last steps so that results below 2^-1022
// round once into the subnormals.  ~0.6 ulp (measured against long double in tests/); x < -746 gives 0, NaN gives NaN.
KC_HD double exp_nonpos(double x)
{
    const double LOG2E = 1.44269504088896338700e+00;
    const double LN2_HI = 6.93147180559890330187e-01;   // 42 significant bits
    const double LN2_LO = 5.49792301870837115524e-14;
    double xc = x < -746.0 ? -746.0 : x;  // NaN falls through
    int k;
    double kd = rint_magic_i(xc * LOG2E, k);
    double rh = kfma(-kd, LN2_HI, xc);             // exact
    double rl = -kd * LN2_LO;
    double r = rh + rl;
    double p = 1.0 / 6227020800.0;                 // 1/13!
    p = kfma(p, r, 1.0 / 479001600.0);
    p = kfma(p, r, 1.0 / 39916800.0);
    p = kfma(p, r, 1.0 / 3628800.0);
    p = kfma(p, r, 1.0 / 362880.0);
    p = kfma(p, r, 1.0 / 40320.0);
    p = kfma(p, r, 1.0 / 5040.0);
    p = kfma(p, r, 1.0 / 720.0);
    p = kfma(p, r, 1.0 / 120.0);
    p = kfma(p, r, 1.0 / 24.0);
    p = kfma(p, r, 1.0 / 6.0);
    p = kfma(p, r, 0.5);
    double r2 = r * r;
    // exp(r) = (1 + r) + r^2 (1/2 + ...) with the rounding error of 1 + r carried along,
    // so the only full-size rounding is the last addition
    double t1 = kadd(1.0, rh);
    double lo = kadd(kadd(rh, -kadd(t1, -1.0)), rl);
    p = kadd(t1, kfma(p, r2, lo));
    // 2^(k+64) is a normal number for every k >= -1077; the final 2^-64 rounds once
    double s1 = from_bits((uint64_t)(int64_t)(k + 64 + 1023) << 52);
    double y = (p * s1) * 5.42101086242752217004e-20;  // 2^-64
    return x < -746.0 ? 0.0 : y;
}

// ---- sin and cos together -------------------------------------------------------
// fdlibm kernel polynomials on [-pi/4, pi/4]; ~1 ulp.
KC_HD void sincos_kernel(double r, double& s, double& c)
{
    const double S1 = -1.66666666666666324348e-01, S2 = 8.33333333332248946124e-03,
                 S3 = -1.98412698298579493134e-04, S4 = 2.75573137070700676789e-06,
                 S5 = -2.50507602534068634195e-08, S6 = 1.58969099521155010221e-10;
    const double C1 = 4.16666666666666019037e-02, C2 = -1.38888888888741095749e-03,
                 C3 = 2.48015872894767294178e-05, C4 = -2.75573143513906633035e-07,
                 C5 = 2.08757232129817482790e-09, C6 = -1.13596475577881948265e-11;
    double z = r * r;
    double ps = kfma(z, S6, S5);
    double pc = kfma(z, C6, C5);
    ps = kfma(z, ps, S4); pc = kfma(z, pc, C4);
    ps = kfma(z, ps, S3); pc = kfma(z, pc, C3);
    ps = kfma(z, ps, S2); pc = kfma(z, pc, C2);
    ps = kfma(z, ps, S1); pc = kfma(z, pc, C1);
    double rz = r * z;
    s = kfma(rz, ps, r);
    double zz = z * z;
    double hz = kfma(-0.5, z, 1.0);
    c = kfma(zz, pc, hz);
}

// sin(x), cos(x) for |x| up to ~1e5 (three-term Cody-Waite reduction).
KC_HD void sincos_fast(double x, double& s, double& c)
{
    int k;
    double kd = rint_magic_i(x * TWO_OVER_PI, k);
    double r = kfma(-kd, PIO2_HI, x);
    r = kfma(-kd, PIO2_MID, r);
    r = kfma(-kd, PIO2_LO, r);
    double ss, cc;
    sincos_kernel(r, ss, cc);
    double a = (k & 1) ? cc : ss;
    double b = (k & 1) ? ss : cc;
    s = (k & 2) ? -a : a;
    c = ((k + 1) & 2) ? -b : b;
}

// Same for 0 <= x <~ 8 (two reduction terms suffice for |k| <= 5).
KC_HD void sincos_small(double x, double& s, double& c)
{
    int k;
    double kd = rint_magic_i(x * TWO_OVER_PI, k);
    double r = kfma(-kd, PIO2_HI, x);
    r = kfma(-kd, PIO2_MID, r);
    double ss, cc;
    sincos_kernel(r, ss, cc);
    double a = (k & 1) ? cc : ss;
    double b = (k & 1) ? ss : cc;
    s = (k & 2) ? -a : a;
    c = ((k + 1) & 2) ? -b : b;
}

// sin(d), cos(d) for |d| <= 0.03 (the Kepler refinement step): Taylor to d^7.
KC_HD void sincos_tiny(double d, double& s, double& c)
{
    double z = d * d;
    double ps = kfma(z, -1.0 / 5040.0, 1.0 / 120.0);
    double pc = kfma(z, -1.0 / 720.0, 1.0 / 24.0);
    ps = kfma(z, ps, -1.0 / 6.0);
    pc = kfma(z, pc, -0.5);
    double dz = d * z;
    s = kfma(dz, ps, d);
    c = kfma(z, pc, 1.0);
}

}  // namespace kc
