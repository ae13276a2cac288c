// Host instantiation of the device math headers, for CPU-side unit tests only
// (tests/test_device_math_host.py).  It exports scalar helpers, not a
// likelihood path: there is deliberately no way to score a particle with it.
#include "kc_kepler.cuh"

extern "C" {
void hc_kepler_E(int n, const double* M, const double* e, double* E)
{
    for (int i = 0; i < n; i++) E[i] = kc::kepler_E(M[i], e[i]);
}
void hc_true_anomaly(int n, const double* t, const double* P, const double* e,
                     const double* Tp, double* f)
{
    for (int i = 0; i < n; i++) f[i] = kc::true_anomaly(t[i], P[i], e[i], Tp[i]);
}
// v = K (cos(f+w) + e cos w) for planets given as arrays, at times t, mode 0/1
void hc_kepler_term(int mode, int n, const double* t, const double* P, const double* K,
                    const double* phi, const double* e, const double* w, double M0,
                    double* v)
{
    for (int i = 0; i < n; i++) {
        kc::PlanetInv p = kc::planet_invariants(P[i], K[i], phi[i], e[i], w[i], M0);
        v[i] = mode ? kc::kepler_term<1>(p, t[i]) : kc::kepler_term<0>(p, t[i]);
    }
}
void hc_mod_two_pi(int n, const double* M, double* r)
{
    for (int i = 0; i < n; i++) r[i] = kc::mod_two_pi(M[i]);
}
void hc_exp_dd(int n, const double* x, double* y)
{
    for (int i = 0; i < n; i++) y[i] = kc::exp_dd(x[i]);
}
void hc_sincos(int which, int n, const double* x, double* s, double* c)
{
    for (int i = 0; i < n; i++) {
        if (which == 0) kc::sincos_fast(x[i], s[i], c[i]);
        else if (which == 1) kc::sincos_small(x[i], s[i], c[i]);
        else kc::sincos_tiny(x[i], s[i], c[i]);
    }
}
void hc_misc(int which, int n, const double* x, const double* y, double* r)
{
    for (int i = 0; i < n; i++) {
        if (which == 0) r[i] = kc::sqrt_fast(x[i]);
        else if (which == 1) r[i] = kc::pow23_fast(x[i]);
        else if (which == 2) r[i] = kc::rcp_nr2(x[i]);
        else if (which == 3) r[i] = kc::div_fast(x[i], y[i]);
        else if (which == 4) r[i] = kc::div_by(x[i], y[i], 1.0 / y[i]);
        else if (which == 6) r[i] = kc::exp_nonpos(x[i]);
        else r[i] = kc::rcp_nr1(x[i]);
    }
}
}
