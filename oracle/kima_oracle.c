/*
 * TEST INFRASTRUCTURE — not part of the product.  Only tests/, smoke() and
 * bench.py's cpu_baseline / --impl reference legs may load this.
 *
 * Plain-C restatement of kima's per-particle likelihood hot path, function by
 * function, each citing the reference lines it follows.  It exists so that the
 * GPU parity tests have a checker on machines where /root/reference is absent
 * (the GPU box).  PARITY PINNED: tests/test_oracle_vs_reference.py checks this
 * file against oracle/_ref/libkima_ref.so (the unmodified reference sources
 * compiled by oracle/Makefile) and against the committed golden vectors in
 * tests/golden/ that were generated from that library.
 *
 * Arithmetic notes: `mu` is accumulated in `long double` as in the reference
 * (src/RVmodel.h:79-80); the Cholesky follows Eigen's blocked right-looking
 * LLT (block-size rule and unblocked kernel) but not Eigen's SIMD summation
 * order, so GP values agree with the reference to rounding, not bit-for-bit.
 */
#include <math.h>
#include <pthread.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

#include "kima_cuda.h" /* kc_model_flags / kc_layout: the shared layout rule */

/* ------------------------------------------------------------------------ */
/* namespace kepler, /root/reference/src/RVmodel.cpp:1339-1477               */
/* ------------------------------------------------------------------------ */

/* kepler::npy_mod, src/RVmodel.cpp:1353-1378 */
static double ko_npy_mod(double a, double b)
{
    double mod = fmod(a, b);
    if (!b) return mod;
    if (mod) {
        if ((b < 0) != (mod < 0)) mod += b;
    } else {
        mod = copysign(0, b);
    }
    return mod;
}

/* kepler::get_markley_starter, src/RVmodel.cpp:1380-1395 */
static double ko_markley_starter(double M, double ecc, double ome)
{
    const double FACTOR1 = 3 * M_PI / (M_PI - 6 / M_PI);
    const double FACTOR2 = 1.6 / (M_PI - 6 / M_PI);
    double M2 = M * M;
    double alpha = FACTOR1 + FACTOR2 * (M_PI - M) / (1 + ecc);
    double d = 3 * ome + alpha * ecc;
    double alphad = alpha * d;
    double r = (3 * alphad * (d - ome) + M2) * M;
    double q = 2 * alphad * ome - M2;
    double q2 = q * q;
    double w = pow(fabs(r) + sqrt(q2 * q + r * r), 2.0 / 3);
    return (2 * r * w / (w * w + w * q + q2) + M) / d;
}

/* kepler::refine_estimate, src/RVmodel.cpp:1397-1413 */
static double ko_refine_estimate(double M, double ecc, double ome, double E)
{
    double sE = E - sin(E);
    double cE = 1 - cos(E);
    double f_0 = ecc * sE + E * ome - M;
    double f_1 = ecc * cE + ome;
    double f_2 = ecc * (E - sE);
    double f_3 = 1 - f_1;
    double d_3 = -f_0 / (f_1 - 0.5 * f_0 * f_2 / f_1);
    double d_4 = -f_0 / (f_1 + 0.5 * d_3 * f_2 + (d_3 * d_3) * f_3 / 6);
    double d_42 = d_4 * d_4;
    double dE = -f_0 / (f_1 + 0.5 * d_4 * f_2 + d_4 * d_4 * f_3 / 6 - d_42 * d_4 * f_2 / 24);
    return E + dE;
}

/* kepler::kepler, src/RVmodel.cpp:1422-1445 */
double ko_kepler(double M, double ecc)
{
    const double two_pi = 2 * M_PI;
    M = ko_npy_mod(M, two_pi);
    int high = M > M_PI;
    if (high) M = two_pi - M;
    double ome = 1.0 - ecc;
    double E = ko_markley_starter(M, ecc, ome);
    E = ko_refine_estimate(M, ecc, ome, E);
    if (high) E = two_pi - E;
    return E;
}

/* kepler::true_anomaly, src/RVmodel.cpp:1457-1475 */
double ko_true_anomaly(double t, double period, double ecc, double t_peri)
{
    double n = 2. * M_PI / period;
    double M = n * (t - t_peri);
    double E = ko_kepler(M, ecc);
    double cosE = cos(E);
    double f = acos((cosE - ecc) / (1 - ecc * cosE));
    if (E > M_PI) f = 2 * M_PI - f;
    return f;
}

void ko_kepler_n(int n, const double* M, const double* ecc, double* E)
{
    for (int i = 0; i < n; i++) E[i] = ko_kepler(M[i], ecc[i]);
}
void ko_true_anomaly_n(int n, const double* t, const double* P, const double* ecc,
                       const double* Tp, double* f)
{
    for (int i = 0; i < n; i++) f[i] = ko_true_anomaly(t[i], P[i], ecc[i], Tp[i]);
}

/* ------------------------------------------------------------------------ */
/* Data singleton (src/Data.h:14-16,52-108) and flags                        */
/* ------------------------------------------------------------------------ */
typedef struct {
    int N, n_inst, n_ind;
    double *t, *y, *sig;
    int* obsi;
    double* actind; /* [n_ind][N] */
    double M0_epoch, t_middle;
} ko_data;

static ko_data g_data;
static kc_model_flags g_flags;
static kc_layout g_layout;

static void* ko_dup(const void* p, size_t n)
{
    void* q = malloc(n ? n : 1);
    if (p && n) memcpy(q, p, n);
    return q;
}

/* M0_epoch = t[0] (src/Data.cpp:162,280,461); t_middle = t_min + 0.5*(t_max -
 * t_min) (src/Data.h:59-66). */
int ko_set_data(int N, const double* t, const double* y, const double* sig,
                const int* obsi, int n_inst, int n_ind, const double* actind)
{
    free(g_data.t); free(g_data.y); free(g_data.sig); free(g_data.obsi);
    free(g_data.actind);
    g_data.N = N; g_data.n_inst = n_inst; g_data.n_ind = n_ind;
    g_data.t = ko_dup(t, sizeof(double) * N);
    g_data.y = ko_dup(y, sizeof(double) * N);
    g_data.sig = ko_dup(sig, sizeof(double) * N);
    g_data.obsi = malloc(sizeof(int) * (N ? N : 1));
    for (int i = 0; i < N; i++) g_data.obsi[i] = obsi ? obsi[i] : 1;
    g_data.actind = ko_dup(actind, sizeof(double) * N * n_ind);
    g_data.M0_epoch = t[0];
    double tmin = t[0], tmax = t[0];
    for (int i = 1; i < N; i++) {
        if (t[i] < tmin) tmin = t[i];
        if (t[i] > tmax) tmax = t[i];
    }
    g_data.t_middle = tmin + 0.5 * (tmax - tmin);
    kc_layout_from_flags(&g_flags, g_data.n_inst, g_data.n_ind, &g_layout);
    return N;
}

int ko_set_flags(const kc_model_flags* f)
{
    g_flags = *f;
    kc_layout_from_flags(&g_flags, g_data.n_inst, g_data.n_ind, &g_layout);
    return 0;
}

/* ------------------------------------------------------------------------ */
/* RVmodel::calculate_mu, src/RVmodel.cpp:286-395 (from-scratch branch),      */
/* add_known_object :416-432, MA(1) term :378-386.                           */
/* ------------------------------------------------------------------------ */
#define AT(row) (params[(size_t)(row) * ld + b])

void ko_calculate_mu(const double* params, int ld, int b, int np, long double* mu)
{
    const ko_data* D = &g_data;
    const kc_layout* L = &g_layout;
    const int N = D->N;
    const double* t = D->t;

    double background = AT(L->row_vsys);
    for (int i = 0; i < N; i++) mu[i] = background; /* :309 */

    if (g_flags.trend) { /* :311-318 */
        double slope = AT(L->row_trend), quadr = AT(L->row_trend + 1),
               cubic = AT(L->row_trend + 2);
        double tmid = D->t_middle;
        for (int i = 0; i < N; i++)
            mu[i] += slope * (t[i] - tmid) + quadr * pow(t[i] - tmid, 2) +
                     cubic * pow(t[i] - tmid, 3);
    }
    if (g_flags.multi_instrument) { /* :320-329 */
        for (int j = 0; j < L->n_off; j++)
            for (int i = 0; i < N; i++)
                if (D->obsi[i] == j + 1) mu[i] += AT(L->row_off + j);
    }
    if (D->n_ind > 0) { /* :331-338 */
        for (int i = 0; i < N; i++)
            for (int j = 0; j < D->n_ind; j++)
                mu[i] += AT(L->row_beta + j) * D->actind[(size_t)j * N + i];
    }
    if (g_flags.known_object) { /* :340-342 -> :416-432 */
        for (int j = 0; j < L->n_ko; j++) {
            double P = AT(L->row_KO_P + j), K = AT(L->row_KO_K + j),
                   e = AT(L->row_KO_e + j), phi = AT(L->row_KO_phi + j),
                   w = AT(L->row_KO_w + j);
            for (int i = 0; i < N; i++) {
                double ti = t[i];
                double Tp = D->M0_epoch - (P * phi) / (2. * M_PI);
                double f = ko_true_anomaly(ti, P, e, Tp);
                double v = K * (cos(f + w) + e * cos(w));
                mu[i] += v;
            }
        }
    }
    for (int j = 0; j < np; j++) { /* :355-375 */
        double P = g_flags.hyperpriors ? exp(AT(L->row_P + j)) : AT(L->row_P + j);
        double K = AT(L->row_K + j), phi = AT(L->row_phi + j),
               ecc = AT(L->row_ecc + j), omega = AT(L->row_w + j);
        for (int i = 0; i < N; i++) {
            double ti = t[i];
            double Tp = D->M0_epoch - (P * phi) / (2. * M_PI);
            double f = ko_true_anomaly(ti, P, ecc, Tp);
            double v = K * (cos(f + omega) + ecc * cos(omega));
            mu[i] += v;
        }
    }
    if (g_flags.MA) { /* :378-386: the loop starts at the 2nd point */
        double sigmaMA = AT(L->row_MA), tauMA = AT(L->row_MA + 1);
        const double* y = D->y;
        for (int i = 1; i < N; i++)
            mu[i] += sigmaMA * exp(-fabs(t[i - 1] - t[i]) / tauMA) * (y[i - 1] - mu[i - 1]);
    }
}

/* ------------------------------------------------------------------------ */
/* RVmodel::calculate_C (standard kernel), src/RVmodel.cpp:186-216            */
/* C is column-major N x N, full symmetric (Eigen::MatrixXd).                */
/* ------------------------------------------------------------------------ */
void ko_calculate_C(const double* params, int ld, int b, double* C)
{
    const ko_data* D = &g_data;
    const kc_layout* L = &g_layout;
    const int N = D->N;
    const double *t = D->t, *sig = D->sig;
    double eta1 = AT(L->row_eta), eta2 = AT(L->row_eta + 1),
           eta3 = AT(L->row_eta + 2), eta4 = AT(L->row_eta + 3);
    for (int i = 0; i < N; i++) {
        for (int j = i; j < N; j++) {
            double c = eta1 * eta1 * exp(-0.5 * pow((t[i] - t[j]) / eta2, 2) -
                                         2.0 * pow(sin(M_PI * (t[i] - t[j]) / eta3) / eta4, 2));
            if (i == j) {
                double jit = g_flags.multi_instrument ? AT(L->row_jit + D->obsi[i] - 1)
                                                      : AT(L->row_jit);
                c += sig[i] * sig[i] + jit * jit;
                C[(size_t)j * N + i] = c;
            } else {
                C[(size_t)j * N + i] = c;
                C[(size_t)i * N + j] = c;
            }
        }
    }
}

/* ------------------------------------------------------------------------ */
/* Eigen::LLT<MatrixXd,Lower>, eigen/Eigen/src/Cholesky/LLT.h                 */
/* ------------------------------------------------------------------------ */
#define A_(i, j) A[(size_t)(j) * lda + (i)]

/* llt_inplace<Lower>::unblocked, LLT.h:313-337.  Works on the n x n block at
 * A (leading dimension lda).  Returns -1 on success or the failing column. */
static int ko_llt_unblocked(double* A, int lda, int n)
{
    for (int k = 0; k < n; k++) {
        int rs = n - k - 1;
        double x = A_(k, k);
        if (k > 0) {
            double s = 0.0;
            for (int c = 0; c < k; c++) s += A_(k, c) * A_(k, c);
            x -= s;
        }
        if (x <= 0.0) return k;
        A_(k, k) = x = sqrt(x);
        if (k > 0 && rs > 0)
            for (int c = 0; c < k; c++) {
                double a10 = A_(k, c);
                for (int i = k + 1; i < n; i++) A_(i, k) -= A_(i, c) * a10;
            }
        if (rs > 0)
            for (int i = k + 1; i < n; i++) A_(i, k) /= x;
    }
    return -1;
}

/* llt_inplace<Lower>::blocked, LLT.h:339-368 */
int ko_llt_blocked(double* A, int n)
{
    const int lda = n;
    if (n < 32) return ko_llt_unblocked(A, lda, n);
    int blockSize = n / 8;
    blockSize = (blockSize / 16) * 16;
    if (blockSize < 8) blockSize = 8;
    if (blockSize > 128) blockSize = 128;
    for (int k = 0; k < n; k += blockSize) {
        int bs = blockSize < n - k ? blockSize : n - k;
        int rs = n - k - bs;
        double* A11 = &A_(k, k);
        int ret = ko_llt_unblocked(A11, lda, bs);
        if (ret >= 0) return k + ret;
        if (rs > 0) {
            /* A21 <- A21 * A11^-T  (solveInPlace<OnTheRight>, :364) */
            for (int c = 0; c < bs; c++) {
                for (int p = 0; p < c; p++) {
                    double l = A_(k + c, k + p);
                    for (int i = k + bs; i < n; i++) A_(i, k + c) -= A_(i, k + p) * l;
                }
                double d = A_(k + c, k + c);
                for (int i = k + bs; i < n; i++) A_(i, k + c) /= d;
            }
            /* A22 -= A21 * A21^T, lower triangle only (rankUpdate, :365) */
            for (int j = k + bs; j < n; j++)
                for (int p = k; p < k + bs; p++) {
                    double l = A_(j, p);
                    for (int i = j; i < n; i++) A_(i, j) -= A_(i, p) * l;
                }
        }
    }
    return -1;
}

/* LLT::solve = forward then backward substitution (LLT.h:476-484) */
static void ko_llt_solve(const double* A, int n, double* x)
{
    const int lda = n;
    for (int i = 0; i < n; i++) { /* L z = r, column-oriented */
        x[i] /= A_(i, i);
        double xi = x[i];
        for (int r = i + 1; r < n; r++) x[r] -= A_(r, i) * xi;
    }
    for (int i = n - 1; i >= 0; i--) { /* L^T s = z */
        double s = x[i];
        for (int r = i + 1; r < n; r++) s -= A_(r, i) * x[r];
        x[i] = s / A_(i, i);
    }
}

/* ------------------------------------------------------------------------ */
/* kernel = celerite: RVmodel::calculate_C :218-266 (the three complex terms of    */
/* the quasi-periodic approximation) + celerite::solver::CholeskySolver::compute,  */
/* dot_solve, log_determinant (celerite/cpp/include/celerite/solver/cholesky.h     */
/* :41-209, :326-400), J = 6.  Returns dot_solve + log_det; *fail = 1 where the    */
/* reference would throw linalg_exception (Dn < 0).                               */
/* ------------------------------------------------------------------------ */
#define KO_J 6
static double ko_celerite(const double* params, int ld, int b, const double* residual, int* fail)
{
    const ko_data* D = &g_data;
    const kc_layout* L = &g_layout;
    const int N = D->N, J = KO_J;
    const double *x = D->t, *sig = D->sig;
    double eta1 = AT(L->row_eta), eta2 = AT(L->row_eta + 1), eta3 = AT(L->row_eta + 2);
    double extra_sigma = AT(L->row_jit);
    /* :231-237 */
    double wbeat = 1 / eta2, wrot = 2 * M_PI / eta3, amp = eta1 * eta1;
    double c = wbeat, d = wrot, xx = c / d;
    double a = amp / 2, bb = amp * xx / 2, e = amp / 8, f = amp * xx / 4, g = amp * (3. / 8. + 0.001);
    double a_comp[3] = {a, e, g}, b_comp[3] = {bb, f, 0.0}, c_comp[3] = {c, c, c}, d_comp[3] = {d, 2 * d, 0.0};
    double jitter = extra_sigma * extra_sigma;
    double asum = (a_comp[0] + a_comp[1]) + a_comp[2];
    double S[KO_J][KO_J], W0[KO_J], W1[KO_J], phi[KO_J], u[KO_J], fv[KO_J];
    memset(S, 0, sizeof S);
    *fail = 0;
    /* D_ = diag + a_real.sum() + a_comp.sum() + jitter, diag = sig^2 (:254-262, cholesky.h:95) */
    double Dn = ((sig[0] * sig[0] + 0.0) + asum) + jitter;
    {
        double value = 1.0 / Dn, t = x[0];
        for (int j = 0, k = 0; j < 3; ++j, k += 2) {
            double dd = d_comp[j] * t;
            W0[k] = cos(dd) * value;
            W0[k + 1] = sin(dd) * value;
        }
    }
    for (int j = 0; j < J; j++) fv[j] = 0.0;
    double xm1 = residual[0];
    double result = xm1 * (xm1 / Dn); /* cholesky.h:347 */
    double log_det = log(Dn);
    for (int n = 1; n < N; ++n) {
        double t = x[n], dx = t - x[n - 1];
        for (int j = 0, k = 0; j < 3; ++j, k += 2) {
            double aa = a_comp[j], bq = b_comp[j], dd = d_comp[j] * t, cd = cos(dd), sd = sin(dd);
            double value = exp(-c_comp[j] * dx);
            phi[k] = value; phi[k + 1] = value;
            u[k] = aa * cd + bq * sd;
            u[k + 1] = aa * sd - bq * cd;
            W1[k] = cd; W1[k + 1] = sd;
        }
        for (int j = 0; j < J; ++j) {
            double phij = phi[j], xj = Dn * W0[j];
            for (int k = 0; k <= j; ++k) S[k][j] = phij * (phi[k] * (S[k][j] + xj * W0[k]));
        }
        Dn = ((sig[n] * sig[n] + 0.0) + asum) + jitter;
        for (int j = 0; j < J; ++j) {
            double uj = u[j], xj = W1[j];
            for (int k = 0; k < j; ++k) {
                double tmp = u[k] * S[k][j];
                Dn -= 2.0 * (uj * tmp);
                xj -= tmp;
                W1[k] -= uj * S[k][j];
            }
            double tmp = uj * S[j][j];
            Dn -= uj * tmp;
            W1[j] = xj - tmp;
        }
        if (Dn < 0) *fail = 1;
        for (int j = 0; j < J; ++j) W1[j] /= Dn;
        /* dot_solve, forward recursion (cholesky.h:348-357) */
        double xv = residual[n];
        for (int j = 0; j < J; ++j) {
            double value = phi[j] * (fv[j] + W0[j] * xm1);
            fv[j] = value;
            xv -= u[j] * value;
        }
        xm1 = xv;
        result += xv * xv / Dn;
        log_det += log(Dn);
        for (int j = 0; j < J; ++j) W0[j] = W1[j];
    }
    return result + log_det;
}

/* ------------------------------------------------------------------------ */
/* RVmodel::log_likelihood, src/RVmodel.cpp:827-943                           */
/* ------------------------------------------------------------------------ */
typedef struct {
    long double* mu;
    double* C;   /* N*N, GP only */
    double* res; /* N */
} ko_work;

static void ko_work_init(ko_work* w)
{
    int N = g_data.N;
    w->mu = malloc(sizeof(long double) * (N ? N : 1));
    w->res = malloc(sizeof(double) * (N ? N : 1));
    w->C = g_flags.GP ? malloc(sizeof(double) * (size_t)N * N + 8) : NULL;
}
static void ko_work_free(ko_work* w)
{
    free(w->mu); free(w->res); free(w->C);
}

static double ko_log_likelihood(const double* params, int ld, int b, ko_work* w)
{
    const ko_data* D = &g_data;
    const kc_layout* L = &g_layout;
    const int N = D->N;
    const double *y = D->y, *sig = D->sig;
    const long double* mu = w->mu;
    const double halflog2pi = 0.5 * log(2. * M_PI); /* :20 */
    double logL = 0.;

    if (g_flags.GP && g_flags.kernel == 1) { /* :875-882 */
        double* residual = w->res;
        for (int i = 0; i < N; i++) residual[i] = (double)(y[i] - mu[i]);
        int fail = 0;
        double s = ko_celerite(params, ld, b, residual, &fail);
        logL = -0.5 * (s + N * log(2 * M_PI));
        if (fail) logL = INFINITY; /* the reference throws linalg_exception out of calculate_C */
    } else if (g_flags.GP) { /* :841-873 */
        double* residual = w->res;
        for (int i = 0; i < N; i++) residual[i] = (double)(y[i] - mu[i]);
        ko_llt_blocked(w->C, N); /* info() is never checked in the reference (:854) */
        double logDeterminant = 0.;
        for (int i = 0; i < N; i++) logDeterminant += 2. * log(w->C[(size_t)i * N + i]);
        double* solution = malloc(sizeof(double) * (N ? N : 1));
        memcpy(solution, residual, sizeof(double) * N);
        ko_llt_solve(w->C, N, solution);
        double exponent = 0.;
        for (int i = 0; i < N; i++) exponent += residual[i] * solution[i];
        free(solution);
        logL = -0.5 * N * log(2 * M_PI) - 0.5 * logDeterminant - 0.5 * exponent;
    } else if (g_flags.studentt) { /* :891-910 */
        double nu = AT(L->row_nu);
        for (int i = 0; i < N; i++) {
            double jit = g_flags.multi_instrument ? AT(L->row_jit + D->obsi[i] - 1)
                                                  : AT(L->row_jit);
            double var = sig[i] * sig[i] + jit * jit;
            /* y[i]-mu[i] is long double in the reference, so pow/log promote */
            long double r2 = powl((long double)y[i] - mu[i], 2);
            logL += lgamma(0.5 * (nu + 1.)) - lgamma(0.5 * nu) - 0.5 * log(M_PI * nu) -
                    0.5 * log(var) - 0.5 * (nu + 1.) * logl(1. + r2 / var / nu);
        }
    } else { /* :912-929 */
        for (int i = 0; i < N; i++) {
            double jit = g_flags.multi_instrument ? AT(L->row_jit + D->obsi[i] - 1)
                                                  : AT(L->row_jit);
            double var = sig[i] * sig[i] + jit * jit;
            long double r2 = powl((long double)y[i] - mu[i], 2);
            logL += -halflog2pi - 0.5 * log(var) - 0.5 * (r2 / var);
        }
    }
    if (isnan(logL) || isinf(logL)) logL = INFINITY; /* :938-941 */
    return logL;
}

/* calculate_mu -> [calculate_C] -> log_likelihood for one lane */
static double ko_eval_lane(const double* params, int ld, int b, int np, ko_work* w)
{
    ko_calculate_mu(params, ld, b, np, w->mu);
    if (g_flags.GP && g_flags.kernel == 0) ko_calculate_C(params, ld, b, w->C);
    return ko_log_likelihood(params, ld, b, w);
}

/* DNest4::LikelihoodType ctor, DNest4/code/LikelihoodType.cpp:17-24 */
double ko_likelihood_type_value(double v)
{
    if (isinf(v) || isnan(v)) v = -1E300;
    return v;
}

/* ---- batch driver (one pthread per requested thread) -------------------- */
typedef struct {
    int B, ld;
    const double* params;
    const int* np;
    double* logL;
    volatile int* next;
    pthread_mutex_t* mtx;
} ko_job;

static void* ko_thread(void* arg)
{
    ko_job* j = arg;
    ko_work w;
    ko_work_init(&w);
    for (;;) {
        pthread_mutex_lock(j->mtx);
        int b = (*j->next)++;
        pthread_mutex_unlock(j->mtx);
        if (b >= j->B) break;
        j->logL[b] = ko_eval_lane(j->params, j->ld, b, j->np[b], &w);
    }
    ko_work_free(&w);
    return NULL;
}

/* Same contract as ref_eval_batch (oracle/ref_harness.cpp): returns seconds. */
double ko_eval_batch(int B, int ld, const double* params, const int* np,
                     double* logL, int nthreads)
{
    if (nthreads < 1) nthreads = 1;
    if (nthreads > 256) nthreads = 256;
    struct timespec t0, t1;
    clock_gettime(CLOCK_MONOTONIC, &t0);
    volatile int next = 0;
    pthread_mutex_t mtx = PTHREAD_MUTEX_INITIALIZER;
    ko_job job = {B, ld, params, np, logL, &next, &mtx};
    if (nthreads == 1) {
        ko_thread(&job);
    } else {
        pthread_t th[256];
        for (int k = 0; k < nthreads; k++) pthread_create(&th[k], NULL, ko_thread, &job);
        for (int k = 0; k < nthreads; k++) pthread_join(th[k], NULL);
    }
    clock_gettime(CLOCK_MONOTONIC, &t1);
    return (t1.tv_sec - t0.tv_sec) + 1e-9 * (t1.tv_nsec - t0.tv_nsec);
}

int ko_mu(int ld, const double* params, int b, int np, double* mu, double* mu_lo)
{
    int N = g_data.N;
    long double* m = malloc(sizeof(long double) * (N ? N : 1));
    ko_calculate_mu(params, ld, b, np, m);
    for (int i = 0; i < N; i++) {
        double hi = (double)m[i];
        mu[i] = hi;
        if (mu_lo) mu_lo[i] = (double)(m[i] - (long double)hi);
    }
    free(m);
    return N;
}

int ko_cov(int ld, const double* params, int b, int np, double* C)
{
    (void)np;
    ko_calculate_C(params, ld, b, C);
    return g_data.N;
}

/* Exposed for the Cholesky unit tests: factor in place, return Eigen's info. */
int ko_cholesky(double* A, int n) { return ko_llt_blocked(A, n); }
void ko_cholesky_solve(const double* A, int n, double* x) { ko_llt_solve(A, n, x); }

int ko_get_layout(kc_layout* out) { *out = g_layout; return 0; }
