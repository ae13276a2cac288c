// Kernel family 2, kernel = celerite: the O(N J^2) semiseparable GP likelihood.
//
// Replaces, per particle,  calculate_mu() -> calculate_C() (celerite branch) ->
// log_likelihood() (celerite branch) of the reference
// (/root/reference/src/RVmodel.cpp:218-266, 875-882) together with the solver it calls,
// celerite::solver::CholeskySolver<double>::compute / dot_solve / log_determinant
// (/root/reference/celerite/cpp/include/celerite/solver/cholesky.h:41-209, 326-400) for
// the three complex terms of kima's quasi-periodic approximation (J = 6).
//
// Two launches: (1) the residual r = y - mu, a CTA per particle (the model code of family 1,
// MA term included), written epoch-major r[n][lane] so that (2) reads it coalesced;
// (2) the semiseparable recursion, which is sequential in the epoch and 6x6 in state:
// one thread per particle, everything in registers, the particle's factorisation and the
// forward substitution of dot_solve fused in one pass (nothing of the factor is stored).
// The operation order and roundings are the reference's (no FMA contraction), so the two
// differ only by the last-bit differences of sin / cos / exp / log.
#include "kc_model.cuh"

namespace kc {

namespace {

template <int MODE>
__global__ void __launch_bounds__(CTA_THREADS)
celerite_resid_kernel(const KParams p, double* __restrict__ resid, int ldr)
{
    extern __shared__ __align__(16) double ma_buf[];
    __shared__ PlanetInv inv[MAX_PLANETS];
    __shared__ ParticleScalars sc;
    const int b = blockIdx.x, tid = threadIdx.x;
    const int N = p.N;
    if (tid < 32) particle_setup(p, b, inv, &sc, tid);
    __syncthreads();
    if (p.F.MA) {
        ma_residuals<MODE>(p, &sc, inv, ma_buf, ma_buf + N, tid, CTA_THREADS);
        for (int i = tid; i < N; i += CTA_THREADS) resid[(size_t)i * ldr + b] = ma_buf[i];
        return;
    }
    const bool multi = p.F.multi_instrument != 0;
    for (int i0 = tid; i0 < N; i0 += 2 * CTA_THREADS) {
        int i1 = i0 + CTA_THREADS;
        const bool has1 = i1 < N;
        if (!has1) i1 = i0;
        const double t0 = __ldg(p.t + i0), t1 = __ldg(p.t + i1);
        const int in0 = multi ? __ldg(p.obsi + i0) - 1 : 0;
        const int in1 = multi ? __ldg(p.obsi + i1) - 1 : 0;
        double mu0, mu1;
        model_mu2<MODE>(p, &sc, inv, i0, i1, t0, t1, in0, in1, mu0, mu1);
        resid[(size_t)i0 * ldr + b] = __ldg(p.y + i0) - mu0;  // residual(i) = y[i] - mu[i] (:845-847)
        if (has1) resid[(size_t)i1 * ldr + b] = __ldg(p.y + i1) - mu1;
    }
}

constexpr int CJ = 6;

__global__ void __launch_bounds__(32)
celerite_kernel(const KParams p, const double* __restrict__ resid, int ldr)
{
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= p.B) return;
    const int N = p.N;
    const kc_layout& L = p.L;
    const double eta1 = ldp(p, L.row_eta, b), eta2 = ldp(p, L.row_eta + 1, b), eta3 = ldp(p, L.row_eta + 2, b);
    const double extra_sigma = ldp(p, L.row_jit, b);
    // the kernel's coefficients (src/RVmodel.cpp:231-250)
    const double wbeat = kdiv(1.0, eta2), wrot = kdiv(kmul(2.0, PI), eta3), amp = kmul(eta1, eta1);
    const double c = wbeat, d = wrot, x = kdiv(c, d);
    const double a = kdiv(amp, 2.0), bq = kdiv(kmul(amp, x), 2.0), e = kdiv(amp, 8.0), f = kdiv(kmul(amp, x), 4.0);
    const double g = kmul(amp, kadd(3. / 8., 0.001));
    const double a_comp[3] = {a, e, g}, b_comp[3] = {bq, f, 0.0}, d_comp[3] = {d, kmul(2.0, d), 0.0};
    const double jitter = kmul(extra_sigma, extra_sigma);
    const double asum = kadd(kadd(a, e), g);

    double S[CJ][CJ];  // upper triangle used
#pragma unroll
    for (int j = 0; j < CJ; j++)
#pragma unroll
        for (int k = 0; k < CJ; k++) S[k][j] = 0.0;
    double W0[CJ], W1[CJ], u[CJ], fv[CJ];
    bool fail = false;

    // D_ = diag + a_real.sum() + a_comp.sum() + jitter, diag = sig^2 (cholesky.h:95)
    double t_prev = __ldg(p.t);
    double Dn = kadd(kadd(__ldg(p.sig2), asum), jitter);
    {
        const double value = kdiv(1.0, Dn);
#pragma unroll
        for (int j = 0, k = 0; j < 3; ++j, k += 2) {
            double sd, cd;
            sincos(kmul(d_comp[j], t_prev), &sd, &cd);
            W0[k] = kmul(cd, value);
            W0[k + 1] = kmul(sd, value);
        }
    }
#pragma unroll
    for (int j = 0; j < CJ; j++) fv[j] = 0.0;
    double xm1 = resid[b];
    double result = kmul(xm1, kdiv(xm1, Dn));  // cholesky.h:347
    double log_det = log(Dn);

    for (int n = 1; n < N; ++n) {
        const double t = __ldg(p.t + n);
        const double dx = kadd(t, -t_prev);
        t_prev = t;
        // c_comp is the same for the three terms: one exponential per epoch
        const double phi = exp(kmul(-c, dx));
#pragma unroll
        for (int j = 0, k = 0; j < 3; ++j, k += 2) {
            double sd, cd;
            sincos(kmul(d_comp[j], t), &sd, &cd);
            u[k] = kadd(kmul(a_comp[j], cd), kmul(b_comp[j], sd));
            u[k + 1] = kadd(kmul(a_comp[j], sd), -kmul(b_comp[j], cd));
            W1[k] = cd;
            W1[k + 1] = sd;
        }
        // S <- phi phi^T o (S + Dn W W^T)   (cholesky.h:150-156)
#pragma unroll
        for (int j = 0; j < CJ; ++j) {
            const double xj = kmul(Dn, W0[j]);
#pragma unroll
            for (int k = 0; k <= j; ++k) S[k][j] = kmul(phi, kmul(phi, kadd(S[k][j], kmul(xj, W0[k]))));
        }
        // the new pivot and row of W   (cholesky.h:158-172)
        Dn = kadd(kadd(__ldg(p.sig2 + n), asum), jitter);
#pragma unroll
        for (int j = 0; j < CJ; ++j) {
            const double uj = u[j];
            double xj = W1[j];
#pragma unroll
            for (int k = 0; k < j; ++k) {
                const double tmp = kmul(u[k], S[k][j]);
                Dn = kadd(Dn, -kmul(2.0, kmul(uj, tmp)));
                xj = kadd(xj, -tmp);
                W1[k] = kadd(W1[k], -kmul(uj, S[k][j]));
            }
            const double tmp = kmul(uj, S[j][j]);
            Dn = kadd(Dn, -kmul(uj, tmp));
            W1[j] = kadd(xj, -tmp);
        }
        if (Dn < 0) fail = true;  // linalg_exception in the reference (cholesky.h:173)
#pragma unroll
        for (int j = 0; j < CJ; ++j) W1[j] = kdiv(W1[j], Dn);
        // dot_solve: forward recursion on the residual (cholesky.h:348-357)
        double xv = resid[(size_t)n * ldr + b];
#pragma unroll
        for (int j = 0; j < CJ; ++j) {
            const double value = kmul(phi, kadd(fv[j], kmul(W0[j], xm1)));
            fv[j] = value;
            xv = kadd(xv, -kmul(u[j], value));
        }
        xm1 = xv;
        result = kadd(result, kdiv(kmul(xv, xv), Dn));
        log_det = kadd(log_det, log(Dn));
#pragma unroll
        for (int j = 0; j < CJ; ++j) W0[j] = W1[j];
    }
    // logL = -0.5 (dot_solve + log_determinant + N log(2 pi))   (src/RVmodel.cpp:877-879)
    double logL = -0.5 * kadd(kadd(result, log_det), kmul((double)N, 1.8378770664093453));
    const double pinf = __longlong_as_double(0x7ff0000000000000ll);
    if (fail || isnan(logL) || isinf(logL)) logL = pinf;  // (:938-941)
    p.logL[b] = logL;
}

// ---- conditional mean on a time grid (pykima's GP_celerite.predict, GP.py:349-355) ----
// K_* (K + sigma)^-1 r by the solver's own O(N + M) recursions: CholeskySolver::compute (cholesky.h:95-190),
// solve (:217-262: forward sweep, division by D, backward sweep) and predict (:598-690: a forward and a
// backward pass over the sorted prediction times).  One thread per particle again; what the backward sweep
// needs of the factor (D, the six W and the forward-solved x of every epoch) goes through HBM scratch laid out
// [N][8][ldr] (coalesced over particles); u and phi are recomputed from the epochs.
constexpr int CF = 8;  // D, W[6], x

__global__ void __launch_bounds__(32)
celerite_predict_kernel(const KParams p, const double* __restrict__ resid, int ldr, double* __restrict__ fac,
                        const double* __restrict__ tpred, int M, double* __restrict__ out)
{
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= p.B) return;
    const int N = p.N;
    const kc_layout& L = p.L;
    const double eta1 = ldp(p, L.row_eta, b), eta2 = ldp(p, L.row_eta + 1, b), eta3 = ldp(p, L.row_eta + 2, b);
    const double extra_sigma = ldp(p, L.row_jit, b);
    const double wbeat = kdiv(1.0, eta2), wrot = kdiv(kmul(2.0, PI), eta3), amp = kmul(eta1, eta1);
    const double c = wbeat, d = wrot, x = kdiv(c, d);
    const double a = kdiv(amp, 2.0), bq = kdiv(kmul(amp, x), 2.0), e = kdiv(amp, 8.0), f = kdiv(kmul(amp, x), 4.0);
    const double g = kmul(amp, kadd(3. / 8., 0.001));
    const double a_comp[3] = {a, e, g}, b_comp[3] = {bq, f, 0.0}, d_comp[3] = {d, kmul(2.0, d), 0.0};
    const double jitter = kmul(extra_sigma, extra_sigma);
    const double asum = kadd(kadd(a, e), g);
    auto F = [&](int n, int k) -> double& { return fac[((size_t)n * CF + k) * ldr + b]; };
    const double* tp = tpred ? tpred : p.t;
    double* row = out + (size_t)b * M;
    const double qnan = __longlong_as_double(0x7ff8000000000000ll);

    // ---- compute + forward sweep (as celerite_kernel), keeping D, W and x of every epoch ----
    double S[CJ][CJ];
#pragma unroll
    for (int j = 0; j < CJ; j++)
#pragma unroll
        for (int k = 0; k < CJ; k++) S[k][j] = 0.0;
    double W0[CJ], W1[CJ], u[CJ], fv[CJ];
    bool fail = false;
    double t_prev = __ldg(p.t);
    double Dn = kadd(kadd(__ldg(p.sig2), asum), jitter);
    {
        const double value = kdiv(1.0, Dn);
#pragma unroll
        for (int j = 0, k = 0; j < 3; ++j, k += 2) {
            double sd, cd;
            sincos(kmul(d_comp[j], t_prev), &sd, &cd);
            W0[k] = kmul(cd, value);
            W0[k + 1] = kmul(sd, value);
        }
    }
#pragma unroll
    for (int j = 0; j < CJ; j++) fv[j] = 0.0;
    double xm1 = resid[b];
    F(0, 0) = Dn;
#pragma unroll
    for (int j = 0; j < CJ; j++) F(0, 1 + j) = W0[j];
    F(0, 7) = xm1;
    for (int n = 1; n < N; ++n) {
        const double t = __ldg(p.t + n);
        const double dx = kadd(t, -t_prev);
        t_prev = t;
        const double phi = exp(kmul(-c, dx));
#pragma unroll
        for (int j = 0, k = 0; j < 3; ++j, k += 2) {
            double sd, cd;
            sincos(kmul(d_comp[j], t), &sd, &cd);
            u[k] = kadd(kmul(a_comp[j], cd), kmul(b_comp[j], sd));
            u[k + 1] = kadd(kmul(a_comp[j], sd), -kmul(b_comp[j], cd));
            W1[k] = cd;
            W1[k + 1] = sd;
        }
#pragma unroll
        for (int j = 0; j < CJ; ++j) {
            const double xj = kmul(Dn, W0[j]);
#pragma unroll
            for (int k = 0; k <= j; ++k) S[k][j] = kmul(phi, kmul(phi, kadd(S[k][j], kmul(xj, W0[k]))));
        }
        Dn = kadd(kadd(__ldg(p.sig2 + n), asum), jitter);
#pragma unroll
        for (int j = 0; j < CJ; ++j) {
            const double uj = u[j];
            double xj = W1[j];
#pragma unroll
            for (int k = 0; k < j; ++k) {
                const double tmp = kmul(u[k], S[k][j]);
                Dn = kadd(Dn, -kmul(2.0, kmul(uj, tmp)));
                xj = kadd(xj, -tmp);
                W1[k] = kadd(W1[k], -kmul(uj, S[k][j]));
            }
            const double tmp = kmul(uj, S[j][j]);
            Dn = kadd(Dn, -kmul(uj, tmp));
            W1[j] = kadd(xj, -tmp);
        }
        if (Dn < 0) fail = true;
#pragma unroll
        for (int j = 0; j < CJ; ++j) W1[j] = kdiv(W1[j], Dn);
        double xv = resid[(size_t)n * ldr + b];
#pragma unroll
        for (int j = 0; j < CJ; ++j) {
            const double value = kmul(phi, kadd(fv[j], kmul(W0[j], xm1)));
            fv[j] = value;
            xv = kadd(xv, -kmul(u[j], value));
        }
        xm1 = xv;
        F(n, 0) = Dn;
#pragma unroll
        for (int j = 0; j < CJ; ++j) {
            F(n, 1 + j) = W1[j];
            W0[j] = W1[j];
        }
        F(n, 7) = xv;
    }
    if (fail || !(Dn == Dn)) {
        for (int m = 0; m < M; m++) row[m] = qnan;
        return;
    }
    // ---- x /= D, backward sweep (cholesky.h:245-255): alpha overwrites x ----
    {
        double xn1 = kdiv(F(N - 1, 7), F(N - 1, 0));
        F(N - 1, 7) = xn1;
#pragma unroll
        for (int j = 0; j < CJ; j++) fv[j] = 0.0;
        for (int n = N - 2; n >= 0; --n) {
            const double t1 = __ldg(p.t + n + 1);
            const double phi = exp(kmul(-c, kadd(t1, -__ldg(p.t + n))));
#pragma unroll
            for (int j = 0, k = 0; j < 3; ++j, k += 2) {
                double sd, cd;
                sincos(kmul(d_comp[j], t1), &sd, &cd);
                u[k] = kadd(kmul(a_comp[j], cd), kmul(b_comp[j], sd));
                u[k + 1] = kadd(kmul(a_comp[j], sd), -kmul(b_comp[j], cd));
            }
            double xv = kdiv(F(n, 7), F(n, 0));
#pragma unroll
            for (int j = 0; j < CJ; ++j) {
                const double value = kmul(phi, kadd(fv[j], kmul(u[j], xn1)));
                fv[j] = value;
                xv = kadd(xv, -kmul(F(n, 1 + j), value));
            }
            F(n, 7) = xv;
            xn1 = xv;
        }
    }
    // ---- predict: forward pass (cholesky.h:613-647) ----
    double Q[CJ];
#pragma unroll
    for (int j = 0; j < CJ; j++) Q[j] = 0.0;
    const double t0 = __ldg(p.t), tN = __ldg(p.t + N - 1);
    int m = 0;
    while (m < M && __ldg(tp + m) <= t0) row[m++] = 0.0;  // pred.setZero(): only the backward pass reaches these
    for (int n = 0; n < N; ++n) {
        const double alphan = F(n, 7);
        const double tref = n < N - 1 ? __ldg(p.t + n + 1) : tN;
        const double tn = __ldg(p.t + n);
        const double tmp = exp(kmul(-c, kadd(tref, -tn)));
#pragma unroll
        for (int j = 0, k = 0; j < 3; ++j, k += 2) {
            double sd, cd;
            sincos(kmul(d_comp[j], tn), &sd, &cd);
            Q[k] = kmul(kadd(Q[k], kmul(alphan, cd)), tmp);
            Q[k + 1] = kmul(kadd(Q[k + 1], kmul(alphan, sd)), tmp);
        }
        while (m < M && (n == N - 1 || __ldg(tp + m) <= tref)) {
            const double xm = __ldg(tp + m);
            const double tq = exp(kmul(-c, kadd(xm, -tref)));
            double pm = 0.0;
#pragma unroll
            for (int j = 0, k = 0; j < 3; ++j, k += 2) {
                double sd, cd;
                sincos(kmul(d_comp[j], xm), &sd, &cd);
                pm = kadd(pm, kmul(kmul(kadd(kmul(a_comp[j], cd), kmul(b_comp[j], sd)), tq), Q[k]));
                pm = kadd(pm, kmul(kmul(kadd(kmul(a_comp[j], sd), -kmul(b_comp[j], cd)), tq), Q[k + 1]));
            }
            row[m] = pm;
            ++m;
        }
    }
    // ---- predict: backward pass (cholesky.h:649-687) ----
    m = M - 1;
    while (m >= 0 && __ldg(tp + m) > tN) --m;
#pragma unroll
    for (int j = 0; j < CJ; j++) Q[j] = 0.0;
    for (int n = N - 1; n >= 0; --n) {
        const double alphan = F(n, 7);
        const double tref = n > 0 ? __ldg(p.t + n - 1) : t0;
        const double tn = __ldg(p.t + n);
        const double tmp = exp(kmul(-c, kadd(tn, -tref)));
#pragma unroll
        for (int j = 0, k = 0; j < 3; ++j, k += 2) {
            double sd, cd;
            sincos(kmul(d_comp[j], tn), &sd, &cd);
            Q[k] = kmul(kadd(Q[k], kmul(alphan, kadd(kmul(a_comp[j], cd), kmul(b_comp[j], sd)))), tmp);
            Q[k + 1] = kmul(kadd(Q[k + 1], kmul(alphan, kadd(kmul(a_comp[j], sd), -kmul(b_comp[j], cd)))), tmp);
        }
        while (m >= 0 && (n == 0 || __ldg(tp + m) > tref)) {
            const double xm = __ldg(tp + m);
            const double tq = exp(kmul(-c, kadd(tref, -xm)));
            double pm = 0.0;
#pragma unroll
            for (int j = 0, k = 0; j < 3; ++j, k += 2) {
                double sd, cd;
                sincos(kmul(d_comp[j], xm), &sd, &cd);
                pm = kadd(pm, kmul(kmul(cd, tq), Q[k]));
                pm = kadd(pm, kmul(kmul(sd, tq), Q[k + 1]));
            }
            row[m] = kadd(row[m], pm);
            --m;
        }
    }
}

}  // namespace

// resid: [N][ldr] doubles, fac: [N][8][ldr] doubles of scratch; d_tpred (null = the data epochs) sorted ascending
cudaError_t launch_celerite_predict(const KParams& p, double* resid, double* fac, int ldr, int kepler_mode,
                                    const double* d_tpred, int nt, double* d_out, cudaStream_t stream, int* n_launches)
{
    if (p.B <= 0 || nt <= 0) return cudaSuccess;
    const size_t ma_smem = p.F.MA ? 2 * sizeof(double) * (size_t)p.N : 0;
    if (ma_smem > 160 * 1024) return cudaErrorInvalidValue;
    cudaError_t e;
    if (kepler_mode == 0) {
        e = cudaFuncSetAttribute(celerite_resid_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ma_smem);
        if (e != cudaSuccess) return e;
        celerite_resid_kernel<0><<<p.B, CTA_THREADS, ma_smem, stream>>>(p, resid, ldr);
    } else {
        e = cudaFuncSetAttribute(celerite_resid_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ma_smem);
        if (e != cudaSuccess) return e;
        celerite_resid_kernel<1><<<p.B, CTA_THREADS, ma_smem, stream>>>(p, resid, ldr);
    }
    celerite_predict_kernel<<<(p.B + 31) / 32, 32, 0, stream>>>(p, resid, ldr, fac, d_tpred, nt, d_out);
    *n_launches += 2;
    return cudaGetLastError();
}

cudaError_t launch_gp_celerite(const KParams& p, double* resid, int ldr, int kepler_mode,
                               cudaStream_t stream, int* n_launches)
{
    if (p.B <= 0) return cudaSuccess;
    const size_t ma_smem = p.F.MA ? 2 * sizeof(double) * (size_t)p.N : 0;
    if (ma_smem > 160 * 1024) return cudaErrorInvalidValue;
    cudaError_t e;
    if (kepler_mode == 0) {
        e = cudaFuncSetAttribute(celerite_resid_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ma_smem);
        if (e != cudaSuccess) return e;
        celerite_resid_kernel<0><<<p.B, CTA_THREADS, ma_smem, stream>>>(p, resid, ldr);
    } else {
        e = cudaFuncSetAttribute(celerite_resid_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ma_smem);
        if (e != cudaSuccess) return e;
        celerite_resid_kernel<1><<<p.B, CTA_THREADS, ma_smem, stream>>>(p, resid, ldr);
    }
    celerite_kernel<<<(p.B + 31) / 32, 32, 0, stream>>>(p, resid, ldr);
    *n_launches += 2;
    return cudaGetLastError();
}

}  // namespace kc

// ---- allocation-free variant (include/kima_cuda.h: kc_eval_celerite_noalloc) ---------------
// The same two kernels as launch_gp_celerite, with the [N][ldr] residual scratch taken from a
// file-scope device array instead of the cudaMalloc'ed block kc_configure owns, for callers
// that cannot allocate (CUDA graph capture, serving loops).  Every GPU that loads the library
// holds the array (KC_NOALLOC_MAX_DOUBLES * 8 bytes = 4 MiB) until the process exits.
#include <mutex>
#include "kc_ctx.h"

namespace {

// In HOST code this name is only a stand-in for the symbol: the device address comes from
// cudaGetSymbolAddress, once per device, below.
static __device__ __align__(256) double g_noalloc_resid[KC_NOALLOC_MAX_DOUBLES];

constexpr int NOALLOC_MAX_DEVICES = 64;
double* g_noalloc_ptr[NOALLOC_MAX_DEVICES] = {};
std::mutex g_noalloc_mutex;

}  // namespace

extern "C" int kc_eval_celerite_noalloc(kc_ctx* c, int B, const double* d_params, const int32_t* d_np,
                                        int ld, double* d_logL)
{
    int rc = kc::check_ready(c, 0);  // data uploaded, flags configured
    if (rc) return rc;
    if (!(c->F.GP && c->F.kernel == 1))
        return kc::fail(KC_ERR_ARG, "kc_eval_celerite_noalloc needs GP=true with kernel=celerite");
    if (B < 0) return kc::fail(KC_ERR_ARG, "B < 0");
    if (B > 0 && (!d_params || !d_np || !d_logL)) return kc::fail(KC_ERR_ARG, "null device pointer");
    if (B > 0 && ld < B) return kc::fail(KC_ERR_ARG, "ld < B");
    if (c->device < 0 || c->device >= NOALLOC_MAX_DEVICES)
        return kc::fail(KC_ERR_ARG, "device index beyond the scratch table");
    const size_t ldr = ((size_t)B + 31) & ~(size_t)31;
    if ((size_t)c->N * ldr > (size_t)KC_NOALLOC_MAX_DOUBLES)
        return kc::fail(KC_ERR_ARG, "kc_eval_celerite_noalloc: N * roundup(B, 32) = " +
                                        std::to_string((size_t)c->N * ldr) + " doubles of scratch exceed the static " +
                                        std::to_string((size_t)KC_NOALLOC_MAX_DOUBLES) +
                                        " (KC_NOALLOC_MAX_DOUBLES); use kc_eval_device or a smaller batch");
    cudaError_t e = cudaSetDevice(c->device);
    if (e != cudaSuccess) return kc::fail_cuda(e, "cudaSetDevice");
    double* scratch;
    {
        // set-up, once per device, with that device current
        std::lock_guard<std::mutex> lock(g_noalloc_mutex);
        if (!g_noalloc_ptr[c->device]) {
            double* ptr = nullptr;
            e = cudaGetSymbolAddress((void**)&ptr, g_noalloc_resid);
            if (e != cudaSuccess) return kc::fail_cuda(e, "cudaGetSymbolAddress(g_noalloc_resid)");
            if (!ptr) return kc::fail(KC_ERR_CUDA, "cudaGetSymbolAddress returned a null address");
            g_noalloc_ptr[c->device] = ptr;
        }
        scratch = g_noalloc_ptr[c->device];
    }
    if (B == 0) return KC_OK;  // set-up only
    kc::KParams p{};
    p.t = c->d_t; p.y = c->d_y; p.sig2 = c->d_sig2; p.obsi = c->d_obsi; p.actind = c->d_actind;
    p.N = c->N; p.n_inst = c->n_inst; p.n_ind = c->n_ind;
    p.M0_epoch = c->M0_epoch; p.t_middle = c->t_middle;
    p.F = c->F; p.L = c->L;
    p.params = d_params; p.np = d_np; p.ld = ld; p.B = B;
    p.logL = d_logL; p.counter = c->d_counter;
    // nothing in the scratch has to start at zero: the residual kernel writes every entry
    // [0,N) x [0,B) that the recursion kernel reads, whatever the last call left behind
    int n = 0;
    e = kc::launch_gp_celerite(p, scratch, (int)ldr, c->kepler_mode, c->stream, &n);
    if (e != cudaSuccess) return kc::fail_cuda(e, "kc_eval_celerite_noalloc launch");
    c->launches += n;
    return KC_OK;
}
