// Kernel family 1: fused Keplerian sum + white-noise likelihood.
//
// Replaces, per particle,  calculate_mu()  ->  log_likelihood()  (Gaussian and
// Student-t branches) of the reference
// (/root/reference/src/RVmodel.cpp:286-375, 889-931, 938-941).
// mu[] is never written to memory: each lane owns epochs i = lane, lane+32, ...
// sums the Keplerians of its epochs in registers, turns them into residual
// terms immediately and the warp reduces with shuffles.
//
// Tiling: WPP warps per particle.  WPP = 1 (a warp per particle, warps of a
// CTA independent, work fetched from an atomic counter) for large batches;
// WPP = 8 (a CTA per particle) when the batch is too small to fill 148 SMs
// with one warp per particle.
#include "kc_model.cuh"

namespace kc {

constexpr double HALF_LOG_2PI = 0.91893853320467274178;  // 0.5*log(2*pi) (:20)

struct __align__(16) ParticleSlot {
    PlanetInv inv[MAX_PLANETS];
    ParticleScalars sc;
};

__device__ __forceinline__ double warp_sum(double v)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// One epoch's log-likelihood term from its residual and variance.
template <int LIK>
__device__ __forceinline__ double loglik_term(double r, double var, double hnu1, double inv_nu)
{
    if (LIK == 0) return -HALF_LOG_2PI - 0.5 * log(var) - 0.5 * div_fast(r * r, var);
    return -0.5 * log(var) - hnu1 * log1p(div_fast(r * r, var) * inv_nu);
}

// sum_i log(var_i) of a lane's epochs as ONE logarithm: the mantissas are multiplied up (each in [1, 2): no
// overflow below 1023 factors), the exponents added as integers.  `log` is the dearest call of an epoch
// without planets (~60 FP64 issue slots of ~250: the 51Peg and BL2009 shapes spend a fifth of their
// instructions there), and log(prod m_i) + ln 2 sum e_i carries less rounding error than the sum of the
// rounded logarithms the reference forms (:912-929).  Zero, subnormal, negative, infinite or NaN variances
// take the plain way, so the non-finite cases end up where they did.
struct LogProd {
    double m = 1.0, plain = 0.0;
    int e = 0, n = 0;
    __device__ __forceinline__ void mul(double v)
    {
        const int hi = __double2hiint(v);
        const int ex = (hi >> 20) & 0x7ff;
        if (hi < 0 || ex == 0 || ex == 0x7ff) {
            plain += log(v);
            return;
        }
        e += ex - 1023;
        m *= __hiloint2double((hi & 0x000fffff) | 0x3ff00000, __double2loint(v));
        if (++n == 1000) {  // never in practice (a lane has N / 32 epochs): fold the product
            plain += log(m);
            m = 1.0;
            n = 0;
        }
    }
    __device__ __forceinline__ double sum() const
    {
        // ln 2 split so that e * hi is exact (32 significant bits x |e| < 2^21)
        return plain + (log(m) + kfma((double)e, 1.90821492927058770002e-10, (double)e * 6.93147180369123816490e-01));
    }
};

// One epoch's term without its -log(2 pi)/2 - log(var)/2 (lane_loglik adds those once per lane).
template <int LIK>
__device__ __forceinline__ double loglik_quad(double r, double var, double hnu1, double inv_nu)
{
    if (LIK == 0) return -0.5 * div_fast(r * r, var);
    return -hnu1 * log1p(div_fast(r * r, var) * inv_nu);
}

// Sum of the per-epoch log-likelihood terms of this lane's epochs.
// LIK: 0 Gaussian (:912-929), 1 Student-t (:891-910).
template <int MODE, int LIK>
__device__ __forceinline__ double lane_loglik(const KParams& p, const ParticleSlot* slot,
                                              int first, int stride)
{
    const ParticleScalars* sc = &slot->sc;
    const int N = p.N;
    const bool multi = p.F.multi_instrument != 0;
    double acc = 0.0;
    LogProd lp;
    int count = 0;
    const double nu = sc->nu;
    const double hnu1 = 0.5 * (nu + 1.0);
    const double inv_nu = LIK == 1 ? 1.0 / nu : 0.0;
    for (int i0 = first; i0 < N; i0 += 2 * stride) {
        int i1 = i0 + stride;
        const bool has1 = i1 < N;
        if (!has1) i1 = i0;
        double t0 = __ldg(p.t + i0), t1 = __ldg(p.t + i1);
        int in0 = multi ? __ldg(p.obsi + i0) - 1 : 0;
        int in1 = multi ? __ldg(p.obsi + i1) - 1 : 0;
        double mu0, mu1;
        model_mu2<MODE>(p, sc, slot->inv, i0, i1, t0, t1, in0, in1, mu0, mu1);
        double r0 = __ldg(p.y + i0) - mu0, r1 = __ldg(p.y + i1) - mu1;
        double var0 = kadd(__ldg(p.sig2 + i0), sc->jit2[in0]);
        double var1 = kadd(__ldg(p.sig2 + i1), sc->jit2[in1]);
        if (LIK == 0) {
            acc += loglik_quad<LIK>(r0, var0, hnu1, inv_nu);
            lp.mul(var0);
            if (has1) {
                acc += loglik_quad<LIK>(r1, var1, hnu1, inv_nu);
                lp.mul(var1);
            }
            count += has1 ? 2 : 1;
        } else {
            // Student-t: the term keeps its own logarithm (a second one, log1p, dominates; the product form
            // costs this variant 6 registers and an eighth of its occupancy for nothing)
            acc += loglik_term<LIK>(r0, var0, hnu1, inv_nu);
            if (has1) acc += loglik_term<LIK>(r1, var1, hnu1, inv_nu);
        }
    }
    if (LIK == 0) acc = kfma(-HALF_LOG_2PI, (double)count, kfma(-0.5, lp.sum(), acc));
    return acc;
}

// Per-particle constant of the Student-t likelihood (:905-906), added N times
// in the reference.
__device__ __forceinline__ double studentt_const(double nu, int N)
{
    return N * (lgamma(0.5 * (nu + 1.0)) - lgamma(0.5 * nu) - 0.5 * log(PI * nu));
}

__device__ __forceinline__ double finish(double logL)
{
    // NaN or +-inf -> +inf (:938-941)
    return (isnan(logL) || isinf(logL)) ? __longlong_as_double(0x7ff0000000000000ll) : logL;
}

// ---- WPP = 1: a warp per particle ------------------------------------------------
template <int MODE, int LIK>
__global__ void __launch_bounds__(CTA_THREADS)
keplerian_warp_kernel(const KParams p)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    ParticleSlot* slots = reinterpret_cast<ParticleSlot*>(smem_raw);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    ParticleSlot* slot = slots + warp;
    for (;;) {
        int b = 0;
        if (lane == 0) b = atomicAdd(p.counter, 1);
        b = __shfl_sync(0xffffffffu, b, 0);
        if (b >= p.B) break;
        particle_setup(p, b, slot->inv, &slot->sc, lane);
        double acc = lane_loglik<MODE, LIK>(p, slot, lane, 32);
        acc = warp_sum(acc);
        if (lane == 0) {
            if (LIK == 1) acc += studentt_const(slot->sc.nu, p.N);
            p.logL[b] = finish(acc);
        }
        __syncwarp();
    }
}

// ---- WPP = 8: a CTA per particle ---------------------------------------------------
template <int MODE, int LIK>
__global__ void __launch_bounds__(CTA_THREADS)
keplerian_cta_kernel(const KParams p)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    ParticleSlot* slot = reinterpret_cast<ParticleSlot*>(smem_raw);
    __shared__ double partial[CTA_WARPS];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int b = blockIdx.x; b < p.B; b += gridDim.x) {
        if (warp == 0) particle_setup(p, b, slot->inv, &slot->sc, lane);
        __syncthreads();
        double acc = lane_loglik<MODE, LIK>(p, slot, threadIdx.x, CTA_THREADS);
        acc = warp_sum(acc);
        if (lane == 0) partial[warp] = acc;
        __syncthreads();
        if (threadIdx.x == 0) {
            double s = 0.0;
#pragma unroll
            for (int w = 0; w < CTA_WARPS; w++) s += partial[w];
            if (LIK == 1) s += studentt_const(slot->sc.nu, p.N);
            p.logL[b] = finish(s);
        }
        __syncthreads();
    }
}

// ---- MA = true: a CTA per particle, the residual recurrence in shared memory ------------------
// (:378-386).  The model without the MA term goes to shared memory as b = y - mu0 together
// with the recurrence coefficients, one warp runs the scan (kc_model.cuh), then every thread
// turns its epochs' residuals into likelihood terms.
template <int MODE, int LIK>
__global__ void __launch_bounds__(CTA_THREADS)
keplerian_ma_kernel(const KParams p)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    ParticleSlot* slot = reinterpret_cast<ParticleSlot*>(smem_raw);
    double* rbuf = reinterpret_cast<double*>(smem_raw + ((sizeof(ParticleSlot) + 15) & ~(size_t)15));
    double* cbuf = rbuf + p.N;
    __shared__ double partial[CTA_WARPS];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const bool multi = p.F.multi_instrument != 0;
    for (int b = blockIdx.x; b < p.B; b += gridDim.x) {
        if (warp == 0) particle_setup(p, b, slot->inv, &slot->sc, lane);
        __syncthreads();
        ma_residuals<MODE>(p, &slot->sc, slot->inv, rbuf, cbuf, threadIdx.x, CTA_THREADS);
        const double nu = slot->sc.nu;
        const double hnu1 = 0.5 * (nu + 1.0), inv_nu = LIK == 1 ? 1.0 / nu : 0.0;
        double acc = 0.0;
        for (int i = threadIdx.x; i < p.N; i += CTA_THREADS) {
            const int in = multi ? __ldg(p.obsi + i) - 1 : 0;
            acc += loglik_term<LIK>(rbuf[i], kadd(__ldg(p.sig2 + i), slot->sc.jit2[in]), hnu1, inv_nu);
        }
        acc = warp_sum(acc);
        if (lane == 0) partial[warp] = acc;
        __syncthreads();
        if (threadIdx.x == 0) {
            double s = 0.0;
#pragma unroll
            for (int w = 0; w < CTA_WARPS; w++) s += partial[w];
            if (LIK == 1) s += studentt_const(slot->sc.nu, p.N);
            p.logL[b] = finish(s);
        }
        __syncthreads();
    }
}

static size_t ma_smem_bytes(int N)
{
    return ((sizeof(ParticleSlot) + 15) & ~(size_t)15) + 2 * sizeof(double) * (size_t)N;
}

template <int MODE, int LIK>
static cudaError_t launch_variant(const KParams& p, int sm_count, cudaStream_t stream,
                                  int* n_launches)
{
    // One warp per particle once that alone gives every SM >= 16 warps.
    const bool per_warp = p.B >= sm_count * 16;
    cudaError_t e;
    if (p.F.MA) {
        const size_t smem = ma_smem_bytes(p.N);
        if (smem > 227 * 1024) return cudaErrorInvalidValue;  // N > ~14000 epochs
        auto k = keplerian_ma_kernel<MODE, LIK>;
        e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
        int per_sm = 0;
        e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k, CTA_THREADS, smem);
        if (e != cudaSuccess) return e;
        if (per_sm < 1) per_sm = 1;
        int grid = sm_count * per_sm;
        if (grid > p.B) grid = p.B;
        k<<<grid, CTA_THREADS, smem, stream>>>(p);
    } else if (per_warp) {
        size_t smem = sizeof(ParticleSlot) * CTA_WARPS;
        auto k = keplerian_warp_kernel<MODE, LIK>;
        e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
        int per_sm = 0;
        e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k, CTA_THREADS, smem);
        if (e != cudaSuccess) return e;
        if (per_sm < 1) per_sm = 1;
        int grid = sm_count * per_sm;
        int need = (p.B + CTA_WARPS - 1) / CTA_WARPS;
        if (grid > need) grid = need;
        e = cudaMemsetAsync(p.counter, 0, sizeof(int), stream);
        if (e != cudaSuccess) return e;
        k<<<grid, CTA_THREADS, smem, stream>>>(p);
    } else {
        size_t smem = sizeof(ParticleSlot);
        auto k = keplerian_cta_kernel<MODE, LIK>;
        int per_sm = 0;
        e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k, CTA_THREADS, smem);
        if (e != cudaSuccess) return e;
        if (per_sm < 1) per_sm = 1;
        int grid = sm_count * per_sm;
        if (grid > p.B) grid = p.B;
        k<<<grid, CTA_THREADS, smem, stream>>>(p);
    }
    *n_launches += 1;
    return cudaGetLastError();
}

cudaError_t launch_keplerian(const KParams& p, int kepler_mode, int sm_count, cudaStream_t stream,
                             int* n_launches)
{
    if (p.B <= 0) return cudaSuccess;
    const int lik = p.F.studentt ? 1 : 0;
    if (kepler_mode == 0) {
        return lik ? launch_variant<0, 1>(p, sm_count, stream, n_launches)
                   : launch_variant<0, 0>(p, sm_count, stream, n_launches);
    }
    return lik ? launch_variant<1, 1>(p, sm_count, stream, n_launches)
               : launch_variant<1, 0>(p, sm_count, stream, n_launches);
}

// ---- test hooks ---------------------------------------------------------------------
__global__ void kepler_E_kernel(int n, const double* M, const double* e, double* E)
{
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) E[i] = kepler_E(M[i], e[i]);
}
__global__ void true_anomaly_kernel(int n, const double* t, const double* P, const double* e,
                                    const double* Tp, double* f)
{
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) f[i] = true_anomaly(t[i], P[i], e[i], Tp[i]);
}
template <int MODE>
__global__ void __launch_bounds__(CTA_THREADS) model_mu_kernel(const KParams p, int b, double* mu)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    __shared__ ParticleSlot slot;
    if (threadIdx.x < 32) particle_setup(p, b, slot.inv, &slot.sc, threadIdx.x);
    __syncthreads();
    if (p.F.MA) {
        double* rbuf = reinterpret_cast<double*>(smem_raw);
        ma_residuals<MODE>(p, &slot.sc, slot.inv, rbuf, rbuf + p.N, threadIdx.x, CTA_THREADS);
        for (int i = threadIdx.x; i < p.N; i += blockDim.x) mu[i] = p.y[i] - rbuf[i];
        return;
    }
    const bool multi = p.F.multi_instrument != 0;
    for (int i = threadIdx.x; i < p.N; i += blockDim.x) {
        double t0 = p.t[i];
        int in0 = multi ? p.obsi[i] - 1 : 0;
        double m0, m1;
        model_mu2<MODE>(p, &slot.sc, slot.inv, i, i, t0, t0, in0, in0, m0, m1);
        mu[i] = m0;
    }
}

cudaError_t launch_kepler_E(int n, const double* M, const double* e, double* E, cudaStream_t s)
{
    if (n <= 0) return cudaSuccess;
    kepler_E_kernel<<<(n + 255) / 256, 256, 0, s>>>(n, M, e, E);
    return cudaGetLastError();
}
cudaError_t launch_true_anomaly(int n, const double* t, const double* P, const double* e,
                                const double* Tp, double* f, cudaStream_t s)
{
    if (n <= 0) return cudaSuccess;
    true_anomaly_kernel<<<(n + 255) / 256, 256, 0, s>>>(n, t, P, e, Tp, f);
    return cudaGetLastError();
}
cudaError_t launch_model_mu(const KParams& p, int kepler_mode, int lane, double* mu, cudaStream_t s)
{
    const size_t smem = p.F.MA ? 2 * sizeof(double) * (size_t)p.N : 0;
    if (smem > 160 * 1024) return cudaErrorInvalidValue;
    cudaError_t e;
    if (kepler_mode == 0) {
        e = cudaFuncSetAttribute(model_mu_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
        model_mu_kernel<0><<<1, CTA_THREADS, smem, s>>>(p, lane, mu);
    } else {
        e = cudaFuncSetAttribute(model_mu_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
        model_mu_kernel<1><<<1, CTA_THREADS, smem, s>>>(p, lane, mu);
    }
    return cudaGetLastError();
}

}  // namespace kc
