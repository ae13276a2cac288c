// Batched posterior-predictive curves: for every staged parameter
// vector, the deterministic model on a time grid and the GP predictive mean on it.
//
// Replaces the per-sample numpy loops of pykima's post-processing
//   model()                 /root/reference/pykima/results.py:723-797
//   GP.predict() (mean)     /root/reference/pykima/GP.py:96-112, called per sample at
//                           results.py:1313-1318 with y - v_at_t (the full residual)
// with the arithmetic of the C++ likelihood: the Keplerians are the reference's Markley
// solver (src/RVmodel.cpp:1380-1475), the covariance is calculate_C's (:186-216, sig^2 +
// jitter^2 on the diagonal) and its Cholesky factor is the one log_likelihood builds
// (:841-873) — the blocked path leaves L and z = L^-1 r in its factor storage, so
//   alpha = C^-1 r = L^-T z            (one back-substitution per lane)
//   mean(t*) = sum_i k(t* - t_i) alpha_i (the kernel generated on the fly, never stored)
// and GP.predict(return_std=True) (GP.py:113-131) as
//   std(t*) = sqrt(k(t*,t*) - |L^-1 k(t,t*)|^2)   (forward substitution against the same factor)
#include "kc_gp_common.cuh"

namespace kc {

namespace {

// ---- deterministic part ---------------------------------------------------------------------
// One CTA per parameter vector, two grid times per thread (two independent Kepler chains).
// At the data epochs (tpred == null) this is calculate_mu without the MA term, offsets and
// indicator terms included; on a grid it is model(): Keplerians (known objects too) + vsys +
// trend, "instrument offsets are only added if t is None" (results.py:727-728).
template <int MODE>
__global__ void __launch_bounds__(CTA_THREADS)
predict_model_kernel(const KParams p, const double* __restrict__ tpred, int nt, double* __restrict__ out)
{
    __shared__ PlanetInv inv[MAX_PLANETS];
    __shared__ ParticleScalars sc;
    const int b = blockIdx.x, tid = threadIdx.x;
    if (tid < 32) particle_setup(p, b, inv, &sc, tid);
    __syncthreads();
    const bool multi = p.F.multi_instrument != 0;
    double* row = out + (size_t)b * nt;
    for (int k0 = tid; k0 < nt; k0 += 2 * CTA_THREADS) {
        int k1 = k0 + CTA_THREADS;
        const bool has1 = k1 < nt;
        if (!has1) k1 = k0;
        double t0, t1, mu0, mu1;
        if (tpred) {
            t0 = __ldg(tpred + k0);
            t1 = __ldg(tpred + k1);
            mu0 = sc.vsys;
            mu1 = sc.vsys;
            if (p.F.trend) {
                const double a0 = kadd(t0, -p.t_middle), a1 = kadd(t1, -p.t_middle);
                mu0 = kadd(mu0, kadd(kadd(kmul(sc.slope, a0), kmul(sc.quadr, kmul(a0, a0))),
                                     kmul(sc.cubic, kmul(kmul(a0, a0), a0))));
                mu1 = kadd(mu1, kadd(kadd(kmul(sc.slope, a1), kmul(sc.quadr, kmul(a1, a1))),
                                     kmul(sc.cubic, kmul(kmul(a1, a1), a1))));
            }
        } else {
            t0 = __ldg(p.t + k0);
            t1 = __ldg(p.t + k1);
            const int in0 = multi ? __ldg(p.obsi + k0) - 1 : 0;
            const int in1 = multi ? __ldg(p.obsi + k1) - 1 : 0;
            mu0 = model_background(p, &sc, k0, t0, in0);
            mu1 = model_background(p, &sc, k1, t1, in1);
        }
        const int npl = sc.npl;
        for (int j = 0; j < npl; j++) {
            const PlanetInv pl = inv[j];
            mu0 += kepler_term<MODE>(pl, t0);
            mu1 += kepler_term<MODE>(pl, t1);
        }
        row[k0] = mu0;
        if (has1) row[k1] = mu1;
    }
}

// ---- alpha = L^-T z ---------------------------------------------------------------------------
// One CTA per lane of the wave.  Factor storage: W[c * ld + r] = L(r, c) for r >= c, row N
// of every column = z.  Blocks of 32 columns from the last to the first: all warps fold the
// already solved part of alpha into the block's right-hand side (columns of L are contiguous:
// coalesced dot products), then one warp back-substitutes the 32 x 32 triangle with the
// block's columns in its lanes.
constexpr int AB = 32;
constexpr int AB_LD = AB + 1;

__global__ void __launch_bounds__(CTA_THREADS)
gp_alpha_kernel(int N, size_t ld, const double* __restrict__ ws_all, size_t slot_doubles,
                double* __restrict__ alpha_all, int lane0)
{
    extern __shared__ __align__(16) double asm_[];
    const int NP = (N + AB - 1) & ~(AB - 1);
    double* a = asm_;            // [NP]: z on entry, alpha on exit
    double* blk = a + NP;        // [AB][AB_LD]: blk[cc][rr] = L(c0 + rr, c0 + cc)
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const double* W = ws_all + (size_t)blockIdx.x * slot_doubles;
    for (int i = tid; i < N; i += CTA_THREADS) a[i] = W[(size_t)i * ld + N];
    __syncthreads();
    const int nblk = (N + AB - 1) / AB;
    for (int kb = nblk - 1; kb >= 0; kb--) {
        const int c0 = AB * kb;
        const int c1 = min(N, c0 + AB);
        const int nv = c1 - c0;
        for (int cc = warp; cc < nv; cc += CTA_WARPS) {
            const double* col = W + (size_t)(c0 + cc) * ld;
            double s0 = 0.0, s1 = 0.0, s2 = 0.0, s3 = 0.0;
            int r = c1 + lane;
            for (; r + 96 < N; r += 128) {
                s0 = kfma(col[r], a[r], s0);
                s1 = kfma(col[r + 32], a[r + 32], s1);
                s2 = kfma(col[r + 64], a[r + 64], s2);
                s3 = kfma(col[r + 96], a[r + 96], s3);
            }
            for (; r < N; r += 32) s0 = kfma(col[r], a[r], s0);
            double s = (s0 + s1) + (s2 + s3);
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
            if (lane == 0) a[c0 + cc] -= s;  // only rows >= c1 of a[] are read in this phase
        }
        for (int idx = tid; idx < AB * AB; idx += CTA_THREADS) {
            const int cc = idx >> 5, rr = idx & 31;
            double v = (rr == cc) ? 1.0 : 0.0;
            if (rr < nv && cc <= rr) v = W[(size_t)(c0 + cc) * ld + c0 + rr];
            blk[cc * AB_LD + rr] = v;
        }
        __syncthreads();
        if (warp == 0) {
            double x = lane < nv ? a[c0 + lane] : 0.0;
            const double dinv = 1.0 / blk[lane * AB_LD + lane];
            for (int r = nv - 1; r >= 0; r--) {
                const double ar = __shfl_sync(0xffffffffu, x * dinv, r);
                if (lane == r) x = ar;
                else if (lane < r) x = kfma(-blk[lane * AB_LD + r], ar, x);
            }
            if (lane < nv) a[c0 + lane] = x;
        }
        __syncthreads();
    }
    double* out = alpha_all + (size_t)(lane0 + blockIdx.x) * N;
    for (int i = tid; i < N; i += CTA_THREADS) out[i] = a[i];
}

// ---- mean(t*) = K(t*, t) alpha ----------------------------------------------------------------
// grid (time tiles, lanes of the wave); alpha and the epochs in shared memory, two grid
// times per thread through qp_kernel2 (the same kernel function the factorisation used).
__global__ void __launch_bounds__(CTA_THREADS)
gp_mean_kernel(const KParams p, const GpLane* __restrict__ lanes, const double* __restrict__ alpha_all,
               const double* __restrict__ tpred, int nt, double* __restrict__ out, int lane0)
{
    extern __shared__ __align__(16) double msm[];
    const int N = p.N;
    double* a = msm;
    double* ts = msm + N;
    const int tid = threadIdx.x, slot = blockIdx.y;
    const double* al = alpha_all + (size_t)(lane0 + slot) * N;
    for (int i = tid; i < N; i += CTA_THREADS) {
        a[i] = al[i];
        ts[i] = __ldg(p.t + i);
    }
    __syncthreads();
    const GpConsts g = lanes[slot].c;
    const bool failed = lanes[slot].fail != 0;
    const int k0 = blockIdx.x * 2 * CTA_THREADS + tid;
    if (k0 >= nt) return;
    int k1 = k0 + CTA_THREADS;
    const bool has1 = k1 < nt;
    if (!has1) k1 = k0;
    const double* tp = tpred ? tpred : p.t;
    const double t0 = __ldg(tp + k0), t1 = __ldg(tp + k1);
    double m0 = 0.0, m1 = 0.0;
    for (int i = 0; i < N; i++) {
        double c0, c1;
        const double ti = ts[i];
        qp_kernel2(g, kadd(ti, -t0), kadd(ti, -t1), c0, c1);
        m0 = kfma(c0, a[i], m0);
        m1 = kfma(c1, a[i], m1);
    }
    if (failed) {  // C was not positive definite: no factor, no prediction
        m0 = m1 = __longlong_as_double(0x7ff8000000000000ll);
    }
    double* row = out + (size_t)(lane0 + slot) * nt;
    row[k0] = m0;
    if (has1) row[k1] = m1;
}

// ---- std(t*) = sqrt(k(t*,t*) - |L^-1 k(t,t*)|^2) -----------------------------------------------
// GP.predict(return_std=True), /root/reference/pykima/GP.py:113-131 (which forms C^-1 explicitly;
// here the forward substitution against the factor the likelihood left in HBM).
// grid (tiles of TV grid times, lanes of the wave).  The tile's right-hand sides live in shared
// memory, xs[tv][NP]; blocked forward substitution over 32-column blocks of L: the 32 x 32
// triangle by one warp per right-hand side (rows in lanes, shuffles), then every thread folds the
// block's solution into the rows below it (columns of L are contiguous: coalesced, read once per
// tile; the 32 x TV solved values are broadcast from shared memory as 16-byte pairs).
template <int TV>
__global__ void __launch_bounds__(CTA_THREADS)
gp_std_kernel(const KParams p, size_t ld, const double* __restrict__ ws_all, size_t slot_doubles,
              const GpLane* __restrict__ lanes, const double* __restrict__ tpred, int nt,
              double* __restrict__ out, int lane0)
{
    extern __shared__ __align__(16) double vsm[];
    const int N = p.N;
    const int NP = (N + AB - 1) & ~(AB - 1);
    double* xs = vsm;                       // [TV][NP]
    double* blk = xs + (size_t)TV * NP;     // [AB][AB_LD]: blk[cc][rr] = L(c0 + rr, c0 + cc)
    double* vb = blk + AB * AB_LD;          // [AB][TV]: the block's solution, tv contiguous
    double* dinv = vb + AB * TV;            // [AB]
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int slot = blockIdx.y;
    const double* W = ws_all + (size_t)slot * slot_doubles;
    const GpConsts g = lanes[slot].c;
    const bool failed = lanes[slot].fail != 0;
    const int k_first = blockIdx.x * TV;
    const double* tp = tpred ? tpred : p.t;
    // right-hand sides: k(t_r, t*_tv), the kernel function the factorisation used
    for (int idx = tid; idx < (TV / 2) * NP; idx += CTA_THREADS) {
        const int r = idx % NP, tv = 2 * (idx / NP);
        const double tr = __ldg(p.t + min(r, N - 1));
        const double t0 = __ldg(tp + min(k_first + tv, nt - 1)), t1 = __ldg(tp + min(k_first + tv + 1, nt - 1));
        double c0, c1;
        qp_kernel2(g, kadd(tr, -t0), kadd(tr, -t1), c0, c1);
        xs[(size_t)tv * NP + r] = r < N ? c0 : 0.0;
        xs[(size_t)(tv + 1) * NP + r] = r < N ? c1 : 0.0;
    }
    double q[(TV + CTA_WARPS - 1) / CTA_WARPS];
#pragma unroll
    for (int i = 0; i < (TV + CTA_WARPS - 1) / CTA_WARPS; i++) q[i] = 0.0;
    __syncthreads();
    const int nblk = (N + AB - 1) / AB;
    for (int kb = 0; kb < nblk; kb++) {
        const int c0 = AB * kb;
        const int c1 = min(N, c0 + AB);
        const int nv = c1 - c0;
        for (int idx = tid; idx < AB * AB; idx += CTA_THREADS) {
            const int cc = idx >> 5, rr = idx & 31;
            double v = (rr == cc) ? 1.0 : 0.0;
            if (rr < nv && cc <= rr) v = W[(size_t)(c0 + cc) * ld + c0 + rr];
            blk[cc * AB_LD + rr] = v;
            if (rr == cc) dinv[cc] = 1.0 / v;
        }
        __syncthreads();
        // the triangle: warp w takes right-hand sides w, w + 8, ...
#pragma unroll
        for (int i = 0; i < (TV + CTA_WARPS - 1) / CTA_WARPS; i++) {
            const int tv = warp + i * CTA_WARPS;
            if (tv < TV) {  // warp-uniform
                double x = lane < nv ? xs[(size_t)tv * NP + c0 + lane] : 0.0;
                const double di = dinv[lane];
                for (int r = 0; r < nv; r++) {
                    const double ar = __shfl_sync(0xffffffffu, x * di, r);
                    if (lane == r) x = ar;
                    else if (lane > r) x = kfma(-blk[r * AB_LD + lane], ar, x);
                }
                if (lane >= nv) x = 0.0;
                vb[lane * TV + tv] = x;
                double s = x * x;
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
                q[i] += s;
            }
        }
        __syncthreads();
        // rows below the block: x(r, :) -= sum_cc L(r, c0 + cc) v(c0 + cc, :)
        for (int r = c1 + tid; r < N; r += CTA_THREADS) {
            double acc[TV];
#pragma unroll
            for (int tv = 0; tv < TV; tv++) acc[tv] = 0.0;
            const double* Lr = W + (size_t)c0 * ld + r;
#pragma unroll 4
            for (int cc = 0; cc < AB; cc++) {
                const double lv = cc < nv ? Lr[(size_t)cc * ld] : 0.0;
                const double2* v2 = reinterpret_cast<const double2*>(vb + cc * TV);
#pragma unroll
                for (int h = 0; h < TV / 2; h++) {
                    const double2 vv = v2[h];
                    acc[2 * h] = kfma(lv, vv.x, acc[2 * h]);
                    acc[2 * h + 1] = kfma(lv, vv.y, acc[2 * h + 1]);
                }
            }
#pragma unroll
            for (int tv = 0; tv < TV; tv++) xs[(size_t)tv * NP + r] -= acc[tv];
        }
        __syncthreads();
    }
    if (lane == 0) {
#pragma unroll
        for (int i = 0; i < (TV + CTA_WARPS - 1) / CTA_WARPS; i++) {
            const int tv = warp + i * CTA_WARPS;
            const int k = k_first + tv;
            if (tv < TV && k < nt) {
                // var < 0 "because of numerical issues": set to 0 (GP.py:125-129); k(t*,t*) = eta1^2
                double var = kadd(g.eta1sq, -q[i]);
                double sd = var > 0.0 ? sqrt(var) : 0.0;
                if (failed || isnan(var)) sd = __longlong_as_double(0x7ff8000000000000ll);
                out[(size_t)(lane0 + slot) * nt + k] = sd;
            }
        }
    }
}

template <int TV>
static cudaError_t launch_std(const KParams& p, const double* ws, size_t slot_doubles, const GpLane* lanes,
                              int lane0, int n, const GpPredict& pr, cudaStream_t s)
{
    const int N = p.N;
    const size_t NP = ((size_t)N + AB - 1) / AB * AB;
    const size_t smem = sizeof(double) * ((size_t)TV * NP + AB * AB_LD + AB * TV + AB);
    cudaError_t e = cudaFuncSetAttribute(gp_std_kernel<TV>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    gp_std_kernel<TV><<<dim3((pr.nt + TV - 1) / TV, n), CTA_THREADS, smem, s>>>(p, gpb_ld(N), ws, slot_doubles, lanes,
                                                                                pr.t_pred, pr.nt, pr.std_out, lane0);
    return cudaGetLastError();
}

}  // namespace

cudaError_t launch_predict_model(const KParams& p, int kepler_mode, const double* d_tpred, int nt,
                                 double* d_out, cudaStream_t s, int* n_launches)
{
    if (p.B <= 0 || nt <= 0) return cudaSuccess;
    if (kepler_mode == 0) predict_model_kernel<0><<<p.B, CTA_THREADS, 0, s>>>(p, d_tpred, nt, d_out);
    else predict_model_kernel<1><<<p.B, CTA_THREADS, 0, s>>>(p, d_tpred, nt, d_out);
    *n_launches += 1;
    return cudaGetLastError();
}

cudaError_t launch_gp_predict_wave(const KParams& p, const double* ws, size_t slot_doubles,
                                   const GpLane* lanes, int lane0, int n, const GpPredict& pr,
                                   cudaStream_t s, int* n_launches)
{
    const int N = p.N;
    const size_t asmem = sizeof(double) * (((size_t)N + AB - 1) / AB * AB + AB * AB_LD);
    const size_t msmem = sizeof(double) * 2 * (size_t)N;
    if (asmem > 200 * 1024 || msmem > 200 * 1024) return cudaErrorInvalidValue;
    cudaError_t e;
    if (asmem > 48 * 1024) {
        e = cudaFuncSetAttribute(gp_alpha_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)asmem);
        if (e != cudaSuccess) return e;
    }
    if (msmem > 48 * 1024) {
        e = cudaFuncSetAttribute(gp_mean_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)msmem);
        if (e != cudaSuccess) return e;
    }
    if (pr.gp_out) {
        gp_alpha_kernel<<<n, CTA_THREADS, asmem, s>>>(N, gpb_ld(N), ws, slot_doubles, pr.alpha, lane0);
        const int tiles = (pr.nt + 2 * CTA_THREADS - 1) / (2 * CTA_THREADS);
        gp_mean_kernel<<<dim3(tiles, n), CTA_THREADS, msmem, s>>>(p, lanes, pr.alpha, pr.t_pred, pr.nt, pr.gp_out, lane0);
        *n_launches += 2;
    }
    if (pr.std_out) {
        // right-hand sides per tile by what fits beside the tile in shared memory
        const size_t budget = 200 * 1024 / sizeof(double) - (AB * AB_LD + AB * 16 + AB);
        const size_t NP = ((size_t)N + AB - 1) / AB * AB;
        if (16 * NP <= budget) e = launch_std<16>(p, ws, slot_doubles, lanes, lane0, n, pr, s);
        else if (8 * NP <= budget) e = launch_std<8>(p, ws, slot_doubles, lanes, lane0, n, pr, s);
        else if (4 * NP <= budget) e = launch_std<4>(p, ws, slot_doubles, lanes, lane0, n, pr, s);
        else if (2 * NP <= budget) e = launch_std<2>(p, ws, slot_doubles, lanes, lane0, n, pr, s);
        else return cudaErrorInvalidValue;
        if (e != cudaSuccess) return e;
        *n_launches += 1;
    }
    return cudaGetLastError();
}

}  // namespace kc
