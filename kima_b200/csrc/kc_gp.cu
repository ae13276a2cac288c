// Kernel family 2: quasi-periodic GP likelihood.
//
// Replaces, per particle,  calculate_mu() -> calculate_C() -> log_likelihood()
// (GP / standard-kernel branch) of the reference
// (/root/reference/src/RVmodel.cpp:170-216, 286-375, 841-873; Eigen LLT,
// eigen/Eigen/src/Cholesky/LLT.h:313-368).
//
// One CTA factors one particle's covariance with a LEFT-looking blocked
// Cholesky:
//   * C is never stored: block column J is generated from (t_i - t_j) at its
//     first and only touch, S = C[:,J] - L[:,0:J] L[J,0:J]^T, so the only HBM
//     traffic is one write of L and the re-reads of earlier block columns.
//   * The residual r = y - mu rides along as row N of the matrix.  Its factored
//     image is z = L^-1 r, so r^T C^-1 r = |z|^2 and no triangular solves are
//     needed afterwards:  logL = -N/2 log 2pi - sum log L_ii - |z|^2 / 2.
//   * The rank-k update runs on the FP64 tensor cores (mma.sync m8n8k4, SASS
//     DMMA.8x8x4): measured 37.1 TFLOP/s on B200, the same ceiling as the DFMA
//     pipe but one issue slot per 512 flops instead of per 64.
#include "kc_model.cuh"
#include "kc_gp_common.cuh"

namespace kc {

namespace gp {
constexpr int NB = 32;    // block-column width
constexpr int R = 128;    // rows per tile
constexpr int KC = 16;    // k-chunk staged in shared memory
constexpr int A_LD = R + 4;    // 132 = 4 mod 16: conflict-free fragment reads
constexpr int B_LD = NB + 4;   // 36  = 4 mod 16
constexpr int S_LD = NB + 1;
}  // namespace gp

size_t gp_slot_doubles(int N)
{
    size_t ld = ((size_t)N + 1 + 3) & ~(size_t)3;
    return ld * (size_t)N;
}

int gp_default_slots(int N, int sm_count)
{
    size_t bytes = gp_slot_doubles(N) * sizeof(double);
    size_t cap = (size_t)40 << 30;  // keep the workspace under 40 GB of the 180
    int slots = 2 * sm_count;
    if ((size_t)slots * bytes > cap) slots = (int)(cap / bytes);
    return slots < 1 ? 1 : slots;
}

struct GpSmem {
    double As[gp::KC][gp::A_LD];
    double Bs[gp::KC][gp::B_LD];
    double Ss[gp::R][gp::S_LD];
    double Ljj[gp::NB][gp::S_LD];
    double dinv[gp::NB];
    PlanetInv inv[MAX_PLANETS];
    ParticleScalars sc;
    GpConsts g;
    double red[CTA_WARPS];
    double logdet, quad;
    int fail;
    int b;
};

template <int MODE>
__global__ void __launch_bounds__(CTA_THREADS)
gp_loglik_kernel(const KParams p, double* __restrict__ ws_all, size_t slot_doubles)
{
    using namespace gp;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    GpSmem& sm = *reinterpret_cast<GpSmem*>(smem_raw);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int N = p.N, N1 = N + 1;
    const int ld = (N1 + 3) & ~3;
    double* W = ws_all + (size_t)blockIdx.x * slot_doubles;
    const bool multi = p.F.multi_instrument != 0;

    for (;;) {
        if (tid == 0) sm.b = atomicAdd(p.counter, 1);
        __syncthreads();
        const int b = sm.b;
        if (b >= p.B) break;
        if (warp == 0) {
            particle_setup(p, b, sm.inv, &sm.sc, lane);
            if (lane == 0) {
                const ParticleScalars& sc = sm.sc;
                sm.g.eta1sq = kmul(sc.eta1, sc.eta1);
                sm.g.eta2 = sc.eta2; sm.g.r_eta2 = kdiv(1.0, sc.eta2);
                sm.g.eta3 = sc.eta3; sm.g.r_eta3 = kdiv(1.0, sc.eta3);
                sm.g.eta4 = sc.eta4; sm.g.r_eta4 = kdiv(1.0, sc.eta4);
                sm.logdet = 0.0; sm.quad = 0.0; sm.fail = 0;
            }
        }
        __syncthreads();

        // residual row: W(N, i) = y_i - mu_i   (:845-847)
        for (int i0 = tid; i0 < N; i0 += 2 * CTA_THREADS) {
            int i1 = i0 + CTA_THREADS;
            const bool has1 = i1 < N;
            if (!has1) i1 = i0;
            double t0 = __ldg(p.t + i0), t1 = __ldg(p.t + i1);
            int in0 = multi ? __ldg(p.obsi + i0) - 1 : 0;
            int in1 = multi ? __ldg(p.obsi + i1) - 1 : 0;
            double mu0, mu1;
            model_mu2<MODE>(p, &sm.sc, sm.inv, i0, i1, t0, t1, in0, in1, mu0, mu1);
            W[(size_t)i0 * ld + N] = __ldg(p.y + i0) - mu0;
            if (has1) W[(size_t)i1 * ld + N] = __ldg(p.y + i1) - mu1;
        }
        __syncthreads();

        const GpConsts g = sm.g;
        const int nblk = (N + NB - 1) / NB;
        for (int J = 0; J < nblk; J++) {
            const int j0 = J * NB;
            const int nbc = min(NB, N - j0);
            const int rows = N1 - j0;
            const int ntile = (rows + R - 1) / R;
            for (int T = 0; T < ntile; T++) {
                const int r0 = j0 + T * R;
                // ---- P = L[r0:r0+R, 0:j0] * L[j0:j0+NB, 0:j0]^T on the tensor cores ----
                // warp w owns rows [16w, 16w+16) x all NB columns: 2 x 4 fragments
                double acc[2][4][2];
#pragma unroll
                for (int a = 0; a < 2; a++)
#pragma unroll
                    for (int c = 0; c < 4; c++) acc[a][c][0] = acc[a][c][1] = 0.0;
                for (int k0 = 0; k0 < j0; k0 += KC) {
                    // stage A (KC x R) and B (KC x NB); columns of L are contiguous in rows
                    for (int e = tid; e < KC * R; e += CTA_THREADS) {
                        int kk = e / R, r = e % R;
                        int row = r0 + r;
                        sm.As[kk][r] = row < N1 ? W[(size_t)(k0 + kk) * ld + row] : 0.0;
                    }
                    for (int e = tid; e < KC * NB; e += CTA_THREADS) {
                        int kk = e / NB, c = e % NB;
                        int row = j0 + c;
                        sm.Bs[kk][c] = row < N ? W[(size_t)(k0 + kk) * ld + row] : 0.0;
                    }
                    __syncthreads();
#pragma unroll
                    for (int kk = 0; kk < KC; kk += 4) {
                        double af[2], bf[4];
#pragma unroll
                        for (int a = 0; a < 2; a++)
                            af[a] = sm.As[kk + (lane & 3)][warp * 16 + a * 8 + (lane >> 2)];
#pragma unroll
                        for (int c = 0; c < 4; c++) bf[c] = sm.Bs[kk + (lane & 3)][c * 8 + (lane >> 2)];
#pragma unroll
                        for (int a = 0; a < 2; a++)
#pragma unroll
                            for (int c = 0; c < 4; c++) dmma(acc[a][c][0], acc[a][c][1], af[a], bf[c]);
                    }
                    __syncthreads();
                }
                // ---- S = C - P, generating C here (its only touch) ----
#pragma unroll
                for (int a = 0; a < 2; a++) {
                    const int rl = warp * 16 + a * 8 + (lane >> 2);
                    const int row = r0 + rl;
                    const double trow = row < N ? __ldg(p.t + row) : 0.0;
#pragma unroll
                    for (int c = 0; c < 4; c++) {
#pragma unroll
                        for (int h = 0; h < 2; h++) {
                            const int cl = c * 8 + 2 * (lane & 3) + h;
                            const int col = j0 + cl;
                            double v = 0.0;
                            if (col < N && row >= col && row < N1) {
                                if (row == N) {
                                    v = W[(size_t)col * ld + N];
                                } else if (row == col) {
                                    int in = multi ? __ldg(p.obsi + row) - 1 : 0;
                                    // eta1^2 * exp(-0 - 0) + sig^2 + jit^2  (:193-206)
                                    v = kadd(g.eta1sq, kadd(__ldg(p.sig2 + row), sm.sc.jit2[in]));
                                } else {
                                    v = qp_kernel(g, kadd(__ldg(p.t + col), -trow));  // t[i]-t[j], i<j (:193)
                                }
                                v -= acc[a][c][h];
                            }
                            sm.Ss[rl][cl] = v;
                        }
                    }
                }
                __syncthreads();
                // ---- diagonal block: unblocked Cholesky by warp 0 (LLT.h:313-337) ----
                if (T == 0) {
                    if (warp == 0) {
                        double ld_acc = 0.0;
                        for (int c = 0; c < nbc; c++) {
                            double d = sm.Ss[c][c];
                            if (!(d > 0.0)) { if (lane == 0) sm.fail = 1; break; }
                            double lcc = sqrt(d);
                            double inv = kdiv(1.0, lcc);
                            if (lane == c) {
                                sm.Ss[c][c] = lcc;
                                sm.dinv[c] = inv;
                                ld_acc += log(lcc);
                            }
                            double x = 0.0;
                            if (lane > c && lane < nbc) {
                                x = sm.Ss[lane][c] * inv;
                                sm.Ss[lane][c] = x;
                            }
                            __syncwarp();
                            if (lane > c && lane < nbc)
                                for (int c2 = c + 1; c2 <= lane; c2++)
                                    sm.Ss[lane][c2] = kfma(-x, sm.Ss[c2][c], sm.Ss[lane][c2]);
                            __syncwarp();
                        }
#pragma unroll
                        for (int o = 16; o > 0; o >>= 1) ld_acc += __shfl_xor_sync(0xffffffffu, ld_acc, o);
                        if (lane == 0) sm.logdet += ld_acc;
                        // keep L_JJ for the remaining tiles of this block column
                        for (int c = 0; c < NB; c++) {
                            double v = 0.0;
                            if (lane < nbc && c < nbc && c <= lane) v = sm.Ss[lane][c];
                            if (lane >= nbc && c == lane) v = 1.0;
                            sm.Ljj[lane][c] = v;
                        }
                        if (lane >= nbc) sm.dinv[lane] = 1.0;
                    }
                    __syncthreads();
                    if (sm.fail) break;
                }
                // ---- X L_JJ^T = S for the rows below the diagonal block; store L ----
                if (tid < R) {
                    const int rl = tid;
                    const int row = r0 + rl;
                    if (row < N1) {
                        const bool in_diag = (T == 0) && rl < nbc;
                        double x[NB];
#pragma unroll
                        for (int c = 0; c < NB; c++) x[c] = sm.Ss[rl][c];
                        if (!in_diag) {
#pragma unroll
                            for (int c = 0; c < NB; c++) {
                                double v = x[c];
#pragma unroll
                                for (int q = 0; q < c; q++) v = kfma(-x[q], sm.Ljj[c][q], v);
                                x[c] = v * sm.dinv[c];
                            }
                        }
#pragma unroll
                        for (int c = 0; c < NB; c++)
                            if (c < nbc) W[(size_t)(j0 + c) * ld + row] = x[c];
                        if (row == N) {
                            double q = 0.0;
#pragma unroll
                            for (int c = 0; c < NB; c++)
                                if (c < nbc) q = kfma(x[c], x[c], q);
                            sm.quad += q;
                        }
                    }
                }
                __syncthreads();
            }
            if (sm.fail) break;
        }
        if (tid == 0) {
            double logL;
            if (sm.fail) {
                logL = __longlong_as_double(0x7ff0000000000000ll);
            } else {
                // -0.5 N log(2 pi) - 0.5 * (2 sum log L_ii) - 0.5 r^T C^-1 r   (:858-870)
                logL = -0.5 * N * 1.8378770664093453 - sm.logdet - 0.5 * sm.quad;
                if (isnan(logL) || isinf(logL)) logL = __longlong_as_double(0x7ff0000000000000ll);
            }
            p.logL[b] = logL;
        }
        __syncthreads();
    }
}

cudaError_t launch_gp(const KParams& p, const GpWorkspace& w, int kepler_mode, int sm_count,
                      cudaStream_t stream, int* n_launches)
{
    if (p.B <= 0) return cudaSuccess;
    size_t smem = sizeof(GpSmem);
    auto k = kepler_mode == 0 ? gp_loglik_kernel<0> : gp_loglik_kernel<1>;
    cudaError_t e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    int grid = w.slots < p.B ? w.slots : p.B;
    e = cudaMemsetAsync(p.counter, 0, sizeof(int), stream);
    if (e != cudaSuccess) return e;
    k<<<grid, CTA_THREADS, smem, stream>>>(p, w.ws, w.slot_doubles);
    *n_launches += 1;
    (void)sm_count;
    return cudaGetLastError();
}

// ---- test hook: the full covariance matrix of one particle (calculate_C) ------------
__global__ void __launch_bounds__(CTA_THREADS) gp_cov_kernel(const KParams p, int b, double* C)
{
    __shared__ ParticleScalars sc;
    __shared__ PlanetInv dummy[1];
    __shared__ GpConsts g;
    if (threadIdx.x < 32) {
        // only the scalars are needed; np is forced to 0 planets worth of work
        const kc_layout& L = p.L;
        if (threadIdx.x == 0) {
            sc.eta1 = ldp(p, L.row_eta, b); sc.eta2 = ldp(p, L.row_eta + 1, b);
            sc.eta3 = ldp(p, L.row_eta + 2, b); sc.eta4 = ldp(p, L.row_eta + 3, b);
            g.eta1sq = kmul(sc.eta1, sc.eta1);
            g.eta2 = sc.eta2; g.r_eta2 = kdiv(1.0, sc.eta2);
            g.eta3 = sc.eta3; g.r_eta3 = kdiv(1.0, sc.eta3);
            g.eta4 = sc.eta4; g.r_eta4 = kdiv(1.0, sc.eta4);
        }
        if (threadIdx.x < KC_MAX_INST) {
            double j2 = 0.0;
            if (threadIdx.x < L.n_jit) { double jt = ldp(p, L.row_jit + threadIdx.x, b); j2 = kmul(jt, jt); }
            sc.jit2[threadIdx.x] = j2;
        }
    }
    (void)dummy;
    __syncthreads();
    const int N = p.N;
    const bool multi = p.F.multi_instrument != 0;
    for (size_t e = (size_t)blockIdx.x * blockDim.x + threadIdx.x; e < (size_t)N * N;
         e += (size_t)gridDim.x * blockDim.x) {
        int i = (int)(e % N), j = (int)(e / N);  // column-major like Eigen::MatrixXd
        int lo = i < j ? i : j, hi = i < j ? j : i;
        double v;
        if (i == j) {
            int in = multi ? p.obsi[i] - 1 : 0;
            v = kadd(g.eta1sq, kadd(p.sig2[i], sc.jit2[in]));
        } else {
            v = qp_kernel(g, kadd(p.t[lo], -p.t[hi]));  // the reference fills (i<j) and mirrors
        }
        C[e] = v;
    }
}

cudaError_t launch_gp_cov(const KParams& p, int lane, double* C, cudaStream_t s)
{
    int grid = (int)(((size_t)p.N * p.N + CTA_THREADS - 1) / CTA_THREADS);
    if (grid > 4096) grid = 4096;
    gp_cov_kernel<<<grid, CTA_THREADS, 0, s>>>(p, lane, C);
    return cudaGetLastError();
}

}  // namespace kc
