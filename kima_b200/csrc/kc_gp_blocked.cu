// Kernel family 2, large-N path: batched blocked Cholesky of the quasi-periodic GP
// covariance, one launch per block column over ALL particles in flight.
//
// Replaces, per particle,  calculate_mu() -> calculate_C() -> log_likelihood()
// (GP / standard-kernel branch) of the reference
// (/root/reference/src/RVmodel.cpp:170-216, 286-375, 841-873; Eigen LLT,
// eigen/Eigen/src/Cholesky/LLT.h:313-368).
//
// Algorithm (left-looking, block column width NB = 64):
//   prep      r = y - mu as row N of the factor storage W (so z = L^-1 r falls out of
//             the factorisation as row N of L and r^T C^-1 r = |z|^2: no solves)
//   for J:    diag    S_JJ = C_JJ - L[J,0:j0] L[J,0:j0]^T, unblocked Cholesky in shared
//                     memory, L_JJ and 1/diag to HBM, log det accumulated
//             update  one CTA per (128-row tile, particle):
//                     S = C[tile,J] - L[tile,0:j0] L[J,0:j0]^T   (DMMA, cp.async pipeline)
//                     X = S L_JJ^-T (forward substitution, a row per thread), X -> HBM
//   finish    logL = -N/2 log 2pi - sum log L_ii - |z|^2/2, non-finite -> +inf
//
// C is never stored: every entry is generated from (t_i - t_j) inside the epilogue of
// the tile that consumes it.  HBM traffic per particle is one write of L (4 N^2 B) and
// N^3/(6 NB) * 8 B of panel re-reads; the FLOPs are the N^3/3 of the factorisation, all
// of the O(N^3) part on the FP64 tensor cores (mma.sync m8n8k4, SASS DMMA).
#include "kc_model.cuh"
#include "kc_gp_common.cuh"

namespace kc {

namespace gpb {
constexpr int NB = 64;        // block-column width
constexpr int R = 128;        // rows per update tile
constexpr int KC = 16;        // k-chunk per pipeline stage
constexpr int STAGES = 4;
constexpr int A_LD = R + 4;   // 132 = 4 mod 16: conflict-free fragment reads (half-warp = 4 k x 4 rows)
constexpr int B_LD = NB + 4;  // 68  = 4 mod 16
constexpr int S_LD = NB + 1;  // row-per-thread access in the substitution
constexpr int LT_LD = NB + 2; // 16-byte aligned rows for paired broadcast reads
constexpr int STAGE_DOUBLES = KC * (A_LD + B_LD);
constexpr int PIPE_DOUBLES = STAGES * STAGE_DOUBLES;
constexpr int EPI_DOUBLES = R * S_LD + NB * LT_LD + NB;
static_assert(EPI_DOUBLES <= PIPE_DOUBLES, "epilogue buffers alias the pipeline stages");
constexpr int D_STAGE_DOUBLES = KC * B_LD;
constexpr int D_SMEM_DOUBLES = STAGES * D_STAGE_DOUBLES + 2 * NB * S_LD;
constexpr int PAD_DOUBLES = 256;  // tiles may read (never use) up to R rows past the last column
}  // namespace gpb

static size_t gpb_ld(int N) { return ((size_t)N + 1 + 15) & ~(size_t)15; }
size_t gpb_slot_doubles(int N) { return gpb_ld(N) * (size_t)N + gpb::PAD_DOUBLES; }
size_t gpb_lane_bytes() { return sizeof(GpLane); }

__device__ __forceinline__ void cp_async16(void* smem, const void* gmem)
{
    unsigned s = (unsigned)__cvta_generic_to_shared(smem);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(s), "l"(gmem) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N_>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N_) : "memory"); }

// ---- prep: per-lane constants + the residual row ------------------------------------------
template <int MODE>
__global__ void __launch_bounds__(CTA_THREADS)
gpb_prep_kernel(const KParams p, double* __restrict__ ws_all, size_t slot_doubles,
                GpLane* __restrict__ lanes, int lane0)
{
    __shared__ PlanetInv inv[MAX_PLANETS];
    __shared__ ParticleScalars sc;
    const int slot = blockIdx.x;
    const int b = lane0 + slot;
    const int tid = threadIdx.x;
    const int N = p.N;
    const size_t ld = ((size_t)N + 1 + 15) & ~(size_t)15;
    double* W = ws_all + (size_t)slot * slot_doubles;
    if (tid < 32) {
        particle_setup(p, b, inv, &sc, tid);
        if (tid == 0) {
            GpLane& g = lanes[slot];
            g.c = gp_consts(sc);
            g.logdet = 0.0;
            g.quad = 0.0;
            g.fail = 0;
        }
        if (tid < KC_MAX_INST) lanes[slot].jit2[tid] = sc.jit2[tid];
    }
    __syncthreads();
    const bool multi = p.F.multi_instrument != 0;
    for (int i0 = tid; i0 < N; i0 += 2 * CTA_THREADS) {
        int i1 = i0 + CTA_THREADS;
        const bool has1 = i1 < N;
        if (!has1) i1 = i0;
        double t0 = __ldg(p.t + i0), t1 = __ldg(p.t + i1);
        int in0 = multi ? __ldg(p.obsi + i0) - 1 : 0;
        int in1 = multi ? __ldg(p.obsi + i1) - 1 : 0;
        double mu0, mu1;
        model_mu2<MODE>(p, &sc, inv, i0, i1, t0, t1, in0, in1, mu0, mu1);
        W[(size_t)i0 * ld + N] = __ldg(p.y + i0) - mu0;  // residual = y - mu (:845-847)
        if (has1) W[(size_t)i1 * ld + N] = __ldg(p.y + i1) - mu1;
    }
}

// ---- diag: S_JJ, its Cholesky factor, log det ------------------------------------------------
// 8 warps = 2 k-halves x 4 row groups; row group w owns fragment rows {w, 7-w} of the
// lower triangle (9 of the 36 lower 8x8 fragments each, so the groups are balanced).
__global__ void __launch_bounds__(CTA_THREADS, 2)
gpb_diag_kernel(const KParams p, double* __restrict__ ws_all, size_t slot_doubles,
                GpLane* __restrict__ lanes, int j0)
{
    using namespace gpb;
    extern __shared__ __align__(16) double dsm[];
    double* pipe = dsm;                                // STAGES x [KC][B_LD]
    double* Ss = dsm + STAGES * D_STAGE_DOUBLES;       // [NB][S_LD]
    double* Sr = Ss + NB * S_LD;                       // second k-half's partial sums
    const int slot = blockIdx.x;
    GpLane& gl = lanes[slot];
    if (gl.fail) return;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int N = p.N;
    const size_t ld = ((size_t)N + 1 + 15) & ~(size_t)15;
    double* W = ws_all + (size_t)slot * slot_doubles;
    const int nbc = min(NB, N - j0);
    const int grp = warp >> 2, w4 = warp & 3;
    const int fr[2] = {w4, 7 - w4};

    double acc[2][8][2];
#pragma unroll
    for (int a = 0; a < 2; a++)
#pragma unroll
        for (int c = 0; c < 8; c++) acc[a][c][0] = acc[a][c][1] = 0.0;

    const int nk = j0 / KC;
    auto load_stage = [&](int s, int it) {
        double* Bs = pipe + s * D_STAGE_DOUBLES;
        const double* src = W + (size_t)(it * KC) * ld + j0;
        // KC x NB doubles = 512 16-byte units, 2 per thread
#pragma unroll
        for (int u = tid; u < KC * NB / 2; u += CTA_THREADS) {
            int kk = u / (NB / 2), r2 = (u % (NB / 2)) * 2;
            cp_async16(Bs + kk * B_LD + r2, src + (size_t)kk * ld + r2);
        }
    };
#pragma unroll
    for (int s = 0; s < STAGES - 1; s++) {
        if (s < nk) load_stage(s, s);
        cp_async_commit();
    }
    for (int it = 0; it < nk; it++) {
        cp_async_wait<STAGES - 2>();
        __syncthreads();
        if (it + STAGES - 1 < nk) load_stage((it + STAGES - 1) % STAGES, it + STAGES - 1);
        cp_async_commit();
        const double* Bs = pipe + (it % STAGES) * D_STAGE_DOUBLES;
#pragma unroll
        for (int k4 = 0; k4 < KC / 8; k4++) {
            const int kk = grp * (KC / 2) + k4 * 4 + (lane & 3);
            double bf[8];
#pragma unroll
            for (int c = 0; c < 8; c++) bf[c] = Bs[kk * B_LD + c * 8 + (lane >> 2)];
#pragma unroll
            for (int a = 0; a < 2; a++) {
                const double af = Bs[kk * B_LD + fr[a] * 8 + (lane >> 2)];
#pragma unroll
                for (int c = 0; c < 8; c++)
                    if (c <= fr[a]) dmma(acc[a][c][0], acc[a][c][1], af, bf[c]);
            }
        }
    }
    cp_async_wait<0>();

    // ---- S = C_JJ - (sum of the two k-halves) ----
    if (grp == 1) {
#pragma unroll
        for (int a = 0; a < 2; a++)
#pragma unroll
            for (int c = 0; c < 8; c++)
                if (c <= fr[a]) {
                    const int r = fr[a] * 8 + (lane >> 2), cc = c * 8 + 2 * (lane & 3);
                    Sr[r * S_LD + cc] = acc[a][c][0];
                    Sr[r * S_LD + cc + 1] = acc[a][c][1];
                }
    }
    __syncthreads();
    if (grp == 0) {
        const GpConsts g = gl.c;
        const bool multi = p.F.multi_instrument != 0;
#pragma unroll
        for (int a = 0; a < 2; a++) {
            const int r = fr[a] * 8 + (lane >> 2);
            const int row = j0 + r;
            const double trow = row < N ? __ldg(p.t + row) : 0.0;
#pragma unroll
            for (int c = 0; c < 8; c++)
                if (c <= fr[a]) {
#pragma unroll
                    for (int h = 0; h < 2; h++) {
                        const int cc = c * 8 + 2 * (lane & 3) + h;
                        const int col = j0 + cc;
                        double v;
                        if (r >= nbc || cc >= nbc) {
                            v = (r == cc) ? 1.0 : 0.0;  // identity padding of a partial block
                        } else if (cc > r) {
                            v = 0.0;
                        } else {
                            if (r == cc) {
                                int in = multi ? __ldg(p.obsi + row) - 1 : 0;
                                // eta1^2 * exp(-0 - 0) + sig^2 + jit^2  (:193-206)
                                v = kadd(g.eta1sq, kadd(__ldg(p.sig2 + row), gl.jit2[in]));
                            } else {
                                v = qp_kernel(g, kadd(__ldg(p.t + col), -trow));
                            }
                            v -= acc[a][c][h] + Sr[r * S_LD + cc];
                        }
                        Ss[r * S_LD + cc] = v;
                    }
                }
        }
    }

    // ---- unblocked right-looking Cholesky of the 64x64 block (LLT.h:313-337) ----
    // Column c of Ss stays unscaled and read-only once step c has begun; the scaled
    // column goes straight to HBM, so one barrier per column suffices.
    double logdet = 0.0;
    bool failed = false;
    for (int c = 0; c < NB; c++) {
        __syncthreads();
        const double d = Ss[c * S_LD + c];
        if (!(d > 0.0)) { failed = true; break; }  // NaN or non-positive pivot: uniform exit
        const double l = sqrt(d);
        const double inv = kdiv(1.0, l);
        if (tid == 0 && c < nbc) {
            logdet += log(l);
            W[(size_t)(j0 + c) * ld + j0 + c] = l;
            gl.dinv[c] = inv;
        }
        if (tid == 0 && c >= nbc) gl.dinv[c] = 1.0;
        // rows r > c: store L(r,c); update the trailing lower triangle
        for (int r = c + 1 + (tid >> 3); r < NB; r += CTA_THREADS / 8) {
            const double lr = Ss[r * S_LD + c] * inv;
            if ((tid & 7) == 0 && r < nbc) W[(size_t)(j0 + c) * ld + j0 + r] = lr;
            for (int c2 = c + 1 + (tid & 7); c2 <= r; c2 += 8) {
                const double lc = Ss[c2 * S_LD + c] * inv;
                Ss[r * S_LD + c2] = kfma(-lr, lc, Ss[r * S_LD + c2]);
            }
        }
    }
    if (tid == 0) {
        if (failed) gl.fail = 1;
        else gl.logdet += logdet;
    }
}

// ---- update: the panel below the diagonal block ----------------------------------------------
__global__ void __launch_bounds__(CTA_THREADS, 2)
gpb_update_kernel(const KParams p, double* __restrict__ ws_all, size_t slot_doubles,
                  GpLane* __restrict__ lanes, int j0, int first_row)
{
    using namespace gpb;
    extern __shared__ __align__(16) double usm[];
    const int slot = blockIdx.y;
    GpLane& gl = lanes[slot];
    if (gl.fail) return;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int N = p.N;
    const size_t ld = ((size_t)N + 1 + 15) & ~(size_t)15;
    double* W = ws_all + (size_t)slot * slot_doubles;
    const int nbc = min(NB, N - j0);
    const int r0 = first_row + blockIdx.x * R;
    const int row_lo = j0 + nbc;  // rows below the diagonal block; row N is the residual
    const int wr = warp >> 1, wc = warp & 1;

    double acc[4][4][2];
#pragma unroll
    for (int a = 0; a < 4; a++)
#pragma unroll
        for (int c = 0; c < 4; c++) acc[a][c][0] = acc[a][c][1] = 0.0;

    const int nk = j0 / KC;
    auto load_stage = [&](int s, int it) {
        double* As = usm + s * STAGE_DOUBLES;
        double* Bs = As + KC * A_LD;
        const double* col0 = W + (size_t)(it * KC) * ld;
        // A: KC x R doubles = 1024 16-byte units (4 per thread); B: KC x NB = 512 (2 per thread)
#pragma unroll
        for (int u = tid; u < KC * R / 2; u += CTA_THREADS) {
            int kk = u / (R / 2), r2 = (u % (R / 2)) * 2;
            cp_async16(As + kk * A_LD + r2, col0 + (size_t)kk * ld + r0 + r2);
        }
#pragma unroll
        for (int u = tid; u < KC * NB / 2; u += CTA_THREADS) {
            int kk = u / (NB / 2), c2 = (u % (NB / 2)) * 2;
            cp_async16(Bs + kk * B_LD + c2, col0 + (size_t)kk * ld + j0 + c2);
        }
    };
#pragma unroll
    for (int s = 0; s < STAGES - 1; s++) {
        if (s < nk) load_stage(s, s);
        cp_async_commit();
    }
    for (int it = 0; it < nk; it++) {
        cp_async_wait<STAGES - 2>();
        __syncthreads();
        if (it + STAGES - 1 < nk) load_stage((it + STAGES - 1) % STAGES, it + STAGES - 1);
        cp_async_commit();
        const double* As = usm + (it % STAGES) * STAGE_DOUBLES;
        const double* Bs = As + KC * A_LD;
#pragma unroll
        for (int k4 = 0; k4 < KC; k4 += 4) {
            const int kk = k4 + (lane & 3);
            double af[4], bf[4];
#pragma unroll
            for (int a = 0; a < 4; a++) af[a] = As[kk * A_LD + wr * 32 + a * 8 + (lane >> 2)];
#pragma unroll
            for (int c = 0; c < 4; c++) bf[c] = Bs[kk * B_LD + wc * 32 + c * 8 + (lane >> 2)];
#pragma unroll
            for (int a = 0; a < 4; a++)
#pragma unroll
                for (int c = 0; c < 4; c++) dmma(acc[a][c][0], acc[a][c][1], af[a], bf[c]);
        }
    }
    cp_async_wait<0>();
    __syncthreads();  // every warp is done with the stages: the epilogue buffers alias them

    double* Ss = usm;                    // [R][S_LD]
    double* LT = usm + R * S_LD;         // [NB][LT_LD]: LT[c][c2] = L_JJ(c2, c)
    double* dinv = LT + NB * LT_LD;      // [NB]

    // ---- S = C - P, generating C here (its only touch) ----
    {
        const GpConsts g = gl.c;
#pragma unroll
        for (int a = 0; a < 4; a++) {
            const int rl = wr * 32 + a * 8 + (lane >> 2);
            const int row = r0 + rl;
            const bool row_ok = row >= row_lo && row <= N;
            const double trow = (row_ok && row < N) ? __ldg(p.t + row) : 0.0;
#pragma unroll
            for (int c = 0; c < 4; c++) {
#pragma unroll
                for (int h = 0; h < 2; h++) {
                    const int cl = wc * 32 + c * 8 + 2 * (lane & 3) + h;
                    const int col = j0 + cl;
                    double v = 0.0;
                    if (row_ok && cl < nbc) {
                        if (row == N) v = W[(size_t)col * ld + N];
                        else v = qp_kernel(g, kadd(__ldg(p.t + col), -trow));  // t[i]-t[j], i<j (:193)
                        v -= acc[a][c][h];
                    }
                    Ss[rl * S_LD + cl] = v;
                }
            }
        }
    }
    // L_JJ (transposed) and the reciprocal diagonal, identity-padded for a partial block
    for (int e = tid; e < NB * NB; e += CTA_THREADS) {
        const int c = e / NB, c2 = e % NB;  // consecutive threads -> consecutive rows of one column
        double v = 0.0;
        if (c2 > c && c2 < nbc) v = W[(size_t)(j0 + c) * ld + j0 + c2];
        LT[c * LT_LD + c2] = v;
    }
    if (tid < NB) dinv[tid] = gl.dinv[tid];
    __syncthreads();

    // ---- X L_JJ^T = S by forward substitution, a row per thread; store L ----
    if (tid < R) {
        const int row = r0 + tid;
        if (row >= row_lo && row <= N) {
            double* srow = Ss + tid * S_LD;
            double q = 0.0;
            {
                double x[32];
#pragma unroll
                for (int c = 0; c < 32; c++) x[c] = srow[c];
#pragma unroll
                for (int c = 0; c < 32; c++) {
                    x[c] *= dinv[c];
#pragma unroll
                    for (int c2 = c + 1; c2 < 32; c2++) x[c2] = kfma(-x[c], LT[c * LT_LD + c2], x[c2]);
                }
#pragma unroll
                for (int c = 0; c < 32; c++) {
                    srow[c] = x[c];
                    if (c < nbc) {
                        W[(size_t)(j0 + c) * ld + row] = x[c];
                        q = kfma(x[c], x[c], q);
                    }
                }
            }
            {
                double y[32];
#pragma unroll
                for (int c = 0; c < 32; c++) y[c] = srow[32 + c];
                for (int k = 0; k < 32; k++) {
                    const double xk = srow[k];
                    const double* lt = LT + k * LT_LD + 32;
#pragma unroll
                    for (int c = 0; c < 32; c++) y[c] = kfma(-xk, lt[c], y[c]);
                }
#pragma unroll
                for (int c = 0; c < 32; c++) {
                    y[c] *= dinv[32 + c];
#pragma unroll
                    for (int c2 = c + 1; c2 < 32; c2++)
                        y[c2] = kfma(-y[c], LT[(32 + c) * LT_LD + 32 + c2], y[c2]);
                }
#pragma unroll
                for (int c = 0; c < 32; c++)
                    if (32 + c < nbc) {
                        W[(size_t)(j0 + 32 + c) * ld + row] = y[c];
                        q = kfma(y[c], y[c], q);
                    }
            }
            if (row == N) gl.quad += q;  // |z|^2 of this block column
        }
    }
}

__global__ void gpb_finish_kernel(const KParams p, const GpLane* __restrict__ lanes, int lane0, int n)
{
    int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n) return;
    const GpLane& g = lanes[s];
    double logL;
    const double pinf = __longlong_as_double(0x7ff0000000000000ll);
    if (g.fail) {
        logL = pinf;
    } else {
        // -0.5 N log(2 pi) - 0.5 * (2 sum log L_ii) - 0.5 r^T C^-1 r   (:858-870)
        logL = -0.5 * p.N * 1.8378770664093453 - g.logdet - 0.5 * g.quad;
        if (isnan(logL) || isinf(logL)) logL = pinf;
    }
    p.logL[lane0 + s] = logL;
}

cudaError_t launch_gp_blocked(const KParams& p, const GpWorkspace& w, int kepler_mode,
                              cudaStream_t stream, int* n_launches)
{
    using namespace gpb;
    if (p.B <= 0) return cudaSuccess;
    const size_t usmem = sizeof(double) * PIPE_DOUBLES;
    const size_t dsmem = sizeof(double) * D_SMEM_DOUBLES;
    cudaError_t e;
    e = cudaFuncSetAttribute(gpb_update_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)usmem);
    if (e != cudaSuccess) return e;
    e = cudaFuncSetAttribute(gpb_diag_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dsmem);
    if (e != cudaSuccess) return e;
    const int N = p.N;
    const int nblk = (N + NB - 1) / NB;
    for (int lane0 = 0; lane0 < p.B; lane0 += w.slots) {
        const int n = min(w.slots, p.B - lane0);
        if (kepler_mode == 0)
            gpb_prep_kernel<0><<<n, CTA_THREADS, 0, stream>>>(p, w.ws, w.slot_doubles, w.lanes, lane0);
        else
            gpb_prep_kernel<1><<<n, CTA_THREADS, 0, stream>>>(p, w.ws, w.slot_doubles, w.lanes, lane0);
        *n_launches += 1;
        for (int J = 0; J < nblk; J++) {
            const int j0 = J * NB;
            const int nbc = min(NB, N - j0);
            gpb_diag_kernel<<<n, CTA_THREADS, dsmem, stream>>>(p, w.ws, w.slot_doubles, w.lanes, j0);
            // rows below the diagonal block, up to and including the residual row N
            const int first_row = nbc == NB ? j0 + NB : j0;
            const int ntile = (N + 1 - first_row + R - 1) / R;
            dim3 grid(ntile, n);
            gpb_update_kernel<<<grid, CTA_THREADS, usmem, stream>>>(p, w.ws, w.slot_doubles, w.lanes, j0,
                                                                   first_row);
            *n_launches += 2;
        }
        gpb_finish_kernel<<<(n + 127) / 128, 128, 0, stream>>>(p, w.lanes, lane0, n);
        *n_launches += 1;
        e = cudaGetLastError();
        if (e != cudaSuccess) return e;
    }
    return cudaGetLastError();
}

}  // namespace kc
