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
constexpr int STAGES = 3;
constexpr int A_LD = R + 4;   // 132 = 4 mod 16: conflict-free fragment reads (half-warp = 4 k x 4 rows)
constexpr int B_LD = NB + 4;  // 68  = 4 mod 16
constexpr int S_LD = NB + 1;  // diag kernel: row-per-lane access
constexpr int STAGE_DOUBLES = KC * (A_LD + B_LD);
constexpr int PIPE_DOUBLES = STAGES * STAGE_DOUBLES;
static_assert(NB * A_LD <= PIPE_DOUBLES, "the S tile aliases the pipeline stages");
constexpr int U_SMEM_DOUBLES = PIPE_DOUBLES + NB * B_LD + NB;  // + L_JJ^T + 1/diag, prefetched
constexpr int D_THREADS = 128;
constexpr int D_STAGES = 3;
constexpr int H_LD = 34;      // 32x32 half-block buffers of the diag kernel
constexpr int D_STAGE_DOUBLES = KC * B_LD;
constexpr int D_SMEM_DOUBLES = D_STAGES * D_STAGE_DOUBLES + NB * S_LD;
static_assert(2 * 32 * H_LD <= D_STAGES * D_STAGE_DOUBLES, "half-block buffers alias the diag pipeline");
constexpr int PAD_DOUBLES = 256;  // tiles may read (never use) up to R rows past the last column
}  // namespace gpb

// What one block-column launch works on.
struct GpbStep {
    int j0;         // first column of the block column
    int n;          // lanes (= slots) in flight
    int first_row;  // update kernel: first row of the first tile
    int ntile;      // update kernel: row tiles per lane
};

size_t gpb_ld(int N) { return ((size_t)N + 1 + 15) & ~(size_t)15; }
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

// Barrier over the 128 threads (warps 0-3) that factor a diagonal block, whether they are
// the whole diag CTA or half of an update CTA.
__device__ __forceinline__ void bar_diag() { asm volatile("bar.sync 1, 128;" ::: "memory"); }

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
    if (p.F.MA) {
        // residual with the MA(1) term (:378-386): recurrence in shared memory, then row N
        extern __shared__ __align__(16) double ma_buf[];
        ma_residuals<MODE>(p, &sc, inv, ma_buf, ma_buf + N, tid, CTA_THREADS);
        for (int i = tid; i < N; i += CTA_THREADS) W[(size_t)i * ld + N] = ma_buf[i];
        return;
    }
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
// 128 threads.  K loop: warp w owns fragment rows {w, 7-w} of the lower triangle (9 of the
// 36 lower 8x8 fragments each, so the warps are balanced); the A fragment of row f is the B
// fragment of column f, so 8 shared-memory reads feed 9 DMMAs.
// Factorisation: 64x64 = two 32x32 Cholesky steps with the rows of the block in the lanes
// of ONE warp (pivot and column broadcast by shuffles, no block-wide barrier per column),
// a row-per-lane triangular solve for L21 and a 4-warp rank-32 update in between.

// K loop of the diag kernel for the warp that owns fragment rows F0 and 7-F0 (all
// fragment indices compile-time: 9 DMMAs per k-step, no predication), then P -> Ss.
template <int F0>
__device__ __forceinline__ void diag_kloop(const double* __restrict__ W, size_t ld, int j0,
                                           double* pipe, double* Ss, int tid, int lane)
{
    using namespace gpb;
    constexpr int F1 = 7 - F0;
    double acc0[F0 + 1][2], acc1[F1 + 1][2];
#pragma unroll
    for (int c = 0; c <= F0; c++) acc0[c][0] = acc0[c][1] = 0.0;
#pragma unroll
    for (int c = 0; c <= F1; c++) acc1[c][0] = acc1[c][1] = 0.0;
    const int nk = j0 / KC;
    auto load_stage = [&](int s, int it) {
        double* Bs = pipe + s * D_STAGE_DOUBLES;
        const double* src = W + (size_t)(it * KC) * ld + j0;
        // KC x NB doubles = 512 16-byte units, 4 per thread
#pragma unroll
        for (int u = tid; u < KC * NB / 2; u += D_THREADS) {
            int kk = u / (NB / 2), r2 = (u % (NB / 2)) * 2;
            cp_async16(Bs + kk * B_LD + r2, src + (size_t)kk * ld + r2);
        }
    };
#pragma unroll
    for (int s = 0; s < D_STAGES - 1; s++) {
        if (s < nk) load_stage(s, s);
        cp_async_commit();
    }
    for (int it = 0; it < nk; it++) {
        cp_async_wait<D_STAGES - 2>();
        bar_diag();
        if (it + D_STAGES - 1 < nk) load_stage((it + D_STAGES - 1) % D_STAGES, it + D_STAGES - 1);
        cp_async_commit();
        const double* Bs = pipe + (it % D_STAGES) * D_STAGE_DOUBLES;
#pragma unroll
        for (int k4 = 0; k4 < KC; k4 += 4) {
            const int kk = k4 + (lane & 3);
            double bf[F1 + 1];
#pragma unroll
            for (int c = 0; c <= F1; c++) bf[c] = Bs[kk * B_LD + c * 8 + (lane >> 2)];
#pragma unroll
            for (int c = 0; c <= F0; c++) dmma(acc0[c][0], acc0[c][1], bf[F0], bf[c]);
#pragma unroll
            for (int c = 0; c <= F1; c++) dmma(acc1[c][0], acc1[c][1], bf[F1], bf[c]);
        }
    }
    cp_async_wait<0>();
#pragma unroll
    for (int c = 0; c <= F0; c++) {
        const int r = F0 * 8 + (lane >> 2), cc = c * 8 + 2 * (lane & 3);
        Ss[r * S_LD + cc] = acc0[c][0];
        Ss[r * S_LD + cc + 1] = acc0[c][1];
    }
#pragma unroll
    for (int c = 0; c <= F1; c++) {
        const int r = F1 * 8 + (lane >> 2), cc = c * 8 + 2 * (lane & 3);
        Ss[r * S_LD + cc] = acc1[c][0];
        Ss[r * S_LD + cc + 1] = acc1[c][1];
    }
}

// Factor the diagonal block that starts at column j0 (threads 0..127 of the CTA; `dsm` is
// D_SMEM_DOUBLES of shared memory they own).  Called by the diag kernel for the first block
// column and by the update kernel's first row tile for every later one: that CTA has just
// produced the 64 rows of L the block needs, and folding the (latency-bound) factorisation
// into the update launch lets the other CTA on the SM keep the FP64 pipe busy meanwhile.
__device__ __forceinline__ void diag_block(const KParams& p, double* __restrict__ W, size_t ld,
                                           GpLane& gl, int j0, double* dsm, int tid)
{
    using namespace gpb;
    double* pipe = dsm;                                // D_STAGES x [KC][B_LD]
    double* Ss = dsm + D_STAGES * D_STAGE_DOUBLES;     // [NB][S_LD]
    double* LTs = pipe;                                // after the K loop: [32][H_LD], LTs[c][c2] = L11(c2,c)
    double* T21 = pipe + 32 * H_LD;                    // [32][H_LD]: L21 row-major
    __shared__ double dinv_s[NB];
    __shared__ __align__(16) double colbuf_s[CHOL_COLBUF];
    __shared__ int fail_s;
    const int warp = tid >> 5, lane = tid & 31;
    const int N = p.N;
    const int nbc = min(NB, N - j0);
    if (tid == 0) fail_s = 0;
    switch (warp) {
        case 0: diag_kloop<0>(W, ld, j0, pipe, Ss, tid, lane); break;
        case 1: diag_kloop<1>(W, ld, j0, pipe, Ss, tid, lane); break;
        case 2: diag_kloop<2>(W, ld, j0, pipe, Ss, tid, lane); break;
        default: diag_kloop<3>(W, ld, j0, pipe, Ss, tid, lane); break;
    }
    bar_diag();
    {
        const GpConsts g = gl.c;
        const bool multi = p.F.multi_instrument != 0;
        const int r = tid & (NB - 1);
        const int row = j0 + r;
        const double trow = row < N ? __ldg(p.t + row) : 0.0;
        for (int cc = tid >> 6; cc < NB; cc += D_THREADS / NB) {
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
                    v = qp_kernel(g, kadd(__ldg(p.t + j0 + cc), -trow));
                }
                v -= Ss[r * S_LD + cc];
            }
            Ss[r * S_LD + cc] = v;
        }
    }
    bar_diag();

    double logdet = 0.0;
#pragma unroll 1
    for (int half = 0; half < 2; half++) {
        const int h0 = 32 * half;
        const int nv = max(0, min(32, nbc - h0));
        if (warp == 0) {
            double s[32];
#pragma unroll
            for (int c = 0; c < 32; c++) s[c] = Ss[(h0 + lane) * S_LD + h0 + c];
            if (!chol32_rows_in_lanes(s, lane, nv, logdet, dinv_s + h0,
                                      W + (size_t)(j0 + h0) * ld + j0 + h0, ld,
                                      half == 0 ? LTs : nullptr, H_LD, colbuf_s) && lane == 0)
                fail_s = 1;
        }
        bar_diag();
        if (half == 0) {
            if (warp == 1) {
                // L21 = S21 L11^-T: row 32+lane, forward substitution
                double t[32];
#pragma unroll
                for (int c = 0; c < 32; c++) t[c] = Ss[(32 + lane) * S_LD + c];
#pragma unroll
                for (int c = 0; c < 32; c++) {
                    t[c] *= dinv_s[c];
#pragma unroll
                    for (int c2 = c + 1; c2 < 32; c2++) t[c2] = kfma(-t[c], LTs[c * H_LD + c2], t[c2]);
                }
#pragma unroll
                for (int c = 0; c < 32; c++) {
                    T21[lane * H_LD + c] = t[c];
                    if (32 + lane < nbc) W[(size_t)(j0 + c) * ld + j0 + 32 + lane] = t[c];
                }
            }
            bar_diag();
            {
                // S22 -= L21 L21^T: warp w takes columns [8w, 8w+8), lane = row
                double t[32];
#pragma unroll
                for (int k = 0; k < 32; k++) t[k] = T21[lane * H_LD + k];
#pragma unroll
                for (int q = 0; q < 8; q++) {
                    const int c2 = 8 * warp + q;
                    const double* o = T21 + c2 * H_LD;
                    double a0 = 0.0, a1 = 0.0;
#pragma unroll
                    for (int k = 0; k < 32; k += 2) {
                        a0 = kfma(t[k], o[k], a0);
                        a1 = kfma(t[k + 1], o[k + 1], a1);
                    }
                    Ss[(32 + lane) * S_LD + 32 + c2] -= a0 + a1;
                }
            }
            bar_diag();
        }
    }
    if (tid < NB) gl.dinv[(j0 / NB) & 1][tid] = dinv_s[tid];  // double-buffered by block parity
    if (tid == 0) {
        if (fail_s) gl.fail = 1;
        else gl.logdet += logdet;
    }
}

__global__ void __launch_bounds__(gpb::D_THREADS, 3)
gpb_diag_kernel(const KParams p, double* __restrict__ ws_all, size_t slot_doubles,
                GpLane* __restrict__ lanes, const GpbStep w)
{
    extern __shared__ __align__(16) double dsm_k[];
    const int slot = blockIdx.x;
    GpLane& gl = lanes[slot];
    if (gl.fail) return;
    const size_t ld = ((size_t)p.N + 1 + 15) & ~(size_t)15;
    diag_block(p, ws_all + (size_t)slot * slot_doubles, ld, gl, w.j0, dsm_k, threadIdx.x);
}


// ---- update: the panel below the diagonal block ----------------------------------------------
__global__ void __launch_bounds__(CTA_THREADS, 2)
gpb_update_kernel(const KParams p, double* __restrict__ ws_all, size_t slot_doubles,
                  GpLane* __restrict__ lanes, const GpbStep w)
{
    using namespace gpb;
    extern __shared__ __align__(16) double usm[];
    const int j0 = w.j0, first_row = w.first_row;
    const int slot = blockIdx.y;
    const int tile = blockIdx.x;
    GpLane& gl = lanes[slot];
    if (gl.fail) return;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int N = p.N;
    const size_t ld = ((size_t)N + 1 + 15) & ~(size_t)15;
    double* W = ws_all + (size_t)slot * slot_doubles;
    const int nbc = min(NB, N - j0);
    const int r0 = first_row + tile * R;
    const int row_lo = j0 + nbc;  // rows below the diagonal block; row N is the residual
    const int wr = warp >> 1, wc = warp & 1;

    double acc[4][4][2];
#pragma unroll
    for (int a = 0; a < 4; a++)
#pragma unroll
        for (int c = 0; c < 4; c++) acc[a][c][0] = acc[a][c][1] = 0.0;

    // L_JJ (as stored: LT[c][c2] = L_JJ(c2, c), only c2 > c is ever read) and 1/diag ride in
    // as the oldest cp.async group, hidden behind the K loop
    double* LT = usm + PIPE_DOUBLES;     // [NB][B_LD]
    double* dinv = LT + NB * B_LD;       // [NB]
    for (int u = tid; u < nbc * (NB / 2); u += CTA_THREADS) {
        const int c = u / (NB / 2), c2 = (u % (NB / 2)) * 2;
        cp_async16(LT + c * B_LD + c2, W + (size_t)(j0 + c) * ld + j0 + c2);
    }
    if (tid < NB / 2) cp_async16(dinv + 2 * tid, &gl.dinv[(j0 / NB) & 1][2 * tid]);
    cp_async_commit();

    const int nk = j0 / KC;
    const bool warp_rows_live = r0 + wr * 32 <= N;
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
        // a warp whose 32 rows are all past row N (partial last tile): loads and barriers only
#pragma unroll
        for (int k4 = 0; k4 < (warp_rows_live ? KC : 0); k4 += 4) {
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

    double* SsT = usm;                   // [NB][A_LD]: the tile, column-major (S, then X in place)

    // P (fragment layout) -> shared memory, so that the rest is rolled loops over the tile
#pragma unroll
    for (int a = 0; a < 4; a++)
#pragma unroll
        for (int c = 0; c < 4; c++)
#pragma unroll
            for (int h = 0; h < 2; h++)
                SsT[(wc * 32 + c * 8 + 2 * (lane & 3) + h) * A_LD + wr * 32 + a * 8 + (lane >> 2)] = acc[a][c][h];
    __syncthreads();

    // ---- S = C - P, generating C here (its only touch) ----
    {
        const GpConsts g = gl.c;
        const int rl = tid & (R - 1);
        const int row = r0 + rl;
        const bool row_ok = row >= row_lo && row <= N;
        const double trow = (row_ok && row < N) ? __ldg(p.t + row) : 0.0;
        // two columns per pass (cl, cl + 2): two independent chains per thread
        for (int cl = tid >> 7; cl < NB; cl += 2 * CTA_THREADS / R) {
            const int cl1 = cl + CTA_THREADS / R;
            const int col = j0 + cl, col1 = j0 + cl1;
            double v0 = 0.0, v1 = 0.0;
            if (row_ok) {
                if (row == N) {
                    if (cl < nbc) v0 = W[(size_t)col * ld + N];
                    if (cl1 < nbc) v1 = W[(size_t)col1 * ld + N];
                } else {
                    // t[i]-t[j], i<j (:193); columns past a partial block are generated and dropped
                    const double tc0 = __ldg(p.t + min(col, N - 1)), tc1 = __ldg(p.t + min(col1, N - 1));
                    qp_kernel2(g, kadd(tc0, -trow), kadd(tc1, -trow), v0, v1);
                }
                v0 = cl < nbc ? v0 - SsT[cl * A_LD + rl] : 0.0;
                v1 = cl1 < nbc ? v1 - SsT[cl1 * A_LD + rl] : 0.0;
            }
            SsT[cl * A_LD + rl] = v0;
            SsT[cl1 * A_LD + rl] = v1;
        }
    }
    __syncthreads();

    // ---- X L_JJ^T = S: each warp owns 16 rows.  Blocked forward substitution over four
    //      16-column blocks: the off-diagonal part on the tensor cores, the 16x16 diagonal
    //      blocks by plain substitution (a row per lane).  Only warp-level syncs. ----
    {
        const int rb = warp * 16;
        for (int jb = 0; jb < 4; jb++) {
            if (jb > 0) {
                double u[2][2][2];
#pragma unroll
                for (int m = 0; m < 2; m++)
#pragma unroll
                    for (int n = 0; n < 2; n++) u[m][n][0] = u[m][n][1] = 0.0;
                for (int kb = 0; kb < jb * 16; kb += 4) {
                    const int kk = kb + (lane & 3);
                    double af[2], bf[2];
#pragma unroll
                    for (int m = 0; m < 2; m++) af[m] = SsT[kk * A_LD + rb + m * 8 + (lane >> 2)];
#pragma unroll
                    for (int n = 0; n < 2; n++) bf[n] = LT[kk * B_LD + jb * 16 + n * 8 + (lane >> 2)];
#pragma unroll
                    for (int m = 0; m < 2; m++)
#pragma unroll
                        for (int n = 0; n < 2; n++) dmma(u[m][n][0], u[m][n][1], af[m], bf[n]);
                }
#pragma unroll
                for (int m = 0; m < 2; m++)
#pragma unroll
                    for (int n = 0; n < 2; n++)
#pragma unroll
                        for (int h = 0; h < 2; h++) {
                            double* q = SsT + (jb * 16 + n * 8 + 2 * (lane & 3) + h) * A_LD + rb + m * 8 + (lane >> 2);
                            *q -= u[m][n][h];
                        }
                __syncwarp();
            }
            if (lane < 16) {
                double x[16];
                double* col0 = SsT + (jb * 16) * A_LD + rb + lane;
                const double* l0 = LT + (jb * 16) * B_LD + jb * 16;
#pragma unroll
                for (int c = 0; c < 16; c++) x[c] = col0[c * A_LD];
#pragma unroll
                for (int c = 0; c < 16; c++) {
                    x[c] *= dinv[jb * 16 + c];
#pragma unroll
                    for (int c2 = c + 1; c2 < 16; c2++) x[c2] = kfma(-x[c], l0[c * B_LD + c2], x[c2]);
                }
#pragma unroll
                for (int c = 0; c < 16; c++) col0[c * A_LD] = x[c];
            }
            __syncwarp();
        }
        // store the warp's 16 rows: lanes 0-15 an even column, lanes 16-31 the odd one
        const int rl = rb + (lane & 15);
        const int row = r0 + rl;
        const bool ok = row >= row_lo && row <= N;
        double q = 0.0;
        for (int cp = 0; cp < NB / 2; cp++) {
            const int c = 2 * cp + (lane >> 4);
            if (ok && c < nbc) {
                const double x = SsT[c * A_LD + rl];
                W[(size_t)(j0 + c) * ld + row] = x;
                q = kfma(x, x, q);
            }
        }
        q += __shfl_xor_sync(0xffffffffu, q, 16);
        if (row == N && lane < 16) gl.quad += q;  // |z|^2 of this block column
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

cudaError_t launch_gp_blocked(const KParams& p, const GpWorkspace& w, int kepler_mode, int sm_count,
                              cudaStream_t stream, int* n_launches, const GpPredict* pred)
{
    using namespace gpb;
    if (p.B <= 0) return cudaSuccess;
    const size_t usmem = sizeof(double) * U_SMEM_DOUBLES;
    const size_t dsmem = sizeof(double) * D_SMEM_DOUBLES;
    cudaError_t e;
    e = cudaFuncSetAttribute(gpb_update_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)usmem);
    if (e != cudaSuccess) return e;
    e = cudaFuncSetAttribute(gpb_diag_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dsmem);
    if (e != cudaSuccess) return e;
    const int N = p.N;
    const size_t ma_smem = p.F.MA ? 2 * sizeof(double) * (size_t)N : 0;
    if (ma_smem > 160 * 1024) return cudaErrorInvalidValue;
    if (ma_smem > 48 * 1024) {
        e = cudaFuncSetAttribute(gpb_prep_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ma_smem);
        if (e != cudaSuccess) return e;
        e = cudaFuncSetAttribute(gpb_prep_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ma_smem);
        if (e != cudaSuccess) return e;
    }
    const int nblk = (N + NB - 1) / NB;
    // The lanes of a wave are cut into up to nstreams groups, each factorised on its own
    // stream with its own share of the slots: the groups drift apart, and one group's
    // latency-bound diagonal blocks (and launch tails) run beside another's update tiles.
    int ng = w.nstreams < 1 ? 1 : w.nstreams;
    while (ng > 1 && (p.B < ng * sm_count || w.slots < ng)) ng--;  // small batches: one group
    // lanes go in gridDim.y of the update kernel (and of kc_predict's mean kernel): at most 65535
    const int gslots = min(w.slots / ng, 65535);
    const int per_group = (p.B + ng - 1) / ng;
    if (ng > 1) {
        e = cudaEventRecord(w.fork, stream);
        if (e != cudaSuccess) return e;
    }
    for (int g = 0; g < ng; g++) {
        cudaStream_t st_g = g == 0 ? stream : w.aux[g - 1];
        if (g > 0) {
            e = cudaStreamWaitEvent(st_g, w.fork, 0);
            if (e != cudaSuccess) return e;
        }
        double* ws_g = w.ws + (size_t)g * gslots * w.slot_doubles;
        GpLane* lanes_g = w.lanes + (size_t)g * gslots;
        const int lane_end = min(p.B, (g + 1) * per_group);
        // waves of equal size (342 + 341 + 341 lanes rather than 476 + 476 + 72)
        const int cnt = lane_end - g * per_group;
        const int nwaves = (cnt + gslots - 1) / gslots;
        const int per_wave = nwaves > 0 ? (cnt + nwaves - 1) / nwaves : 0;
        for (int lane0 = g * per_group; lane0 < lane_end; lane0 += per_wave) {
            const int n = min(per_wave, lane_end - lane0);
            if (kepler_mode == 0)
                gpb_prep_kernel<0><<<n, CTA_THREADS, ma_smem, st_g>>>(p, ws_g, w.slot_doubles, lanes_g, lane0);
            else
                gpb_prep_kernel<1><<<n, CTA_THREADS, ma_smem, st_g>>>(p, ws_g, w.slot_doubles, lanes_g, lane0);
            *n_launches += 1;
            for (int J = 0; J < nblk; J++) {
                GpbStep st;
                st.j0 = J * NB;
                st.n = n;
                const int nbc = min(NB, N - st.j0);
                // rows below the diagonal block, up to and including the residual row N
                st.first_row = nbc == NB ? st.j0 + NB : st.j0;
                st.ntile = (N + 1 - st.first_row + R - 1) / R;
                gpb_diag_kernel<<<n, D_THREADS, dsmem, st_g>>>(p, ws_g, w.slot_doubles, lanes_g, st);
                gpb_update_kernel<<<dim3(st.ntile, n), CTA_THREADS, usmem, st_g>>>(p, ws_g, w.slot_doubles, lanes_g, st);
                *n_launches += 2;
            }
            gpb_finish_kernel<<<(n + 127) / 128, 128, 0, st_g>>>(p, lanes_g, lane0, n);
            *n_launches += 1;
            e = cudaGetLastError();
            if (e != cudaSuccess) return e;
            if (pred) {  // the wave's factors are about to be overwritten by the next one
                e = launch_gp_predict_wave(p, ws_g, w.slot_doubles, lanes_g, lane0, n, *pred, st_g, n_launches);
                if (e != cudaSuccess) return e;
            }
        }
        if (g > 0) {
            e = cudaEventRecord(w.join[g - 1], st_g);
            if (e != cudaSuccess) return e;
            e = cudaStreamWaitEvent(stream, w.join[g - 1], 0);
            if (e != cudaSuccess) return e;
        }
    }
    return cudaGetLastError();
}

}  // namespace kc
