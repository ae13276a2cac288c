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
//                     S = C[tile,J] - L[tile,0:j0] L[J,0:j0]^T   (DMMA, TMA + mbarrier pipeline)
//                     X = S L_JJ^-T (forward substitution, a row per thread), X -> HBM
//   finish    logL = -N/2 log 2pi - sum log L_ii - |z|^2/2, non-finite -> +inf
//
// C is never stored: every entry is generated inside the epilogue of the tile that consumes it, from
// (t_i - t_j) and the epochs' phases (kc_gp_common.cuh: gp_phase / qp_entry; the prep kernel leaves
// the 2N sines and cosines of a lane in two extra columns of its factor storage).  HBM traffic per particle is one write of L (4 N^2 B) and
// N^3/(6 NB) * 8 B of panel re-reads; the FLOPs are the N^3/3 of the factorisation, all
// of the O(N^3) part on the FP64 tensor cores (mma.sync m8n8k4, SASS DMMA).
#include <cuda.h>  // CUtensorMap (types only: the encoder is fetched through cudaGetDriverEntryPoint)
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
constexpr int S_LD = NB + 4;  // diag kernel: S_JJ column-major, element (r, c) at Ss[c * S_LD + r]; 68 = 4 mod 16
constexpr int STAGE_DOUBLES = KC * (A_LD + B_LD);
constexpr int PIPE_DOUBLES = STAGES * STAGE_DOUBLES;
static_assert(NB * A_LD <= PIPE_DOUBLES, "the S tile aliases the pipeline stages");
// the epilogue's buffers (the S tile, then L_JJ^T and 1/diag behind it) alias the pipeline stages:
// L_JJ is fetched once the K loop has drained them, its latency hidden behind the covariance generation
constexpr int EPI_DOUBLES = NB * A_LD + NB * B_LD + NB;
constexpr int U_SMEM_DOUBLES = PIPE_DOUBLES > EPI_DOUBLES ? PIPE_DOUBLES : EPI_DOUBLES;
constexpr int D_THREADS = 128;
constexpr int D_STAGES = 3;
constexpr int H_LD = 34;      // 32x32 half-block buffers of the diag kernel
constexpr int D_STAGE_DOUBLES = KC * B_LD;
constexpr int D_SMEM_DOUBLES = D_STAGES * D_STAGE_DOUBLES + NB * S_LD;
static_assert(32 * H_LD <= D_STAGES * D_STAGE_DOUBLES, "the L21 buffer aliases the diag pipeline");
constexpr int PAD_DOUBLES = 256;  // tiles may read (never use) up to R rows past the last column
}  // namespace gpb

// What one block-column launch works on.
struct GpbStep {
    int j0;         // first column of the block column
    int slot0;      // first factor-storage slot of the lane group (third coordinate of the tensor maps)
    int n;          // lanes (= slots) in flight
    int first_row;  // update kernel: first row of the first tile
    int ntile;      // update kernel: row tiles per lane
};

size_t gpb_ld(int N) { return ((size_t)N + 1 + 15) & ~(size_t)15; }
// a whole number of columns per slot: the slot stride of the tensor maps is then a multiple of the column stride
// columns 0 .. N-1: the factor (row N of each: the residual / z); columns N, N+1: sin and cos of the epochs' phases
// (gp_phase), written by the prep kernel and read by every covariance generation of the lane
size_t gpb_slot_doubles(int N) { return gpb_ld(N) * ((size_t)N + 2 + (gpb::PAD_DOUBLES + gpb_ld(N) - 1) / gpb_ld(N)); }

// The two TMA descriptors over the factor storage of a context (see the plumbing comment below).
cudaError_t gpb_make_tensor_maps(GpWorkspace& w, int N)
{
    using namespace gpb;
    typedef CUresult (*encode_fn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult q;
    cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q);
    if (e != cudaSuccess) return e;
    if (!fn || q != cudaDriverEntryPointSuccess) return cudaErrorNotSupported;
    const cuuint64_t ld = gpb_ld(N);
    const cuuint64_t dims[3] = {ld, (cuuint64_t)N, (cuuint64_t)w.slots};
    const cuuint64_t strides[2] = {ld * sizeof(double), (cuuint64_t)w.slot_doubles * sizeof(double)};
    const cuuint32_t estr[3] = {1, 1, 1};
    const cuuint32_t rows[2] = {A_LD, B_LD};
    unsigned char* dst[2] = {w.tmap_a, w.tmap_b};
    for (int i = 0; i < 2; i++) {
        alignas(64) CUtensorMap tm;
        const cuuint32_t box[3] = {rows[i], KC, 1};
        CUresult r = reinterpret_cast<encode_fn>(fn)(&tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 3, w.ws, dims, strides, box, estr,
                                                      CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                                                      CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r != CUDA_SUCCESS) return cudaErrorInvalidValue;
        static_assert(sizeof(CUtensorMap) == sizeof(w.tmap_a), "CUtensorMap is 128 bytes");
        memcpy(dst[i], &tm, sizeof(tm));
    }
    return cudaSuccess;
}
size_t gpb_lane_bytes() { return sizeof(GpLane); }

// ---- TMA + mbarrier plumbing (sm_100a: SASS UTMALDG / UBLKCP / SYNCS) ------------------------
// Panels of L travel from HBM to shared memory as boxes of two tensor maps over the whole factor
// storage, seen as a (row, column, slot) tensor of doubles with strides (8, 8 ld, 8 slot_doubles)
// bytes: box A = A_LD rows x KC columns, box B = B_LD rows x KC columns.  The box heights ARE the
// padded leading dimensions of the shared-memory stages (132 and 68 = 4 mod 16: conflict-free DMMA
// fragment reads), so a dense box lands in the padded layout; the 4 extra rows are read and
// ignored, rows past the tensor come back as zeros.  One elected thread issues, consumers wait on
// the stage's "full" mbarrier and release it through its "empty" mbarrier: no CTA-wide barrier
// inside the K loop and no per-thread copy instructions.
__device__ __forceinline__ void tma_load_3d(void* dst, const CUtensorMap* tm, uint64_t* bar, int row, int col, int slot)
{
    asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
                 ::"r"(smem_u32(dst)), "l"(reinterpret_cast<uint64_t>(tm)), "r"(smem_u32(bar)), "r"(row), "r"(col), "r"(slot)
                 : "memory");
}
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
    {
        const GpConsts g = gp_consts(sc);
        for (int i = tid; i < N; i += CTA_THREADS) {
            double s, c;
            gp_phase(g, __ldg(p.t + i), s, c);
            W[(size_t)N * ld + i] = s;
            W[(size_t)(N + 1) * ld + i] = c;
        }
    }
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
__device__ __forceinline__ void diag_kloop(const CUtensorMap* tmB, int slot_g, uint64_t* full, uint64_t* empty,
                                           int j0, double* pipe, double* Ss, int tid, int lane)
{
    using namespace gpb;
    constexpr int F1 = 7 - F0;
    double acc0[F0 + 1][2], acc1[F1 + 1][2];
#pragma unroll
    for (int c = 0; c <= F0; c++) acc0[c][0] = acc0[c][1] = 0.0;
#pragma unroll
    for (int c = 0; c <= F1; c++) acc1[c][0] = acc1[c][1] = 0.0;
    const int nk = j0 / KC;
    for (int it = 0; it < nk; it++) {
        const int st = it % D_STAGES;
        if (tid == 0) {
            const int it2 = it + D_STAGES - 1;
            if (it2 < nk) {
                const int s2 = it2 % D_STAGES;
                if (it2 >= D_STAGES) mbar_wait(empty + s2, ((it2 / D_STAGES) & 1) ^ 1);
                mbar_expect_tx(full + s2, D_STAGE_DOUBLES * 8);
                tma_load_3d(pipe + s2 * D_STAGE_DOUBLES, tmB, full + s2, j0, it2 * KC, slot_g);
            }
        }
        mbar_wait(full + st, (it / D_STAGES) & 1);
        const double* Bs = pipe + st * D_STAGE_DOUBLES;
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
        __syncwarp();
        if (lane == 0) mbar_arrive(empty + st);
    }
#pragma unroll
    for (int c = 0; c <= F0; c++) {
        const int r = F0 * 8 + (lane >> 2), cc = c * 8 + 2 * (lane & 3);
        Ss[cc * S_LD + r] = acc0[c][0];
        Ss[(cc + 1) * S_LD + r] = acc0[c][1];
    }
#pragma unroll
    for (int c = 0; c <= F1; c++) {
        const int r = F1 * 8 + (lane >> 2), cc = c * 8 + 2 * (lane & 3);
        Ss[cc * S_LD + r] = acc1[c][0];
        Ss[(cc + 1) * S_LD + r] = acc1[c][1];
    }
}

// Factor the diagonal block that starts at column j0 (threads 0..127 of the CTA; `dsm` is
// D_SMEM_DOUBLES of shared memory they own).  Called by the diag kernel for the first block
// column and by the update kernel's first row tile for every later one: that CTA has just
// produced the 64 rows of L the block needs, and folding the (latency-bound) factorisation
// into the update launch lets the other CTA on the SM keep the FP64 pipe busy meanwhile.
__device__ __forceinline__ void diag_block(const KParams& p, double* __restrict__ W, size_t ld,
                                           GpLane& gl, int j0, double* dsm, int tid,
                                           const CUtensorMap* tmB, int slot_g)
{
    using namespace gpb;
    double* pipe = dsm;                                // D_STAGES x [KC][B_LD]
    double* Ss = dsm + D_STAGES * D_STAGE_DOUBLES;     // [NB][S_LD], column-major
    double* T21 = pipe;                                // after the K loop: [32][H_LD], L21 row-major
    __shared__ double dinv_s[NB];
    __shared__ double etab[64];        // 2^(j/64) for exp_tab
    __shared__ double col_t[NB], col_s[NB], col_c[NB];  // epochs and phases of the block's columns
    __shared__ int fail_s;
    const int warp = tid >> 5, lane = tid & 31;
    const int N = p.N;
    const int nbc = min(NB, N - j0);
    __shared__ __align__(8) uint64_t dbars[2 * D_STAGES];
    exp_tab_load(etab, tid);
    if (tid >= 64) {
        const int cc = tid - 64, col = min(j0 + cc, N - 1);
        col_t[cc] = __ldg(p.t + col);
        col_s[cc] = W[(size_t)N * ld + col];
        col_c[cc] = W[(size_t)(N + 1) * ld + col];
    }
    uint64_t* full = dbars;
    uint64_t* empty = dbars + D_STAGES;
    if (tid == 0) {
        fail_s = 0;
        for (int s = 0; s < D_STAGES; s++) {
            mbar_init(full + s, 1);
            mbar_init(empty + s, D_THREADS / 32);
        }
        mbar_fence_init();
    }
    bar_diag();
    if (tid == 0) {
        const int nk = j0 / KC;
        for (int s = 0; s < D_STAGES - 1 && s < nk; s++) {
            mbar_expect_tx(full + s, D_STAGE_DOUBLES * 8);
            tma_load_3d(pipe + s * D_STAGE_DOUBLES, tmB, full + s, j0, s * KC, slot_g);
        }
    }
    switch (warp) {
        case 0: diag_kloop<0>(tmB, slot_g, full, empty, j0, pipe, Ss, tid, lane); break;
        case 1: diag_kloop<1>(tmB, slot_g, full, empty, j0, pipe, Ss, tid, lane); break;
        case 2: diag_kloop<2>(tmB, slot_g, full, empty, j0, pipe, Ss, tid, lane); break;
        default: diag_kloop<3>(tmB, slot_g, full, empty, j0, pipe, Ss, tid, lane); break;
    }
    bar_diag();
    {
        const GpConsts g = gl.c;
        const bool multi = p.F.multi_instrument != 0;
        const int r = tid & (NB - 1);
        const int row = j0 + r;
        const double trow = col_t[min(r, nbc - 1)], srow = col_s[min(r, nbc - 1)], crow = col_c[min(r, nbc - 1)];
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
                    // t[i]-t[j], i<j (:193): i is the column
                    v = qp_entry(g, etab, kadd(col_t[cc], -trow), col_s[cc], col_c[cc], srow, crow);
                }
                v -= Ss[cc * S_LD + r];
            }
            Ss[cc * S_LD + r] = v;
        }
    }
    bar_diag();

    double logdet = 0.0;
#pragma unroll 1
    for (int half = 0; half < 2; half++) {
        const int h0 = 32 * half;
        const int nv = max(0, min(32, nbc - h0));
        if (warp == 0) {
            // L11 / L22 stay in Ss (in place) and go to HBM
            if (!chol32_panels<false>(Ss + h0 * S_LD + h0, S_LD, lane, nv, logdet, dinv_s + h0,
                                      W + (size_t)(j0 + h0) * ld + j0 + h0, ld) && lane == 0)
                fail_s = 1;
        }
        bar_diag();
        if (half == 0) {
            if (warp == 1) {
                // L21 = S21 L11^-T: row 32+lane, forward substitution
                double t[32];
#pragma unroll
                for (int c = 0; c < 32; c++) t[c] = Ss[c * S_LD + 32 + lane];
#pragma unroll
                for (int c = 0; c < 32; c++) {
                    t[c] *= dinv_s[c];
#pragma unroll
                    for (int c2 = c + 1; c2 < 32; c2++) t[c2] = kfma(-t[c], Ss[c * S_LD + c2], t[c2]);  // L11(c2, c)
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
                    Ss[(32 + c2) * S_LD + 32 + lane] -= a0 + a1;
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
                GpLane* __restrict__ lanes, const GpbStep w, const __grid_constant__ CUtensorMap tmB)
{
    extern __shared__ __align__(128) double dsm_k[];
    const int slot = blockIdx.x;
    GpLane& gl = lanes[slot];
    if (gl.fail) return;
    const size_t ld = ((size_t)p.N + 1 + 15) & ~(size_t)15;
    diag_block(p, ws_all + (size_t)slot * slot_doubles, ld, gl, w.j0, dsm_k, threadIdx.x, &tmB, w.slot0 + slot);
}


// ---- update: the panel below the diagonal block ----------------------------------------------
__global__ void __launch_bounds__(CTA_THREADS, 2)
gpb_update_kernel(const KParams p, double* __restrict__ ws_all, size_t slot_doubles,
                  GpLane* __restrict__ lanes, const GpbStep w, const __grid_constant__ CUtensorMap tmA,
                  const __grid_constant__ CUtensorMap tmB)
{
    using namespace gpb;
    extern __shared__ __align__(128) double usm[];
    __shared__ __align__(8) uint64_t ubars[2 * STAGES + 1];
    __shared__ double etab[64];                          // 2^(j/64) for exp_tab
    __shared__ double col_t[NB], col_s[NB], col_c[NB];   // epochs and phases of block column J
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
    // on their own mbarrier after the K loop, hidden behind the covariance generation
    double* LT = usm + NB * A_LD;        // [NB][B_LD], behind the S tile
    double* dinv = LT + NB * B_LD;       // [NB]
    uint64_t* full = ubars;
    uint64_t* empty = ubars + STAGES;
    uint64_t* aux = ubars + 2 * STAGES;
    const int nk = j0 / KC;
    const int slot_g = w.slot0 + slot;
    const bool warp_rows_live = r0 + wr * 32 <= N;
    constexpr uint32_t STAGE_BYTES = STAGE_DOUBLES * 8;
    auto issue_stage = [&](int s, int it) {
        double* As = usm + s * STAGE_DOUBLES;
        mbar_expect_tx(full + s, STAGE_BYTES);
        tma_load_3d(As, &tmA, full + s, r0, it * KC, slot_g);            // A_LD rows x KC columns of the row tile
        tma_load_3d(As + KC * A_LD, &tmB, full + s, j0, it * KC, slot_g);  // B_LD rows x KC columns of block row J
    };
    if (tid == 0) {
        for (int s = 0; s < STAGES; s++) {
            mbar_init(full + s, 1);
            mbar_init(empty + s, CTA_WARPS);
        }
        mbar_init(aux, 1);
        mbar_fence_init();
    }
    exp_tab_load(etab, tid);
    if (tid >= 64 && tid < 64 + NB) {
        const int cc = tid - 64, col = min(j0 + cc, N - 1);  // columns past a partial block: generated and dropped
        col_t[cc] = __ldg(p.t + col);
        col_s[cc] = W[(size_t)N * ld + col];
        col_c[cc] = W[(size_t)(N + 1) * ld + col];
    }
    __syncthreads();
    if (tid == 0)
        for (int s = 0; s < STAGES - 1 && s < nk; s++) issue_stage(s, s);
    for (int it = 0; it < nk; it++) {
        const int st = it % STAGES;
        if (tid == 0) {
            // refill the stage every warp released at the end of iteration it - 1
            const int it2 = it + STAGES - 1;
            if (it2 < nk) {
                const int s2 = it2 % STAGES;
                if (it2 >= STAGES) mbar_wait(empty + s2, ((it2 / STAGES) & 1) ^ 1);
                issue_stage(s2, it2);
            }
        }
        mbar_wait(full + st, (it / STAGES) & 1);
        const double* As = usm + st * STAGE_DOUBLES;
        const double* Bs = As + KC * A_LD;
        // a warp whose 32 rows are all past row N (partial last tile): waits and releases only
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
        __syncwarp();
        if (lane == 0) mbar_arrive(empty + st);
    }
    __syncthreads();  // every warp is done with the stages: the epilogue buffers alias them
    if (tid == 0) {
        mbar_expect_tx(aux, (NB / KC) * KC * B_LD * 8 + NB * 8);
        for (int q = 0; q < NB / KC; q++) tma_load_3d(LT + q * KC * B_LD, &tmB, aux, j0, j0 + q * KC, slot_g);
        bulk_load(dinv, &gl.dinv[(j0 / NB) & 1][0], NB * 8, aux);
    }

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
        const bool gen = row_ok && row < N;
        const double trow = gen ? __ldg(p.t + row) : 0.0;
        const double srow = gen ? W[(size_t)N * ld + row] : 0.0, crow = gen ? W[(size_t)(N + 1) * ld + row] : 0.0;
        // four columns per pass (cl, cl + 2, cl + 4, cl + 6): four independent chains per thread
        constexpr int CS = CTA_THREADS / R;  // 2
        for (int cl = tid >> 7; cl < NB; cl += 4 * CS) {
            double v[4] = {0.0, 0.0, 0.0, 0.0};
            if (row_ok) {
                if (row == N) {
#pragma unroll
                    for (int q = 0; q < 4; q++)
                        if (cl + q * CS < nbc) v[q] = W[(size_t)(j0 + cl + q * CS) * ld + N];
                } else {
                    // t[i]-t[j], i<j (:193): i is the column
#pragma unroll
                    for (int q = 0; q < 4; q++) {
                        const int cc = cl + q * CS;
                        v[q] = qp_entry(g, etab, kadd(col_t[cc], -trow), col_s[cc], col_c[cc], srow, crow);
                    }
                }
#pragma unroll
                for (int q = 0; q < 4; q++) v[q] = cl + q * CS < nbc ? v[q] - SsT[(cl + q * CS) * A_LD + rl] : 0.0;
            }
#pragma unroll
            for (int q = 0; q < 4; q++) SsT[(cl + q * CS) * A_LD + rl] = v[q];
        }
    }
    mbar_wait(aux, 0);  // L_JJ^T and 1/diag have landed (issued before the generation)
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
    if (!w.tma_ok) return cudaErrorNotReady;  // kc_configure builds the tensor maps
    CUtensorMap tmA, tmB;
    memcpy(&tmA, w.tmap_a, sizeof(tmA));
    memcpy(&tmB, w.tmap_b, sizeof(tmB));
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
                st.slot0 = g * gslots;
                st.n = n;
                const int nbc = min(NB, N - st.j0);
                // rows below the diagonal block, up to and including the residual row N
                st.first_row = nbc == NB ? st.j0 + NB : st.j0;
                st.ntile = (N + 1 - st.first_row + R - 1) / R;
                gpb_diag_kernel<<<n, D_THREADS, dsmem, st_g>>>(p, ws_g, w.slot_doubles, lanes_g, st, tmB);
                gpb_update_kernel<<<dim3(st.ntile, n), CTA_THREADS, usmem, st_g>>>(p, ws_g, w.slot_doubles, lanes_g, st, tmA, tmB);
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
