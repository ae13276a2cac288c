// Kernel family 2, small-N path: the whole GP likelihood of one particle inside one CTA,
// the factor resident in shared memory (N <= ~208: CoRoT7-sized data sets).
//
// Replaces, per particle,  calculate_mu() -> calculate_C() -> log_likelihood()
// (GP / standard-kernel branch) of the reference
// (/root/reference/src/RVmodel.cpp:170-216, 286-375, 841-873; Eigen LLT,
// eigen/Eigen/src/Cholesky/LLT.h:313-368).  One launch, one CTA per particle, no HBM
// traffic beyond the particle's parameters and one double out.
//
// Storage: the lower triangle of C / L in packed 32-wide block columns; block column J is
// column-major [32][ld_J] over rows 32J .. N, where row N is the residual r = y - mu (so
// z = L^-1 r falls out of the factorisation as row N of L and r^T C^-1 r = |z|^2).
//
// Warp roles (16 warps in one CTA per SM; 8, 5 or 4 warps in 2, 3 or 4 CTAs per SM when the factor
// is small enough, so that the serial chains of several particles overlap):
//   warp 0      "diag": the 32x32 Cholesky of each diagonal block (chol32_panels: 8-column panels
//               with the rows in lanes, trailing update on the tensor cores) -- a serial chain of
//               N column steps, the critical path
//   warps 1..   "workers", left-looking per block column J:
//       U_J  S[:,J] = C[:,J] - L[:,0:j0] L[J,0:j0]^T   8x32 strips on the FP64 tensor cores
//       T_J  X = S L_JJ^-T                              a row per lane, 8-column panels + DMMA
//     and, whenever they would wait for the diag warp, they generate covariance entries of
//     later block columns from a shared work queue: the generation (the largest consumer of
//     issue slots at this size) hides under the factorisation chain.
#include "kc_model.cuh"
#include "kc_gp_common.cuh"
#ifdef KC_GPS_PROFILE
#include <cstdio>
#endif

namespace kc {

namespace gps {
constexpr int T = 32;
constexpr int WARPS_PER_SM = 16;      // 128 registers per thread: 1 x 16, 2 x 8, 3 x 5 or 4 x 4 warps fill the register file
constexpr int MAXBLK = 8;
constexpr int GW = 4;              // columns per generation item: two qp_kernel2 calls, four independent chains per lane
constexpr int GQ = T / GW;         // generation items per 32x32 tile
constexpr int SLACK = 64;          // doubles past the last block column (strips read, never use, rows > N)
constexpr size_t SMEM_LIMIT = 227 * 1024;

struct Hdr {
    PlanetInv inv[MAX_PLANETS];
    ParticleScalars sc;
    GpConsts g;
    double dinv[T];
    double etab[64];           // 2^(j/64) for exp_tab
    double quad;
    int base[MAXBLK];
    int ld[MAXBLK];
    int gen_end[MAXBLK];
    int gen_next, u_done, d_done, col_done;
    int tile_cnt[MAXBLK + 1];  // block column 0: finished generation items per 32-row tile
};

__host__ __device__ inline int ld_for(int rows) { return ((rows + 11) & ~15) + 4; }  // >= rows, = 4 mod 16
__host__ __device__ inline int npad(int N) { return (N + 15) & ~15; }
__host__ __device__ inline size_t hdr_bytes() { return (sizeof(Hdr) + 15) & ~(size_t)15; }
}  // namespace gps

// Bytes of dynamic shared memory the small path needs for N epochs (0 = does not fit).
size_t gps_smem_bytes(int N)
{
    using namespace gps;
    const int nb = (N + T - 1) / T;
    if (nb > MAXBLK) return 0;
    size_t d = 4 * (size_t)npad(N) + SLACK;  // epochs, noise variances, sin and cos of the phases
    for (int J = 0; J < nb; J++) d += (size_t)T * ld_for(N + 1 - T * J);
    size_t bytes = hdr_bytes() + d * sizeof(double);
    return bytes <= SMEM_LIMIT ? bytes : 0;
}

namespace {

using namespace gps;

// switches of the schedule (each measured on a B200, profiles/gps_variants_r02.txt)
#ifndef KC_GPS_ISOLATE
#define KC_GPS_ISOLATE 0    // 16-warp CTAs: warps 4, 8, 12 (the diag warp's scheduler) stay idle
#endif
#ifndef KC_GPS_LOOKAHEAD
#define KC_GPS_LOOKAHEAD 1  // U_J: the strips of the diagonal block first, then the diag warp starts
#endif
#ifndef KC_GPS_TILEGEN
#define KC_GPS_TILEGEN 0    // block column 0: factorisation / substitution start per generated tile
#endif
template <int WORKERS>
__device__ __forceinline__ void worker_bar() { asm volatile("bar.sync 1, %0;" ::"n"(WORKERS * 32) : "memory"); }
__device__ __forceinline__ void red_release_add(int* p, int v)
{
    asm volatile("red.release.cta.shared.add.s32 [%0], %1;" ::"r"((unsigned)__cvta_generic_to_shared(p)), "r"(v) : "memory");
}
// warp-uniform read of a flag / counter in shared memory
__device__ __forceinline__ int ld_uniform(const int* p, int lane)
{
    int v = 0;
    if (lane == 0) v = ld_acquire_cta(p);
    return __shfl_sync(0xffffffffu, v, 0);
}

// One generation item: rows 32I .. 32I+31 (a row per lane) x GW columns of block column J.
__device__ __noinline__ void gen_item(Hdr& H, const double* tS, const double* nS, double* M,
                                         int N, int item, int nb, int lane)
{
    int J = 0;
    while (item >= H.gen_end[J]) J++;
    const int local = item - (J ? H.gen_end[J - 1] : 0);
    const int I = J + local / GQ, q = local % GQ;
    const int j0 = T * J;
    const int row = T * I + lane;
    const int c0 = j0 + GW * q;
    const bool row_ok = row < N;
    const int NP = npad(N);
    const double* sS = nS + NP;   // sin, cos of the phases (gp_phase), behind the noise variances
    const double* cS = sS + NP;
    const double trow = tS[row_ok ? row : N - 1], srow = sS[row_ok ? row : N - 1], crow = cS[row_ok ? row : N - 1];
    double* dst = M + H.base[J] + (size_t)(GW * q) * H.ld[J] + (row - j0);
    const int ldj = H.ld[J];
    const GpConsts g = H.g;
#pragma unroll
    for (int cc = 0; cc < GW; cc += 2) {
        const int ca = c0 + cc, cb = ca + 1;
        // t[i]-t[j] with i<j (:193): i is the column (smaller index) of the lower triangle
        const int ia = ca < N ? ca : N - 1, ib = cb < N ? cb : N - 1;
        double va = qp_entry(g, H.etab, kadd(tS[ia], -trow), sS[ia], cS[ia], srow, crow);
        double vb = qp_entry(g, H.etab, kadd(tS[ib], -trow), sS[ib], cS[ib], srow, crow);
        if (row_ok) {
            // eta1^2 * exp(-0 - 0) + sig^2 + jit^2 on the diagonal (:199-206)
            if (ca == row) va = kadd(g.eta1sq, nS[row]);
            if (cb == row) vb = kadd(g.eta1sq, nS[row]);
            if (ca <= row && ca < N) dst[(size_t)cc * ldj] = va;
            if (cb <= row && cb < N) dst[(size_t)(cc + 1) * ldj] = vb;
        }
    }
#if KC_GPS_TILEGEN
    if (J == 0) {
        __syncwarp();
        if (lane == 0) red_release_add(&H.tile_cnt[I], 1);
    }
#endif
}

__device__ __forceinline__ int claim(Hdr& H, int lane)
{
    int it = 0;
    if (lane == 0) it = atomicAdd(&H.gen_next, 1);
    return __shfl_sync(0xffffffffu, it, 0);
}

// U_J for one 8-row strip: S -= L[strip, 0:j0] L[J, 0:j0]^T, accumulated in registers over
// all earlier block columns, one read-modify-write of the strip at the end.  8 rows (one
// fragment row) per item: at these sizes the number of items per step, not the operand
// reuse, decides how many of the four tensor pipes are busy.
__device__ __forceinline__ void update_strip(const Hdr& H, double* M, int N, int J, int strip, int lane)
{
    const int j0 = T * J;
    const int r0 = j0 + 8 * strip;
    double acc[4][2];
#pragma unroll
    for (int n = 0; n < 4; n++) acc[n][0] = acc[n][1] = 0.0;
    for (int Jk = 0; Jk < J; Jk++) {
        const int ldk = H.ld[Jk];
        const double* Acol = M + H.base[Jk] + (r0 - T * Jk) + (lane >> 2) + (size_t)(lane & 3) * ldk;
        const double* Bcol = M + H.base[Jk] + (j0 - T * Jk) + (lane >> 2) + (size_t)(lane & 3) * ldk;
#pragma unroll
        for (int k4 = 0; k4 < T; k4 += 4) {
            double bf[4];
            const double af = Acol[(size_t)k4 * ldk];
#pragma unroll
            for (int n = 0; n < 4; n++) bf[n] = Bcol[(size_t)k4 * ldk + 8 * n];
#pragma unroll
            for (int n = 0; n < 4; n++) dmma(acc[n][0], acc[n][1], af, bf[n]);
        }
    }
    double* S = M + H.base[J];
    const int ldj = H.ld[J];
    const int row = r0 + (lane >> 2);
    if (row <= N) {
#pragma unroll
        for (int n = 0; n < 4; n++)
#pragma unroll
            for (int h = 0; h < 2; h++) {
                double* q = S + (size_t)(8 * n + 2 * (lane & 3) + h) * ldj + (row - j0);
                *q -= acc[n][h];
            }
    }
}

// T_J for 32 rows below the diagonal block: X L_JJ^T = S in place, a row per lane, following the
// diag warp panel by panel (8 columns).  Within a panel: plain substitution on the 8 columns of
// the lane's row in registers (L_JJ read as shared-memory broadcasts); the columns to the right
// of the panel are then updated on the tensor cores, X[:, rest] -= X[:, panel] L_JJ[rest, panel]^T,
// fragments read-modify-written in shared memory: 4x fewer issue slots than a 31-wide register
// window per column.
#ifndef KC_GPS_SLEEP_T
#define KC_GPS_SLEEP_T 20
#endif
#ifndef KC_GPS_SLEEP_D
#define KC_GPS_SLEEP_D 40
#endif
#ifndef KC_GPS_SLEEP_F
#define KC_GPS_SLEEP_F 40
#endif
__device__ __forceinline__ void trsm_tile(Hdr& H, double* M, int N, int J, int nbc, int tile, int lane)
{
    const int j0 = T * J;
    const int row0 = nbc + T * tile;  // first row of the tile, relative to j0
    const int row = j0 + row0 + lane;
    const bool ok = row <= N;
    const int ldj = H.ld[J];
    double* S = M + H.base[J];
    double* mine = S + (ok ? row0 + lane : nbc);
    double* tileS = S + row0;
    const int fr = lane >> 2, fk = lane & 3;
    double q = 0.0;
#pragma unroll 1
    for (int c0 = 0; c0 < nbc; c0 += CHOL_PANEL) {
        const int want = j0 + min(c0 + CHOL_PANEL, nbc);
        while (ld_uniform(&H.col_done, lane) < want) __nanosleep(KC_GPS_SLEEP_T);
        double x[CHOL_PANEL];
#pragma unroll
        for (int k = 0; k < CHOL_PANEL; k++) x[k] = mine[(size_t)min(c0 + k, nbc - 1) * ldj];
#pragma unroll
        for (int k = 0; k < CHOL_PANEL; k++) {
            const int c = c0 + k;
            if (c < nbc) {  // warp-uniform
                const double x0 = x[k] * H.dinv[c];
                x[k] = x0;
                q = kfma(x0, x0, q);
                const double* lc = S + (size_t)c * ldj + c0;  // lc[k2] = L_JJ(c0 + k2, c)
#pragma unroll
                for (int k2 = k + 1; k2 < CHOL_PANEL; k2++) x[k2] = kfma(-x0, lc[k2], x[k2]);
            }
        }
#pragma unroll
        for (int k = 0; k < CHOL_PANEL; k++)
            if (ok && c0 + k < nbc) mine[(size_t)(c0 + k) * ldj] = x[k];
        __syncwarp();
        if (c0 + CHOL_PANEL < nbc) {
            // A fragments: X(tile row 8 Ir + fr, panel column 4 h + fk); columns past a partial block are zero
            double af[4][2];
#pragma unroll
            for (int Ir = 0; Ir < 4; Ir++)
#pragma unroll
                for (int h = 0; h < 2; h++) {
                    const int c = c0 + 4 * h + fk;
                    af[Ir][h] = c < nbc ? tileS[(size_t)c * ldj + 8 * Ir + fr] : 0.0;
                }
#pragma unroll 1
            for (int cb = c0 + CHOL_PANEL; cb < nbc; cb += 8) {
                // B fragment: L_JJ(cb + fr, panel column 4 h + fk)
                double bf[2];
#pragma unroll
                for (int h = 0; h < 2; h++) {
                    const int c = c0 + 4 * h + fk;
                    bf[h] = (c < nbc && cb + fr < nbc) ? S[(size_t)c * ldj + cb + fr] : 0.0;
                }
#pragma unroll
                for (int Ir = 0; Ir < 4; Ir++) {
                    double u0 = 0.0, u1 = 0.0;
                    dmma(u0, u1, af[Ir][0], bf[0]);
                    dmma(u0, u1, af[Ir][1], bf[1]);
                    const int col = cb + 2 * fk;
                    const int r = j0 + row0 + 8 * Ir + fr;
                    double* qd = tileS + (size_t)col * ldj + 8 * Ir + fr;
                    if (r <= N) {
                        if (col < nbc) qd[0] -= u0;
                        if (col + 1 < nbc) qd[ldj] -= u1;
                    }
                }
            }
            __syncwarp();
        }
    }
    if (ok && row == N) H.quad += q;  // |z|^2 of this block column; one writer per step
}

// NW warps per CTA (1 diag + NW-1 workers); 16 / NW CTAs share an SM when the factor is small
// enough, so that the serial chains of several particles overlap.
template <int MODE, int NW>
__global__ void __launch_bounds__(NW * 32, WARPS_PER_SM / NW) gps_kernel(const KParams p)
{
    constexpr int THREADS = NW * 32;
    constexpr bool ISO = KC_GPS_ISOLATE && NW == 16;
    constexpr int WORKERS = ISO ? NW - 4 : NW - 1;
    extern __shared__ __align__(16) unsigned char gps_smem[];
    Hdr& H = *reinterpret_cast<Hdr*>(gps_smem);
    const int N = p.N;
    const int NP = npad(N);
    double* tS = reinterpret_cast<double*>(gps_smem + hdr_bytes());
    double* nS = tS + NP;
    double* sS = nS + NP;
    double* cS = sS + NP;
    double* M = cS + NP;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int b = blockIdx.x;
    const int nb = (N + T - 1) / T;

    if (warp == 0) {
        particle_setup(p, b, H.inv, &H.sc, lane);
        if (lane == 0) {
            H.g = gp_consts(H.sc);
            int off = 0, ge = 0;
            for (int J = 0; J < nb; J++) {
                H.base[J] = off;
                H.ld[J] = ld_for(N + 1 - T * J);
                off += T * H.ld[J];
                ge += (nb - J) * GQ;
                H.gen_end[J] = ge;
            }
            H.gen_next = 0; H.u_done = 0; H.d_done = 0; H.col_done = 0;
            H.quad = 0.0;
            for (int I = 0; I <= MAXBLK; I++) H.tile_cnt[I] = 0;
        }
    }
    __syncthreads();
    {
        const bool multi = p.F.multi_instrument != 0;
        for (int i = tid; i < N; i += THREADS) {
            const double ti = __ldg(p.t + i);
            tS[i] = ti;
            const int in = multi ? __ldg(p.obsi + i) - 1 : 0;
            nS[i] = kadd(__ldg(p.sig2 + i), H.sc.jit2[in]);
            gp_phase(H.g, ti, sS[i], cS[i]);
        }
        exp_tab_load(H.etab, tid);
        // residual = y - mu (:845-847) into row N of the block column that owns epoch i
        if (p.F.MA) {
            // with the MA(1) term (:378-386): recurrence in the (still unused) factor storage
            __syncthreads();
            double* rbuf = M;  // nothing of the factor storage is live yet
            ma_residuals<MODE>(p, &H.sc, H.inv, rbuf, rbuf + NP, tid, THREADS);
            double rv[(MAXBLK * T + THREADS - 1) / THREADS];
            int k = 0;
            for (int i = tid; i < N; i += THREADS) rv[k++] = rbuf[i];
            __syncthreads();
            k = 0;
            for (int i = tid; i < N; i += THREADS) {
                const int J = i / T;
                M[H.base[J] + (size_t)(i % T) * H.ld[J] + (N - T * J)] = rv[k++];
            }
        } else
        for (int i0 = tid; i0 < N; i0 += 2 * THREADS) {
            int i1 = i0 + THREADS;
            const bool has1 = i1 < N;
            if (!has1) i1 = i0;
            const double t0 = __ldg(p.t + i0), t1 = __ldg(p.t + i1);
            const int in0 = multi ? __ldg(p.obsi + i0) - 1 : 0;
            const int in1 = multi ? __ldg(p.obsi + i1) - 1 : 0;
            double mu0, mu1;
            model_mu2<MODE>(p, &H.sc, H.inv, i0, i1, t0, t1, in0, in1, mu0, mu1);
            {
                const int J = i0 / T;
                M[H.base[J] + (size_t)(i0 % T) * H.ld[J] + (N - T * J)] = __ldg(p.y + i0) - mu0;
            }
            if (has1) {
                const int J = i1 / T;
                M[H.base[J] + (size_t)(i1 % T) * H.ld[J] + (N - T * J)] = __ldg(p.y + i1) - mu1;
            }
        }
    }
    __syncthreads();

    const int total_items = H.gen_end[nb - 1];
#ifdef KC_GPS_PROFILE
    __shared__ long long prof_t[2][MAXBLK][4];
    const bool prof = (b == 0) && lane == 0 && warp < 2;
    long long tp = clock64();
#define PROF(tag)                                                                          \
    if (prof) { long long tn = clock64(); prof_t[warp][J][tag] = tn - tp; tp = tn; }
#else
#define PROF(tag)
#endif
    double logdet = 0.0;
    bool ok = true;

    if (warp == 0) {
        // ---- diag role ----
        for (int J = 0; J < nb; J++) {
            const int j0 = T * J;
#if KC_GPS_TILEGEN
            if (J == 0) while (ld_uniform(&H.tile_cnt[0], lane) < GQ) __nanosleep(KC_GPS_SLEEP_D);
            else
#endif
            while (ld_uniform(&H.u_done, lane) < J + 1) __nanosleep(KC_GPS_SLEEP_D);
            PROF(0);
            const int nv = min(T, N - j0);
            if (!chol32_panels<true>(M + H.base[J], H.ld[J], lane, nv, logdet, H.dinv, nullptr, 0,
                                     &H.col_done, T * J))
                ok = false;
            __syncwarp();
            if (lane == 0) st_release_cta(&H.d_done, J + 1);
            PROF(1);
        }
    } else if (!(ISO && (warp & 3) == 0)) {
        // ---- worker role ----
        const int wid = ISO ? warp - 1 - (warp >> 2) : warp - 1;
        for (int J = 0; J < nb; J++) {
            const int j0 = T * J;
            const int nbc = min(T, N - j0);
            // every entry of block column J exists before it is updated
            while (ld_uniform(&H.gen_next, lane) < H.gen_end[J]) {
                const int it = claim(H, lane);
                if (it < total_items) gen_item(H, tS, nS, M, N, it, nb, lane);
            }
#if KC_GPS_TILEGEN
            if (J > 0)   // block column 0 is followed tile by tile (tile_cnt), not as a whole
#endif
            worker_bar<WORKERS>();
            PROF(0);
            if (J > 0) {
                const int nstrips = (N + 1 - j0 + 7) / 8;
#if KC_GPS_LOOKAHEAD
                // the four strips of the diagonal block first: the diag warp starts while the
                // rows below are still being updated
                const int ndiag = min(4, nstrips);
                for (int s = wid; s < ndiag; s += WORKERS) update_strip(H, M, N, J, s, lane);
                worker_bar<WORKERS>();
                if (wid == 0 && lane == 0) st_release_cta(&H.u_done, J + 1);
                for (int s = ndiag + wid; s < nstrips; s += WORKERS) update_strip(H, M, N, J, s, lane);
                worker_bar<WORKERS>();
#else
                for (int s = wid; s < nstrips; s += WORKERS) update_strip(H, M, N, J, s, lane);
                worker_bar<WORKERS>();
#endif
            }
            PROF(1);
            if (wid == 0 && lane == 0) st_release_cta(&H.u_done, J + 1);
            // T_J follows the diag warp column by column (one 32-row tile per warp); the
            // other workers generate ahead until the block is factored
            const int first = j0 + nbc;
            const int ntiles = (N + 1 - first + T - 1) / T;
            if (wid < ntiles) {
                for (int tl = wid; tl < ntiles; tl += WORKERS) {
#if KC_GPS_TILEGEN
                    // every item of block column 0 has been claimed by now: the tile completes without us
                    if (J == 0 && tl + 1 < nb)
                        while (ld_uniform(&H.tile_cnt[tl + 1], lane) < GQ) __nanosleep(KC_GPS_SLEEP_T);
#endif
                    trsm_tile(H, M, N, J, nbc, tl, lane);
                }
            } else {
                while (ld_uniform(&H.d_done, lane) < J + 1) {
                    if (ld_uniform(&H.gen_next, lane) < total_items) {
                        const int it = claim(H, lane);
                        if (it < total_items) gen_item(H, tS, nS, M, N, it, nb, lane);
                    } else {
                        __nanosleep(KC_GPS_SLEEP_F);
                    }
                }
            }
            PROF(2);
            worker_bar<WORKERS>();
            PROF(3);
        }
    }
    __syncthreads();
#ifdef KC_GPS_PROFILE
    if (b == 0 && tid == 0)
        for (int J = 0; J < nb; J++)
            printf("J%d diag: wait_u %lld chol %lld | worker: gen_pre %lld U %lld fill %lld T %lld\n", J,
                   prof_t[0][J][0], prof_t[0][J][1], prof_t[1][J][0], prof_t[1][J][1], prof_t[1][J][2], prof_t[1][J][3]);
#endif
    if (tid == 0) {
        const double pinf = __longlong_as_double(0x7ff0000000000000ll);
        // -0.5 N log(2 pi) - 0.5 * (2 sum log L_ii) - 0.5 r^T C^-1 r   (:858-870)
        double logL = -0.5 * N * 1.8378770664093453 - logdet - 0.5 * H.quad;
        if (!ok || isnan(logL) || isinf(logL)) logL = pinf;  // (:938-941)
        p.logL[b] = logL;
    }
}

}  // namespace

template <int MODE, int NW>
static cudaError_t launch_one(const KParams& p, size_t smem, cudaStream_t stream)
{
    cudaError_t e = cudaFuncSetAttribute(gps_kernel<MODE, NW>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    gps_kernel<MODE, NW><<<p.B, NW * 32, smem, stream>>>(p);
    return cudaGetLastError();
}

cudaError_t launch_gp_small(const KParams& p, int kepler_mode, cudaStream_t stream, int* n_launches)
{
    if (p.B <= 0) return cudaSuccess;
    const size_t smem = gps_smem_bytes(p.N);
    if (!smem) return cudaErrorInvalidValue;
    // CTAs per SM by shared-memory footprint (228 KB per SM, 1 KB reserved per CTA)
    const size_t per_sm = 228 * 1024;
    const int fit = (int)(per_sm / (smem + 1024));
    cudaError_t e;
    if (fit >= 4) e = kepler_mode == 0 ? launch_one<0, 4>(p, smem, stream) : launch_one<1, 4>(p, smem, stream);
    else if (fit == 3) e = kepler_mode == 0 ? launch_one<0, 5>(p, smem, stream) : launch_one<1, 5>(p, smem, stream);
    else if (fit == 2) e = kepler_mode == 0 ? launch_one<0, 8>(p, smem, stream) : launch_one<1, 8>(p, smem, stream);
    else e = kepler_mode == 0 ? launch_one<0, 16>(p, smem, stream) : launch_one<1, 16>(p, smem, stream);
    *n_launches += 1;
    return e;
}

}  // namespace kc
