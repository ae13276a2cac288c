// Kernel family 2, small-N path: the whole GP likelihood of one particle inside one CTA,
// the factor resident in shared memory (N <= 211: CoRoT7-sized data sets).
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
//   warp 0      "diag": the planet invariants, then the 32x32 Cholesky of each diagonal block
//               (chol32_panels: 8-column panels with the rows in lanes, trailing update on the
//               tensor cores) -- a serial chain of N column steps, the critical path
//   warps 1..   "workers", left-looking per block column J:
//       U_J  S[:,J] = C[:,J] - L[:,0:j0] L[J,0:j0]^T   8x32 strips on the FP64 tensor cores; the
//            strips of the diagonal tile first, as soon as that tile exists and T_{J-1} has
//            finished the rows of block J (flags, not barriers)
//       T_J  X = S L_JJ^-T                              a row per lane, 8-column panels + DMMA,
//            panel by panel behind the diag warp
//     and, whenever they would wait, items of a work queue: covariance entries of later block
//     columns (gen_item; the generator of kc_gp_common.cuh) and, behind block column 0, the
//     residuals (mean_item).
// Around it: gps_order_kernel hands the lanes of a batch to the SMs in order of planet count;
// gps_classify_kernel / gps_hit_kernel / gps_kernel<.., CACHE = true> are the factor cache
// (kima_cuda.h: kc_gp_cache_enable).  DESIGN.md sections 3.2.1 and 3.2.4 have the measurements.
#include "kc_model.cuh"
#include "kc_gp_common.cuh"
#include <cstdio>

namespace kc {

namespace gps {
constexpr int T = 32;
constexpr int WARPS_PER_SM = 16;      // 128 registers per thread: 1 x 16, 2 x 8, 3 x 5 or 4 x 4 warps fill the register file
constexpr int MAXBLK = 8;
// Columns per generation item = independent qp_entry chains per lane: 8 where a particle has 7 or more
// workers (3-7 % faster at N = 100 .. 208 than 4), 4 in the 4- and 5-warp CTAs of the smallest sizes, where an
// 8-column item keeps one of three workers away from the factorisation for too long
// (profiles/gps_variants_r02.txt).
__host__ __device__ constexpr int gw_for(int nw) { return nw >= 8 ? 8 : 4; }
constexpr int SLACK = 64;          // doubles past the last block column (strips read, never use, rows > N)
constexpr size_t SMEM_LIMIT = 227 * 1024;

struct Hdr {
    PlanetInv inv[MAX_PLANETS];
    ParticleScalars sc;
    GpConsts g;
    alignas(16) double dinv[MAXBLK][T];    // 1 / diag(L_JJ) of every block (T_J may still read while block J+1 is factored; the factor cache keeps them)
    double etab[64];           // 2^(j/64) for exp_tab
    double quad;
    int base[MAXBLK];
    int ld[MAXBLK];
    int gen_end[MAXBLK];
    int gen_next, u_done, d_done, col_done;
    int gen0;                  // generation items of block column 0; the residual items sit behind them in the queue
    int setup_done, mean_cnt;  // planet invariants ready / finished residual items
    int victim, slot;          // factor cache: entry of the slot this particle's factor goes to (-1: not kept)
    int t0_done;               // J + 1 once T_J has finished the tile below the diagonal block (the rows of block J + 1)
    int pre_cnt[MAXBLK];       // strips of the diagonal tile of block column J updated one block column behind
    int tile_cnt[MAXBLK + 1];  // block column 0: finished generation items per 32-row tile
    int dcnt[MAXBLK];          // block column J > 0: finished generation items of its diagonal tile
    int ccnt[MAXBLK];          // finished queue items of block column J (its covariance entries; for J = 0 the residuals too)
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

// Doubles of one factor-cache entry: the block columns as they sit in shared memory, then 1/diag of every block.
size_t gps_cache_img_doubles(int N)
{
    using namespace gps;
    const int nb = (N + T - 1) / T;
    size_t d = (size_t)nb * T;
    for (int J = 0; J < nb; J++) d += (size_t)T * ld_for(N + 1 - T * J);
    return d;
}

namespace {

using namespace gps;

// switches of the schedule (each measured on a B200, profiles/gps_variants_r02.txt)
#ifndef KC_GPS_ISOLATE
#define KC_GPS_ISOLATE 0    // 16-warp CTAs: warps 4, 8, 12 (the diag warp's scheduler) stay idle
#endif
#ifndef KC_GPS_ENDBAR
#define KC_GPS_ENDBAR 0     // CTAs of fewer than 16 warps: all workers meet at the end of every step
#endif
#ifndef KC_GPS_ROLLED
#define KC_GPS_ROLLED(nw) ((nw) == 8 || (nw) == 5)
#endif
#ifndef KC_GPS_GENONLY
#define KC_GPS_GENONLY 0    // 16-warp CTAs: warps 4, 8, 12 only take queue items (covariance entries, residuals): no tensor-core work beside the diag warp
#endif
#ifndef KC_GPS_SETUP_BY_DIAG
#define KC_GPS_SETUP_BY_DIAG 1  // the planet invariants by the diag warp (idle until the first tile exists), not by a worker
#endif
#ifndef KC_GPS_SLEEP_T
#define KC_GPS_SLEEP_T 20
#endif
#ifndef KC_GPS_SLEEP_D
#define KC_GPS_SLEEP_D 40
#endif
#ifndef KC_GPS_SLEEP_F
#define KC_GPS_SLEEP_F 40
#endif
#ifdef KC_GPS_PROFILE
__device__ long long g_prof_twait[1024], g_prof_tsub[1024], g_prof_tdmma[1024];
__device__ unsigned long long g_prof_cta[3 * 32768];  // per lane: globaltimer at entry and exit (ns), SM id
__device__ __forceinline__ unsigned long long prof_globaltimer()
{
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}   // per CTA (mod 1024): cycles warp 1 waited for col_done inside trsm_tile
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
template <int GW>
__device__ __noinline__ void gen_item(Hdr& H, const double* tS, const double* nS, double* M,
                                         int N, int item, int nb, int lane)
{
    constexpr int GQ = T / GW;  // generation items per 32x32 tile
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
    if (I > J) {
        // a tile below the diagonal: every column is an epoch before every row, nothing to select --
        // GW independent chains per lane, one predicate on the stores (rows past N - 1: row N is the residual)
        double v[GW];
#pragma unroll
        for (int k = 0; k < GW; k++)
            v[k] = qp_entry(g, H.etab, kadd(tS[c0 + k], -trow), sS[c0 + k], cS[c0 + k], srow, crow);
        if (row_ok) {
#pragma unroll
            for (int k = 0; k < GW; k++) dst[(size_t)k * ldj] = v[k];
        }
    } else
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
    __syncwarp();
    if (lane == 0) {
        // block column 0 is consumed tile by tile (the diag warp and T_0 start as soon as their tile exists);
        // U_J's look-ahead strips start as soon as the diagonal tile of block column J exists
        if (J == 0) red_release_add(&H.tile_cnt[I], 1);
        else if (I == J) red_release_add(&H.dcnt[J], 1);
        red_release_add(&H.ccnt[J], 1);
    }
}

// y - mu at epoch i (:845-847): ONE compiled body for the full path's queue items and for the hit kernel
// (inlined copies may contract mu += K * (...) differently, and a hit has to return the bits of a miss)
template <int MODE>
__device__ __noinline__ double residual_at(const KParams& p, const ParticleScalars* sc, const PlanetInv* inv, int i)
{
    const double ti = __ldg(p.t + i);
    const int in = p.F.multi_instrument ? __ldg(p.obsi + i) - 1 : 0;
    double mu = model_background(p, sc, i, ti, in);
    const int npl = sc->npl;
    for (int j = 0; j < npl; j++) mu += kepler_term<MODE>(inv[j], ti);
    return __ldg(p.y + i) - mu;
}
__device__ __noinline__ void particle_setup_shared(const KParams& p, int b, PlanetInv* inv, ParticleScalars* sc, int lane)
{
    particle_setup(p, b, inv, sc, lane);
}

// One residual item: y - mu (:845-847) at 32 epochs, an epoch per lane, into row N of the block columns
// that own them.  Needs the planet invariants (setup_done); same summation order as model_mu2.
template <int MODE>
__device__ __noinline__ void mean_item(const KParams& p, Hdr& H, double* M, int N, int m, int lane)
{
    while (ld_uniform(&H.setup_done, lane) == 0) __nanosleep(KC_GPS_SLEEP_F);
    const int i = T * m + lane;
    if (i < N) {
        const int J = i / T;
        M[H.base[J] + (size_t)(i % T) * H.ld[J] + (N - T * J)] = residual_at<MODE>(p, &H.sc, H.inv, i);
    }
    __syncwarp();
    if (lane == 0) {
        red_release_add(&H.mean_cnt, 1);
        red_release_add(&H.ccnt[0], 1);
    }
}

// Whatever the queue holds at position `it`: covariance entries, or (behind block column 0) residuals.
template <int MODE, int GW>
__device__ __forceinline__ void work_item(const KParams& p, Hdr& H, const double* tS, const double* nS, double* M,
                                          int N, int it, int nb, int lane)
{
    if (it >= H.gen0 && it < H.gen_end[0]) mean_item<MODE>(p, H, M, N, it - H.gen0, lane);
    else gen_item<GW>(H, tS, nS, M, N, it, nb, lane);
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
__device__ __forceinline__ void update_strip(const Hdr& H, double* M, int N, int J, int strip, int lane, int k_begin,
                                             int k_end)
{
    const int j0 = T * J;
    const int r0 = j0 + 8 * strip;
    double acc[4][2];
#pragma unroll
    for (int n = 0; n < 4; n++) acc[n][0] = acc[n][1] = 0.0;
    for (int Jk = k_begin; Jk < k_end; Jk++) {
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

// The same for a full block (32 columns, every block but a partial last one): no column bounds, so the
// panel is straight-line code -- 1/diag, the lane's 8 entries and the 28 entries of the panel's triangle are
// all in flight before the substitution chain starts (8 x (multiply, fused multiply-add) = ~140 cycles, where
// the bounds-checked version above spends ~1200 per panel waiting for its loads one column at a time), and the
// trailing update's blocks are unrolled.  ROLLED keeps the panel loop rolled (a quarter of the code).
template <bool ROLLED>
__device__ __forceinline__ void trsm_tile_full(Hdr& H, double* M, int N, int J, int tile, int lane)
{
    const int j0 = T * J;
    const int row0 = T + T * tile;  // first row of the tile, relative to j0
    const int row = j0 + row0 + lane;
    const bool ok = row <= N;
    const int ldj = H.ld[J];
    double* S = M + H.base[J];
    double* mine = S + (ok ? row0 + lane : T);
    double* tileS = S + row0;
    const int fr = lane >> 2, fk = lane & 3;
    const double* dinv = H.dinv[J];
    // rows past N of the last tile: computed on a clamped address, never stored
    bool rok[4];
#pragma unroll
    for (int Ir = 0; Ir < 4; Ir++) rok[Ir] = j0 + row0 + 8 * Ir + fr <= N;
    double q = 0.0;
    auto panel = [&](const int p) {
        const int c0 = CHOL_PANEL * p;
        while (ld_uniform(&H.col_done, lane) < j0 + c0 + CHOL_PANEL) __nanosleep(KC_GPS_SLEEP_T);
        double x[CHOL_PANEL], dv[CHOL_PANEL], l[CHOL_PANEL][CHOL_PANEL];
#pragma unroll
        for (int k = 0; k < CHOL_PANEL; k++) {
            x[k] = mine[(size_t)(c0 + k) * ldj];
            dv[k] = dinv[c0 + k];
        }
#pragma unroll
        for (int k = 0; k < CHOL_PANEL; k++)
#pragma unroll
            for (int k2 = k + 1; k2 < CHOL_PANEL; k2++) l[k][k2] = S[(size_t)(c0 + k) * ldj + c0 + k2];  // L_JJ(c0 + k2, c0 + k)
#pragma unroll
        for (int k = 0; k < CHOL_PANEL; k++) {
            const double x0 = x[k] * dv[k];
            x[k] = x0;
            q = kfma(x0, x0, q);
#pragma unroll
            for (int k2 = k + 1; k2 < CHOL_PANEL; k2++) x[k2] = kfma(-x0, l[k][k2], x[k2]);
        }
        if (ok) {
#pragma unroll
            for (int k = 0; k < CHOL_PANEL; k++) mine[(size_t)(c0 + k) * ldj] = x[k];
        }
        __syncwarp();
        if (p < T / CHOL_PANEL - 1) {
            // A fragments: X(tile row 8 Ir + fr, panel column 4 h + fk)
            double af[4][2];
#pragma unroll
            for (int Ir = 0; Ir < 4; Ir++)
#pragma unroll
                for (int h = 0; h < 2; h++) af[Ir][h] = tileS[(size_t)(c0 + 4 * h + fk) * ldj + 8 * Ir + fr];
#pragma unroll
            for (int cbi = p + 1; cbi < T / 8; cbi++) {
                const int cb = 8 * cbi;
                // B fragment: L_JJ(cb + fr, panel column 4 h + fk)
                const double b0 = S[(size_t)(c0 + fk) * ldj + cb + fr], b1 = S[(size_t)(c0 + 4 + fk) * ldj + cb + fr];
#pragma unroll
                for (int Ir = 0; Ir < 4; Ir++) {
                    double u0 = 0.0, u1 = 0.0;
                    dmma(u0, u1, af[Ir][0], b0);
                    dmma(u0, u1, af[Ir][1], b1);
                    double* qd = tileS + (size_t)(cb + 2 * fk) * ldj + 8 * Ir + fr;
                    if (rok[Ir]) {
                        qd[0] -= u0;
                        qd[ldj] -= u1;
                    }
                }
            }
            __syncwarp();
        }
    };
    if constexpr (ROLLED) {
#pragma unroll 1
        for (int p = 0; p < T / CHOL_PANEL; p++) panel(p);
    } else {
#pragma unroll
        for (int p = 0; p < T / CHOL_PANEL; p++) panel(p);
    }
    if (row == N) H.quad += q;  // |z|^2 of this block column; one writer per step
}

// T_J for 32 rows below the diagonal block: X L_JJ^T = S in place, a row per lane, following the
// diag warp panel by panel (8 columns).  Within a panel: plain substitution on the 8 columns of
// the lane's row in registers (L_JJ read as shared-memory broadcasts); the columns to the right
// of the panel are then updated on the tensor cores, X[:, rest] -= X[:, panel] L_JJ[rest, panel]^T,
// fragments read-modify-written in shared memory: 4x fewer issue slots than a 31-wide register
// window per column.
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
#ifdef KC_GPS_PROFILE
        long long tw0 = clock64();
#endif
        while (ld_uniform(&H.col_done, lane) < want) __nanosleep(KC_GPS_SLEEP_T);
#ifdef KC_GPS_PROFILE
        long long tw1 = clock64();
        if (threadIdx.x == 32) g_prof_twait[blockIdx.x & 1023] += tw1 - tw0;
#endif
        double x[CHOL_PANEL];
#pragma unroll
        for (int k = 0; k < CHOL_PANEL; k++) x[k] = mine[(size_t)min(c0 + k, nbc - 1) * ldj];
#pragma unroll
        for (int k = 0; k < CHOL_PANEL; k++) {
            const int c = c0 + k;
            if (c < nbc) {  // warp-uniform
                const double x0 = x[k] * H.dinv[J][c];
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
#ifdef KC_GPS_PROFILE
        long long tw2 = clock64();
        if (threadIdx.x == 32) g_prof_tsub[blockIdx.x & 1023] += tw2 - tw1;
#endif
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
#ifdef KC_GPS_PROFILE
        if (threadIdx.x == 32) g_prof_tdmma[blockIdx.x & 1023] += clock64() - tw2;
#endif
    }
    if (row == N) H.quad += q;  // |z|^2 of this block column; one writer per step
}

// ---- factor cache ---------------------------------------------------------------------------------
// Most proposals of a nested-sampling run leave the covariance matrix alone (they move planets, offsets,
// the trend: 62 % of RVmodel::perturb's draws in GP mode, src/RVmodel.cpp:448-544), and C -- hence L --
// depends on eta1..eta4 and the jitters only.  With a cache attached (kc_gp_cache_enable) every lane names a
// slot, the particle it carries; the slot keeps the factors of the last two covariance matrices seen there
// (the particle's state and its last rejected or not-yet-superseded proposal), keyed by those parameters.
//   gps_classify_kernel  one thread per lane: key compare, hit and miss lists, the entry a miss will fill
//   gps_hit_kernel       one WARP per hit: the factor stays in HBM / L2 (161 KB at N = 177, read once);
//                        the warp recomputes the residuals and takes them through exactly the arithmetic
//                        the full path applies to the residual row (same DMMA sequences with the same
//                        operands, same FMA chains), so a hit returns the log-likelihood of a miss bit for
//                        bit and chains do not depend on the cache.  32 hits in flight per SM.
//   gps_kernel           the misses, as without a cache; the factor image then goes to the slot by bulk copy.
__device__ __forceinline__ double cache_key(const KParams& p, int b, int i)
{
    const kc_layout& L = p.L;
    if (i < 4) return ldp(p, L.row_eta + i, b);
    return i - 4 < L.n_jit ? ldp(p, L.row_jit + i - 4, b) : 0.0;
}

__global__ void gps_classify_kernel(const KParams p, const GpsCache gc)
{
    constexpr int KEYLEN = 4 + KC_MAX_INST;
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= p.B) return;
    const int slot = gc.slots[b];
    int cl = -1;  // not cached
    if (slot >= 0) {
        GpsCacheMeta* m = gc.meta + (size_t)slot * 2;
        bool eq0 = m[0].valid != 0, eq1 = m[1].valid != 0;
        for (int i = 0; i < KEYLEN; i++) {
            const long long k = __double_as_longlong(cache_key(p, b, i));
            eq0 = eq0 && __double_as_longlong(m[0].key[i]) == k;
            eq1 = eq1 && __double_as_longlong(m[1].key[i]) == k;
        }
        if ((eq0 || eq1) && gc.debug != 1) {
            cl = eq0 ? 0 : 1;
            m[cl].stamp = gc.stamp;
        } else {
            // the entry to fill: an empty one, else the one used longer ago
            int victim = !m[0].valid ? 0 : !m[1].valid ? 1 : (m[0].stamp - m[1].stamp <= 0 ? 0 : 1);
            if (eq0 || eq1) victim = eq0 ? 0 : 1;  // (debug) refill the entry that matched
            cl = -2 - victim;
        }
    }
    gc.cls[b] = cl;
    if (cl >= 0) gc.hit_list[atomicAdd(gc.cnt + 0, 1)] = b;
    else gc.miss_list[atomicAdd(gc.cnt + 1, 1)] = b;
}

namespace hit {
constexpr int WARPS = 8;
struct WarpState {
    PlanetInv inv[MAX_PLANETS];
    ParticleScalars sc;
    double z[MAXBLK * T];   // the residual row: z[32 J + c] = entry (N, 32 J + c)
};
}  // namespace hit

template <int MODE>
__global__ void __launch_bounds__(hit::WARPS * 32, 4) gps_hit_kernel(const KParams p, const GpsCache gc, const int pre)
{
    extern __shared__ __align__(16) unsigned char hit_smem[];
    __shared__ int base[MAXBLK], ld[MAXBLK];
    const int N = p.N;
    const int nb = (N + T - 1) / T;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int fr = lane >> 2, fk = lane & 3;
    if (tid == 0) {
        int off = 0;
        for (int J = 0; J < nb; J++) {
            base[J] = off;
            ld[J] = ld_for(N + 1 - T * J);
            off += T * ld[J];
        }
    }
    __syncthreads();
    const int item = blockIdx.x * hit::WARPS + warp;
    if (item >= gc.cnt[0]) return;
    const int b = gc.hit_list[item];
    const int slot = gc.slots[b], entry = gc.cls[b];
    hit::WarpState& W = reinterpret_cast<hit::WarpState*>(hit_smem)[warp];
    const double* __restrict__ img = gc.img + ((size_t)slot * 2 + entry) * gc.img_doubles;
    const size_t d_tiles = gc.img_doubles - (size_t)nb * T;
    const double* __restrict__ dinv = img + d_tiles;
    // The image is read once, in a known order, 32x32 blocks at a time.  Thousands of warps are in flight on
    // the chip (L2 holds ~26 KB for each), so a warp asks L2 for one block ahead of its use, no more: lane c
    // the 256 bytes of column c.
    auto prefetch_block = [&](const double* blk, const int ldk) {
        const double* q = blk + (size_t)lane * ldk;
        asm volatile("prefetch.global.L2 [%0];" ::"l"(q));
        asm volatile("prefetch.global.L2 [%0];" ::"l"(q + 16));
        asm volatile("prefetch.global.L2 [%0];" ::"l"(q + 31));
    };
    prefetch_block(img + base[0], ld[0]);  // the first diagonal block, behind the residuals

    particle_setup_shared(p, b, W.inv, &W.sc, lane);
    __syncwarp();
    for (int m = 0; m < nb; m++) {
        const int i = T * m + lane;
        W.z[i] = i < N ? residual_at<MODE>(p, &W.sc, W.inv, i) : 0.0;
    }
    __syncwarp();

    // U_J on the residual row: z_J -= sum_k z_k L(J,k)^T over block columns [kb, ke), the DMMA sequence of
    // update_strip (the A operand carries the row in all eight fragment rows: every row of the output is it)
    auto U = [&](const int J, const int kb, const int ke) {
        const int j0 = T * J;
        double acc[4][2];
#pragma unroll
        for (int n = 0; n < 4; n++) acc[n][0] = acc[n][1] = 0.0;
        for (int Jk = kb; Jk < ke; Jk++) {
            const int ldk = ld[Jk];
            const double* Bp = img + base[Jk] + (j0 - T * Jk) + fr + (size_t)fk * ldk;
            const double* zk = W.z + T * Jk;
            if (Jk + 1 < ke) prefetch_block(img + base[Jk + 1] + (j0 - T * (Jk + 1)), ld[Jk + 1]);
#pragma unroll
            for (int k4 = 0; k4 < T; k4 += 4) {
                double bf[4];
                const double af = zk[k4 + fk];
#pragma unroll
                for (int n = 0; n < 4; n++) bf[n] = __ldcg(Bp + (size_t)k4 * ldk + 8 * n);
#pragma unroll
                for (int n = 0; n < 4; n++) dmma(acc[n][0], acc[n][1], af, bf[n]);
            }
        }
        if (fr == 0) {
#pragma unroll
            for (int n = 0; n < 4; n++)
#pragma unroll
                for (int h = 0; h < 2; h++) W.z[T * J + 8 * n + 2 * fk + h] -= acc[n][h];
        }
        __syncwarp();
    };

    double quad = 0.0;
    for (int J = 0; J < nb; J++) {
        const int j0 = T * J;
        const int nbc = min(T, N - j0);
        if (J >= 1) {
            const int nstrips = (N + 1 - j0 + 7) / 8;
            const int ndiag = min(4, nstrips);
            const int sN = (N - j0) / 8;  // the strip that holds row N
            if (sN < ndiag) {
                // a short last block: the residual row rides with the look-ahead strips of the full path,
                // which 16-warp CTAs update one block column behind
                if (pre && J >= 2) U(J, 0, J - 1);
                U(J, pre ? J - 1 : 0, J);
            } else {
                U(J, 0, J);
            }
        }
        // T_J on the residual row: the arithmetic of trsm_tile_full (full blocks) / trsm_tile (a partial last one)
        const int ldj = ld[J];
        const double* S = img + base[J];
        double* zj = W.z + T * J;
        double q = 0.0;
        // the next step opens with block (J+1, 0) and closes with diagonal block J+1
        if (J + 1 < nb) {
            prefetch_block(img + base[0] + T * (J + 1), ld[0]);
            prefetch_block(img + base[J + 1], ld[J + 1]);
        }
        if (nbc == T) {
#pragma unroll 1
            for (int c0 = 0; c0 < T; c0 += CHOL_PANEL) {
                double x[CHOL_PANEL], dv[CHOL_PANEL], l[CHOL_PANEL][CHOL_PANEL];
#pragma unroll
                for (int k = 0; k < CHOL_PANEL; k++) {
                    x[k] = zj[c0 + k];
                    dv[k] = __ldcg(dinv + T * J + c0 + k);
                }
#pragma unroll
                for (int k = 0; k < CHOL_PANEL; k++)
#pragma unroll
                    for (int k2 = k + 1; k2 < CHOL_PANEL; k2++) l[k][k2] = __ldcg(S + (size_t)(c0 + k) * ldj + c0 + k2);
#pragma unroll
                for (int k = 0; k < CHOL_PANEL; k++) {
                    const double x0 = x[k] * dv[k];
                    x[k] = x0;
                    q = kfma(x0, x0, q);
#pragma unroll
                    for (int k2 = k + 1; k2 < CHOL_PANEL; k2++) x[k2] = kfma(-x0, l[k][k2], x[k2]);
                }
                __syncwarp();
                if (lane == 0) {
#pragma unroll
                    for (int k = 0; k < CHOL_PANEL; k++) zj[c0 + k] = x[k];
                }
                __syncwarp();
                if (c0 + CHOL_PANEL < T) {
                    const double af0 = zj[c0 + fk], af1 = zj[c0 + 4 + fk];
                    for (int cb = c0 + CHOL_PANEL; cb < T; cb += 8) {
                        const double b0 = __ldcg(S + (size_t)(c0 + fk) * ldj + cb + fr);
                        const double b1 = __ldcg(S + (size_t)(c0 + 4 + fk) * ldj + cb + fr);
                        double u0 = 0.0, u1 = 0.0;
                        dmma(u0, u1, af0, b0);
                        dmma(u0, u1, af1, b1);
                        if (fr == 0) {
                            zj[cb + 2 * fk] -= u0;
                            zj[cb + 2 * fk + 1] -= u1;
                        }
                    }
                    __syncwarp();
                }
            }
        } else {
#pragma unroll 1
            for (int c0 = 0; c0 < nbc; c0 += CHOL_PANEL) {
                double x[CHOL_PANEL];
#pragma unroll
                for (int k = 0; k < CHOL_PANEL; k++) x[k] = zj[min(c0 + k, nbc - 1)];
#pragma unroll
                for (int k = 0; k < CHOL_PANEL; k++) {
                    const int c = c0 + k;
                    if (c < nbc) {
                        const double x0 = x[k] * __ldcg(dinv + T * J + c);
                        x[k] = x0;
                        q = kfma(x0, x0, q);
                        const double* lc = S + (size_t)c * ldj + c0;
#pragma unroll
                        for (int k2 = k + 1; k2 < CHOL_PANEL; k2++) x[k2] = kfma(-x0, __ldcg(lc + k2), x[k2]);
                    }
                }
                __syncwarp();
                if (lane == 0) {
#pragma unroll
                    for (int k = 0; k < CHOL_PANEL; k++)
                        if (c0 + k < nbc) zj[c0 + k] = x[k];
                }
                __syncwarp();
                if (c0 + CHOL_PANEL < nbc) {
                    double af[2];
#pragma unroll
                    for (int h = 0; h < 2; h++) {
                        const int c = c0 + 4 * h + fk;
                        af[h] = c < nbc ? zj[c] : 0.0;
                    }
                    for (int cb = c0 + CHOL_PANEL; cb < nbc; cb += 8) {
                        double bf[2];
#pragma unroll
                        for (int h = 0; h < 2; h++) {
                            const int c = c0 + 4 * h + fk;
                            bf[h] = (c < nbc && cb + fr < nbc) ? __ldcg(S + (size_t)c * ldj + cb + fr) : 0.0;
                        }
                        double u0 = 0.0, u1 = 0.0;
                        dmma(u0, u1, af[0], bf[0]);
                        dmma(u0, u1, af[1], bf[1]);
                        const int col = cb + 2 * fk;
                        if (fr == 0) {
                            if (col < nbc) zj[col] -= u0;
                            if (col + 1 < nbc) zj[col + 1] -= u1;
                        }
                    }
                    __syncwarp();
                }
            }
        }
        quad += q;  // |z|^2 block column by block column, the order of the full path
    }
    if (lane == 0) {
        const double pinf = __longlong_as_double(0x7ff0000000000000ll);
        // -0.5 N log(2 pi) - 0.5 * (2 sum log L_ii) - 0.5 r^T C^-1 r   (:858-870)
        double logL = -0.5 * N * 1.8378770664093453 - gc.meta[(size_t)slot * 2 + entry].logdet - 0.5 * quad;
        if (isnan(logL) || isinf(logL)) logL = pinf;  // (:938-941)
        p.logL[b] = logL;
    }
}

// NW warps per CTA (1 diag + NW-1 workers); 16 / NW CTAs share an SM when the factor is small
// enough, so that the serial chains of several particles overlap.
template <int MODE, int NW, bool CACHE>
__global__ void __launch_bounds__(NW * 32, WARPS_PER_SM / NW) gps_kernel(const KParams p, const GpsCache gc)
{
    constexpr int THREADS = NW * 32;
    constexpr int GW = gw_for(NW), GQ = T / GW;
    // 16-warp CTAs (one particle per SM, N > 128): the next diagonal tile is updated one block column behind,
    // off the critical path (+3-4 % at N = 177, 208; with 7 or fewer workers it costs more than it saves)
    constexpr bool PRE = NW >= 16;
    // Two 8-warp or three 5-warp CTAs per SM run measurably faster with the panel loops of the factorisation
    // and of the substitution rolled (N = 100: 1.19 -> 1.10 ms per 8192 lanes, N = 128: 1.37 -> 1.20, N = 96:
    // 0.768 -> 0.731; the unrolled code of several particles in different phases does not fit the instruction
    // cache), one 16-warp or four 4-warp CTAs with them unrolled (N = 177: 2.71 -> 2.62 ms, N = 64: 0.421 ->
    // 0.399; profiles/gps_variants_r02.txt)
    constexpr bool ROLLED = KC_GPS_ROLLED(NW);
    constexpr bool ISO = KC_GPS_ISOLATE && NW == 16;
    constexpr bool GENONLY = KC_GPS_GENONLY && NW == 16 && !ISO;
    constexpr int WORKERS = (ISO || GENONLY) ? NW - 4 : NW - 1;
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
    // One particle per CTA, a CTA per lane of the batch -- or, with a factor cache attached, a grid that fills
    // the SMs once and walks the list of lanes gps_classify_kernel found no factor for (their number is only
    // known on the device).
    int item = blockIdx.x;
    if (CACHE && item >= gc.cnt[1]) return;
    do {  // without a cache: this CTA's lane, once (no back edge)
    const int b = CACHE ? gc.miss_list[item] : (p.order ? p.order[item] : item);
    const int nb = (N + T - 1) / T;
#ifdef KC_GPS_PROFILE
    __shared__ long long prof_c[4];
    if (tid == 0) {
        prof_c[0] = clock64();
        if (b < 32768) {
            unsigned smid;
            asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
            g_prof_cta[3 * b] = prof_globaltimer();
            g_prof_cta[3 * b + 2] = smid;
        }
    }
#endif

    // ---- start-up.  Nothing the factorisation's first steps need waits for the planets: every thread reads
    // the four GP hyperparameters and its epochs' jitter itself (covariance entries need nothing else), the
    // planet invariants are prepared by the last worker warp while the others already generate block column 0,
    // and the residuals y - mu are queue items behind that block column (first needed by the last tile of T_0).
    {
        const kc_layout& L = p.L;
        const GpConsts g = gp_consts(ldp(p, L.row_eta, b), ldp(p, L.row_eta + 1, b), ldp(p, L.row_eta + 2, b),
                                     ldp(p, L.row_eta + 3, b));
        const bool multi = p.F.multi_instrument != 0;
        for (int i = tid; i < N; i += THREADS) {
            const double ti = __ldg(p.t + i);
            tS[i] = ti;
            const int in = multi ? __ldg(p.obsi + i) - 1 : 0;
            double j2 = 0.0;
            if (in < L.n_jit) {
                const double jt = ldp(p, L.row_jit + in, b);
                j2 = kmul(jt, jt);
            }
            nS[i] = kadd(__ldg(p.sig2 + i), j2);
            gp_phase(g, ti, sS[i], cS[i]);
        }
        exp_tab_load(H.etab, tid);
        if (tid == THREADS - 1) {
            H.g = g;
            // residual items: none with the MA(1) term (its recurrence runs before the factorisation, below)
            const int nm = p.F.MA ? 0 : (N + T - 1) / T;
            int off = 0, ge = 0;
            for (int J = 0; J < nb; J++) {
                H.base[J] = off;
                H.ld[J] = ld_for(N + 1 - T * J);
                off += T * H.ld[J];
                ge += (nb - J) * GQ;
                if (J == 0) H.gen0 = ge;
                H.gen_end[J] = ge + nm;
            }
            H.gen_next = 0; H.u_done = 0; H.d_done = 0; H.col_done = 0;
            H.setup_done = 0; H.mean_cnt = 0; H.t0_done = 0;
            for (int I = 0; I < MAXBLK; I++) H.pre_cnt[I] = 0;
            H.quad = 0.0;
            for (int I = 0; I <= MAXBLK; I++) H.tile_cnt[I] = 0;
            for (int I = 0; I < MAXBLK; I++) H.dcnt[I] = H.ccnt[I] = 0;
            H.victim = H.slot = -1;
            if (CACHE) {
                const int cl = gc.cls[b];   // -2 - victim: a cached lane whose factor is kept
                if (cl <= -2) {
                    H.victim = -2 - cl;
                    H.slot = gc.slots[b];
                }
            }
        }
    }
    if (p.F.MA) {
        // residual with the MA(1) term (:378-386): planets first, then the recurrence in the (still unused)
        // factor storage, all threads
        if (warp == 0) particle_setup_shared(p, b, H.inv, &H.sc, lane);
        __syncthreads();
        double* rbuf = M;
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
        if (tid == 0) H.setup_done = 1;
    }
    __syncthreads();
#ifdef KC_GPS_PROFILE
    if (tid == 0) prof_c[1] = clock64();
#endif

    const int total_items = H.gen_end[nb - 1];
#ifdef KC_GPS_PROFILE
    __shared__ long long prof_t[2][MAXBLK][4];
    const bool prof = (b == gridDim.x / 2) && lane == 0 && warp < 2;
    long long tp = clock64();
    if (tid == 0) prof_c[2] = tp;
#define PROF(tag)                                                                          \
    if (prof) { long long tn = clock64(); prof_t[warp][J][tag] = tn - tp; tp = tn; }
    __shared__ long long prof_abs[MAXBLK][6];
#define PABS(k) if (prof) prof_abs[J][k] = clock64() - prof_c[2];
#else
#define PROF(tag)
#define PABS(k)
#endif
    double logdet = 0.0;
    bool ok = true;

    if (warp == 0) {
        // ---- diag role ----
#if KC_GPS_SETUP_BY_DIAG
        if (!p.F.MA) {
            // the planets (needed by the residual items only), while the workers generate the first tile
            particle_setup_shared(p, b, H.inv, &H.sc, lane);
            if (lane == 0) st_release_cta(&H.setup_done, 1);
        }
#endif
        for (int J = 0; J < nb; J++) {
            const int j0 = T * J;
            // the first block as soon as its tile is generated, the others once U_J has updated them
            if (J == 0) while (ld_uniform(&H.tile_cnt[0], lane) < GQ) __nanosleep(KC_GPS_SLEEP_D);
            else while (ld_uniform(&H.u_done, lane) < J + 1) __nanosleep(KC_GPS_SLEEP_D);
            PROF(0);
            PABS(0);
            const int nv = min(T, N - j0);
            bool fine;
            if constexpr (ROLLED)
                fine = chol32_panels_rolled<true>(M + H.base[J], H.ld[J], lane, nv, logdet, H.dinv[J], nullptr, 0,
                                                  &H.col_done, T * J);
            else
                fine = chol32_panels<true>(M + H.base[J], H.ld[J], lane, nv, logdet, H.dinv[J], nullptr, 0,
                                           &H.col_done, T * J);
            if (!fine) ok = false;
            __syncwarp();
            if (lane == 0) st_release_cta(&H.d_done, J + 1);
            PROF(1);
            PABS(1);
        }
    } else if (GENONLY && (warp & 3) == 0) {
        // ---- the diag warp's scheduler mates: queue items only ----
        while (ld_uniform(&H.gen_next, lane) < total_items) {
            const int it = claim(H, lane);
            if (it < total_items) work_item<MODE, GW>(p, H, tS, nS, M, N, it, nb, lane);
        }
    } else if (!(ISO && (warp & 3) == 0)) {
        // ---- worker role ----
        const int wid = (ISO || GENONLY) ? warp - 1 - (warp >> 2) : warp - 1;
        if (!KC_GPS_SETUP_BY_DIAG && warp == NW - 1 && !p.F.MA) {
            // the planets: needed by the residual items only, which sit behind block column 0 in the queue
            particle_setup_shared(p, b, H.inv, &H.sc, lane);
            if (lane == 0) st_release_cta(&H.setup_done, 1);
        }
        for (int J = 0; J < nb; J++) {
            const int j0 = T * J;
            const int nbc = min(T, N - j0);
            if (J == 0) {
                // block column 0 is followed tile by tile (tile_cnt), not as a whole; the residual items are
                // claimed here too
                while (ld_uniform(&H.gen_next, lane) < H.gen_end[0]) {
                    const int it = claim(H, lane);
                    if (it < total_items) work_item<MODE, GW>(p, H, tS, nS, M, N, it, nb, lane);
                }
                PROF(0);
            } else {
                // Only the diagonal tile of block column J stands between T_{J-1} and the next factorisation
                // step: its (up to four) 8-row strips are updated as soon as it is generated, the diag warp
                // starts, and the rest of the block column is generated and updated behind it.
                const int nstrips = (N + 1 - j0 + 7) / 8;
                const int ndiag = min(4, nstrips);
                while (ld_uniform(&H.gen_next, lane) < H.gen_end[J - 1] + GQ) {
                    const int it = claim(H, lane);
                    if (it < total_items) work_item<MODE, GW>(p, H, tS, nS, M, N, it, nb, lane);
                }
                if (wid < ndiag) {
                    // what the strips need: the tile generated, the rows of block J of T_{J-1} (its first tile),
                    // and, one block column behind, the contributions of block columns 0 .. J-2 (below).  The
                    // other tiles of T_{J-1} may still be running: nothing here touches them.
                    while (ld_uniform(&H.dcnt[J], lane) < GQ) __nanosleep(KC_GPS_SLEEP_T);
                    while (ld_uniform(&H.t0_done, lane) < J) __nanosleep(KC_GPS_SLEEP_T);
                    if (PRE && J >= 2)
                        while (ld_uniform(&H.pre_cnt[J], lane) < ndiag) __nanosleep(KC_GPS_SLEEP_T);
                    for (int s = wid; s < ndiag; s += WORKERS) update_strip(H, M, N, J, s, lane, PRE ? J - 1 : 0, J);
                    // the strip warps among themselves (barrier 2), then the diag warp may go
                    asm volatile("bar.sync 2, %0;" ::"r"(min(ndiag, WORKERS) * 32) : "memory");
                    if (wid == 0 && lane == 0) st_release_cta(&H.u_done, J + 1);
                }
                PROF(0);
                PABS(2);
                // before the rest of block column J is updated: every entry of it exists (the other workers
                // have been generating since they left T_{J-1}) and T_{J-1} is complete -- one barrier for both
                while (ld_uniform(&H.gen_next, lane) < H.gen_end[J]) {
                    const int it = claim(H, lane);
                    if (it < total_items) work_item<MODE, GW>(p, H, tS, nS, M, N, it, nb, lane);
                }
                if (GENONLY)  // items still in the hands of warps that are not part of the barrier
                    while (ld_uniform(&H.ccnt[J], lane) < H.gen_end[J] - H.gen_end[J - 1]) __nanosleep(KC_GPS_SLEEP_T);
                worker_bar<WORKERS>();
                for (int s = ndiag + wid; s < nstrips; s += WORKERS) update_strip(H, M, N, J, s, lane, 0, J);
                worker_bar<WORKERS>();
            }
            PROF(1);
            if (PRE && J >= 1 && J + 1 < nb) {
                // Off the critical path, while the diag warp factors block J: the diagonal tile of block column
                // J+1 takes the contributions of block columns 0 .. J-1 now (all final), so that only block
                // column J's is left between T_J and the next factorisation step.  The last workers do it
                // (they have no T tile at the sizes with 7 or more workers).
                const int nd1 = min(4, (N + 1 - (j0 + T) + 7) / 8);
                const int rev = WORKERS - 1 - wid;
                if (rev < nd1) {
                    while (ld_uniform(&H.dcnt[J + 1], lane) < GQ) {
                        // its generation items come first in the queue of block column J+1: help if unclaimed
                        if (ld_uniform(&H.gen_next, lane) < H.gen_end[J] + GQ) {
                            const int it = claim(H, lane);
                            if (it < total_items) work_item<MODE, GW>(p, H, tS, nS, M, N, it, nb, lane);
                        } else {
                            __nanosleep(KC_GPS_SLEEP_T);
                        }
                    }
                    int done = 0;
                    for (int s = rev; s < nd1; s += WORKERS, done++) update_strip(H, M, N, J + 1, s, lane, 0, J);
                    __syncwarp();
                    if (lane == 0) red_release_add(&H.pre_cnt[J + 1], done);
                }
            }
            // T_J follows the diag warp column by column (one 32-row tile per warp); the
            // other workers generate ahead until the block is factored
            PABS(3);
            const int first = j0 + nbc;
            const int ntiles = (N + 1 - first + T - 1) / T;
            if (wid < ntiles) {
                for (int tl = wid; tl < ntiles; tl += WORKERS) {
                    if (J == 0) {
                        // every item of block column 0 has been claimed by now: the tile completes without us
                        if (tl + 1 < nb)
                            while (ld_uniform(&H.tile_cnt[tl + 1], lane) < GQ) __nanosleep(KC_GPS_SLEEP_T);
                        // the last tile holds row N, the residuals
                        if (tl == ntiles - 1)
                            while (ld_uniform(&H.mean_cnt, lane) < H.gen_end[0] - H.gen0) __nanosleep(KC_GPS_SLEEP_T);
                    }
                    if (nbc == T) trsm_tile_full<ROLLED>(H, M, N, J, tl, lane);
                    else trsm_tile(H, M, N, J, nbc, tl, lane);
                    if (tl == 0) {
                        __syncwarp();
                        if (lane == 0) st_release_cta(&H.t0_done, J + 1);
                    }
                }
            } else {
                while (ld_uniform(&H.d_done, lane) < J + 1) {
                    if (ld_uniform(&H.gen_next, lane) < total_items) {
                        const int it = claim(H, lane);
                        if (it < total_items) work_item<MODE, GW>(p, H, tS, nS, M, N, it, nb, lane);
                    } else {
                        __nanosleep(KC_GPS_SLEEP_F);
                    }
                }
            }
            PROF(2);
            PABS(4);
#if KC_GPS_ENDBAR
            if (NW < 16) worker_bar<WORKERS>();
#endif
        }
    }
    __syncthreads();
#ifdef KC_GPS_PROFILE
    if (b == gridDim.x / 2 && tid == 0) {
        printf("N %d warps %d: setup %lld mean+phases %lld factor %lld; warp 1 waited for col_done in T, all J: %lld\n", N, NW,
               prof_c[1] - prof_c[0], prof_c[2] - prof_c[1], clock64() - prof_c[2], g_prof_twait[blockIdx.x & 1023]);
        printf("warp 1 in T, all J: substitution %lld trailing %lld\n", g_prof_tsub[blockIdx.x & 1023], g_prof_tdmma[blockIdx.x & 1023]);
    }
    if (b == gridDim.x / 2 && tid == 0)
        for (int J = 0; J < nb; J++)
            printf("J%d abs: chol %lld .. %lld | warp 1: strips done %lld, T %lld .. %lld\n", J, prof_abs[J][0], prof_abs[J][1],
                   prof_abs[J][2], prof_abs[J][3], prof_abs[J][4]);
    if (b == gridDim.x / 2 && tid == 0)
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
#ifdef KC_GPS_PROFILE
        if (b < 32768) g_prof_cta[3 * b + 1] = prof_globaltimer();
#endif
    }
    if (CACHE && H.slot >= 0) {
        // factor cache, a miss: the factor goes to the slot's older entry (not a failed factorisation)
        GpsCacheMeta& me = gc.meta[(size_t)H.slot * 2 + H.victim];
        double* dst = gc.img + ((size_t)H.slot * 2 + H.victim) * gc.img_doubles;
        const size_t d_tiles = gc.img_doubles - (size_t)nb * T;
        fence_proxy_async();  // the factor was written with ordinary stores: every writer fences, then the barrier
        __syncthreads();
        if (tid == 0) {
            const size_t bytes = d_tiles * sizeof(double);
            for (size_t o = 0; o < bytes; o += 32768)
                bulk_store(reinterpret_cast<unsigned char*>(dst) + o, reinterpret_cast<const unsigned char*>(M) + o,
                           (uint32_t)(bytes - o < 32768 ? bytes - o : 32768));
            bulk_store(dst + d_tiles, &H.dinv[0][0], (uint32_t)((size_t)nb * T * sizeof(double)));
            bulk_store_commit_wait();
        }
        if (tid < 4 + KC_MAX_INST) me.key[tid] = cache_key(p, b, tid);
        if (tid == 0) {
            me.logdet = logdet;
            me.valid = ok ? 1 : 0;
            me.stamp = gc.stamp;
        }
    }
    if (!CACHE) break;
    __syncthreads();  // the next particle of this CTA reuses the shared memory
    item += gridDim.x;
    } while (item < gc.cnt[1]);
}

}  // namespace

template <int MODE, int NW>
static cudaError_t launch_one(const KParams& p, size_t smem, cudaStream_t stream, const GpsCache& gc)
{
    cudaError_t e;
    if (gc.img) {
        // the misses of a cached batch: a persistent grid, one wave of CTAs walking the miss list
        e = cudaFuncSetAttribute(gps_kernel<MODE, NW, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
        int dev = 0, sms = 0, per_sm = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, gps_kernel<MODE, NW, true>, NW * 32, smem);
        if (e != cudaSuccess) return e;
        int grid = sms * (per_sm > 0 ? per_sm : 1);
        if (grid > p.B) grid = p.B;
        gps_kernel<MODE, NW, true><<<grid, NW * 32, smem, stream>>>(p, gc);
        return cudaGetLastError();
    }
    e = cudaFuncSetAttribute(gps_kernel<MODE, NW, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    gps_kernel<MODE, NW, false><<<p.B, NW * 32, smem, stream>>>(p, gc);
    return cudaGetLastError();
}

// developer check: cached against uncached results of one launch, mismatches reported from the device
__global__ void gps_cache_compare_kernel(const KParams p, const double* plain, const GpsCache gc)
{
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= p.B) return;
    const double a = p.logL[b], r = plain[b];
    if (__double_as_longlong(a) != __double_as_longlong(r))
        printf("# factor cache MISMATCH lane %d slot %d class %d: cached %.17g uncached %.17g np %d eta %.17g %.17g %.17g %.17g jit %.17g\n", b,
               gc.slots[b], gc.cls[b], a, r, p.np[b], ldp(p, p.L.row_eta, b), ldp(p, p.L.row_eta + 1, b), ldp(p, p.L.row_eta + 2, b),
               ldp(p, p.L.row_eta + 3, b), ldp(p, p.L.row_jit, b));
}
void launch_gp_cache_compare(const KParams& p, const double* plain, const GpsCache& gc, cudaStream_t s)
{
    gps_cache_compare_kernel<<<(p.B + 127) / 128, 128, 0, s>>>(p, plain, gc);
}

// Lanes in order of planet count, most planets first (one CTA, a counting sort; the order within a count is
// whatever the atomics make it: nothing depends on it).  Why: the kernel's hot code (~60-100 KB over its warp
// roles) does not fit the SM's instruction cache, and SMs that run the same phase at the same time feed each
// other through the shared levels.  A batch whose neighbouring CTAs differ in planet count -- any real one --
// drifts out of step and EVERY particle gets slower, the planet-free ones included (N = 177, Np ~ U{0..5}:
// 61.8 us per particle against 48.6 us for the same lanes sorted, 6.37 vs 5.27 ms per 16 384 lanes,
// profiles/gps_timeline_r02.jsonl); CTAs are handed out in index order, so sorting keeps the SMs in step.
// With `list` (the miss list of a cached batch, *n_list entries): the same over the lanes of that list.
__global__ void __launch_bounds__(1024) gps_order_kernel(const int32_t* __restrict__ np, int B, int* __restrict__ order,
                                                         const int* __restrict__ list, const int* __restrict__ n_list)
{
    constexpr int BUCKETS = 64;
    __shared__ int cnt[BUCKETS], off[BUCKETS];
    const int tid = threadIdx.x;
    if (list) B = *n_list;
    if (tid < BUCKETS) cnt[tid] = 0;
    __syncthreads();
    // warp-aggregated: lanes of a warp with the same count send one atomic (a batch of equal counts would
    // otherwise serialise B atomics on one address)
    const int nround = (B + blockDim.x - 1) / blockDim.x;
    for (int r = 0; r < nround; r++) {
        const int i = r * blockDim.x + tid;
        const int b = i < B ? (list ? list[i] : i) : -1;
        const int key = b >= 0 ? min(max(np[b], 0), BUCKETS - 1) : -1;
        const unsigned peers = __match_any_sync(0xffffffffu, key);
        if (key >= 0 && (tid & 31) == __ffs(peers) - 1) atomicAdd(&cnt[key], __popc(peers));
    }
    __syncthreads();
    if (tid == 0) {
        int o = 0;
        for (int k = BUCKETS - 1; k >= 0; k--) {
            off[k] = o;
            o += cnt[k];
        }
    }
    __syncthreads();
    for (int r = 0; r < nround; r++) {
        const int i = r * blockDim.x + tid;
        const int b = i < B ? (list ? list[i] : i) : -1;
        const int key = b >= 0 ? min(max(np[b], 0), BUCKETS - 1) : -1;
        const unsigned peers = __match_any_sync(0xffffffffu, key);
        const int leader = __ffs(peers) - 1, lane = tid & 31;
        int base = 0;
        if (key >= 0 && lane == leader) base = atomicAdd(&off[key], __popc(peers));
        base = __shfl_sync(0xffffffffu, base, leader);
        if (key >= 0) order[base + __popc(peers & ((1u << lane) - 1))] = b;
    }
}

#ifdef KC_GPS_PROFILE
extern "C" int kc_prof_fetch_cta(unsigned long long* out, int n)
{
    return (int)cudaMemcpyFromSymbol(out, g_prof_cta, sizeof(unsigned long long) * 3 * (size_t)n);
}
#endif

cudaError_t launch_gp_small(const KParams& p0, int kepler_mode, cudaStream_t stream, int* n_launches,
                            const GpsCache* cache, int* order_scratch)
{
    if (p0.B <= 0) return cudaSuccess;
    KParams p = p0;
    p.order = nullptr;
    if (order_scratch && p.B >= 1024 && !(cache && cache->img && cache->slots)) {
        gps_order_kernel<<<1, 1024, 0, stream>>>(p.np, p.B, order_scratch, nullptr, nullptr);
        p.order = order_scratch;
        *n_launches += 1;
    }
    GpsCache gc{};
    // with the MA(1) term the residual recurrence uses the factor storage as scratch: no cache then
    if (cache && cache->img && cache->slots && !p.F.MA) gc = *cache;
    const size_t smem0 = gps_smem_bytes(p.N);
    if (!smem0) return cudaErrorInvalidValue;
    const int fit0 = (int)((size_t)(228 * 1024) / (smem0 + 1024));
    if (gc.img) {
        // hits first (a warp each, the factor read from HBM / L2), then the misses through the full kernel
        cudaError_t e0 = cudaMemsetAsync(gc.cnt, 0, 2 * sizeof(int), stream);
        if (e0 != cudaSuccess) return e0;
        gps_classify_kernel<<<(p.B + 127) / 128, 128, 0, stream>>>(p, gc);
        const size_t hsm = sizeof(hit::WarpState) * hit::WARPS;
        const int pre = fit0 < 2 ? 1 : 0;  // 16-warp CTAs update the look-ahead strips one block column behind
        const int hgrid = (p.B + hit::WARPS - 1) / hit::WARPS;
        if (kepler_mode == 0) {
            e0 = cudaFuncSetAttribute(gps_hit_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)hsm);
            if (e0 != cudaSuccess) return e0;
            gps_hit_kernel<0><<<hgrid, hit::WARPS * 32, hsm, stream>>>(p, gc, pre);
        } else {
            e0 = cudaFuncSetAttribute(gps_hit_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)hsm);
            if (e0 != cudaSuccess) return e0;
            gps_hit_kernel<1><<<hgrid, hit::WARPS * 32, hsm, stream>>>(p, gc, pre);
        }
        *n_launches += 2;
        if (order_scratch && p.B >= 1024) {
            // the misses in order of planet count too: the CTAs of the persistent grid walk the list in step
            gps_order_kernel<<<1, 1024, 0, stream>>>(p.np, p.B, order_scratch, gc.miss_list, gc.cnt + 1);
            gc.miss_list = order_scratch;
            *n_launches += 1;
        }
    }
    const size_t smem = gps_smem_bytes(p.N);
    if (!smem) return cudaErrorInvalidValue;
    // CTAs per SM by shared-memory footprint (228 KB per SM, 1 KB reserved per CTA)
    const size_t per_sm = 228 * 1024;
    const int fit = (int)(per_sm / (smem + 1024));
    cudaError_t e;
    if (fit >= 4) e = kepler_mode == 0 ? launch_one<0, 4>(p, smem, stream, gc) : launch_one<1, 4>(p, smem, stream, gc);
    else if (fit == 3) e = kepler_mode == 0 ? launch_one<0, 5>(p, smem, stream, gc) : launch_one<1, 5>(p, smem, stream, gc);
    else if (fit == 2) e = kepler_mode == 0 ? launch_one<0, 8>(p, smem, stream, gc) : launch_one<1, 8>(p, smem, stream, gc);
    else e = kepler_mode == 0 ? launch_one<0, 16>(p, smem, stream, gc) : launch_one<1, 16>(p, smem, stream, gc);
    *n_launches += 1;
    return e;
}

}  // namespace kc
