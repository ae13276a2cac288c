// Pieces shared by the two GP kernel files: the quasi-periodic kernel with the
// reference's roundings (/root/reference/src/RVmodel.cpp:193-194), the DMMA wrapper
// and the per-lane state of the blocked path.
#pragma once
#include "kc_model.cuh"

namespace kc {

// sin(x) alone, |x| < ~1e5: one polynomial, chosen by quadrant.
__device__ __forceinline__ double sin_core(double x);
__device__ __forceinline__ double sin_fast(double x)
{
    if (fabs(x) > 1.0e5) return sin(x);
    return sin_core(x);
}

// The same without the large-argument guard (callers check |x| <= 1e5 themselves).
__device__ __forceinline__ double sin_core(double x)
{
    const double S1 = -1.66666666666666324348e-01, S2 = 8.33333333332248946124e-03,
                 S3 = -1.98412698298579493134e-04, S4 = 2.75573137070700676789e-06,
                 S5 = -2.50507602534068634195e-08, S6 = 1.58969099521155010221e-10;
    const double C1 = 4.16666666666666019037e-02, C2 = -1.38888888888741095749e-03,
                 C3 = 2.48015872894767294178e-05, C4 = -2.75573143513906633035e-07,
                 C5 = 2.08757232129817482790e-09, C6 = -1.13596475577881948265e-11;
    int k;
    double kd = rint_magic_i(x * TWO_OVER_PI, k);
    double r = kfma(-kd, PIO2_HI, x);
    r = kfma(-kd, PIO2_MID, r);
    r = kfma(-kd, PIO2_LO, r);
    const bool odd = k & 1;
    double z = r * r;
    double p = kfma(z, odd ? C6 : S6, odd ? C5 : S5);
    p = kfma(z, p, odd ? C4 : S4);
    p = kfma(z, p, odd ? C3 : S3);
    p = kfma(z, p, odd ? C2 : S2);
    p = kfma(z, p, odd ? C1 : S1);
    double base = odd ? kfma(-0.5, z, 1.0) : r;
    double mult = odd ? z * z : r * z;
    double v = kfma(mult, p, base);
    return (k & 2) ? -v : v;
}

struct GpConsts {
    double eta1sq;            // eta1*eta1
    double eta2, r_eta2;      // eta and its correctly rounded reciprocal
    double eta3, r_eta3;
    double eta4, r_eta4;
};

// Per-lane state of the blocked path, living beside the factor storage in HBM.
struct __align__(16) GpLane {
    double dinv[2][64];       // 1 / diag(L_JJ), double-buffered by block-column parity (cp.async source)
    GpConsts c;
    double jit2[KC_MAX_INST];
    double logdet, quad;
    int fail, pad;
    double pad2;
};
static_assert(sizeof(GpLane) % 16 == 0, "lanes[] elements stay 16-byte aligned");

__device__ __forceinline__ GpConsts gp_consts(const ParticleScalars& sc)
{
    GpConsts g;
    g.eta1sq = kmul(sc.eta1, sc.eta1);
    g.eta2 = sc.eta2; g.r_eta2 = kdiv(1.0, sc.eta2);
    g.eta3 = sc.eta3; g.r_eta3 = kdiv(1.0, sc.eta3);
    g.eta4 = sc.eta4; g.r_eta4 = kdiv(1.0, sc.eta4);
    return g;
}

// C(i,j) off the diagonal, with the reference's roundings (:193-194):
// eta1*eta1*exp(-0.5*pow(tau/eta2,2) - 2.0*pow(sin(M_PI*tau/eta3)/eta4,2)).
__device__ __forceinline__ double qp_kernel(const GpConsts& g, double tau)
{
    double u = div_by(tau, g.eta2, g.r_eta2);
    double a1 = -0.5 * kmul(u, u);
    double s = sin_fast(div_by(kmul(PI, tau), g.eta3, g.r_eta3));
    double v = div_by(s, g.eta4, g.r_eta4);
    double a2 = 2.0 * kmul(v, v);
    return kmul(g.eta1sq, exp_nonpos(kadd(a1, -a2)));
}

// Two entries at once: two independent, branch-free dependency chains for the FP64 pipe
// (one chain per thread leaves it half idle at 4 warps per scheduler).
__device__ __forceinline__ void qp_kernel2(const GpConsts& g, double tau0, double tau1,
                                           double& c0, double& c1)
{
    double u0 = div_by(tau0, g.eta2, g.r_eta2), u1 = div_by(tau1, g.eta2, g.r_eta2);
    double x0 = div_by(kmul(PI, tau0), g.eta3, g.r_eta3), x1 = div_by(kmul(PI, tau1), g.eta3, g.r_eta3);
    double s0 = sin_core(x0), s1 = sin_core(x1);
    if (!(fabs(x0) <= 1.0e5) || !(fabs(x1) <= 1.0e5)) {  // never taken with any sane eta3
        s0 = sin(x0);
        s1 = sin(x1);
    }
    double v0 = div_by(s0, g.eta4, g.r_eta4), v1 = div_by(s1, g.eta4, g.r_eta4);
    double e0 = kadd(-0.5 * kmul(u0, u0), -2.0 * kmul(v0, v0));
    double e1 = kadd(-0.5 * kmul(u1, u1), -2.0 * kmul(v1, v1));
    c0 = kmul(g.eta1sq, exp_nonpos(e0));
    c1 = kmul(g.eta1sq, exp_nonpos(e1));
}

__device__ __forceinline__ void dmma(double& c0, double& c1, double a, double b)
{
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                 : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}

// Unblocked Cholesky of a 32x32 block (LLT.h:313-337) by one warp, row `lane` of the block
// in s[0..lane].  Rolled loops over the columns: every step consumes s[0] and writes the
// updated row shifted down by one (s[k-1] <- s[k] - L(r,c) L(c+k,c)), so all register
// indices are static and the body stays in the instruction cache; the window of live
// entries narrows every 8 columns (31, 23, 15, 7).  The finished column goes to Lout
// (Lout(c, r) = L(r, c), column-major; HBM or shared) and, transposed, to LT if wanted.
// `progress` (shared memory, may be null): set to progress_base + c + 1 once column c and
// 1/L(c,c) are visible, so that other warps can follow the factorisation column by column.
// CTA-scope release store / acquire load for flags in shared memory (a __threadfence_block()
// is a sequentially consistent MEMBAR.SC.CTA, far more than a producer/consumer flag needs).
__device__ __forceinline__ void st_release_cta(int* p, int v)
{
    asm volatile("st.release.cta.shared.s32 [%0], %1;" ::"r"((unsigned)__cvta_generic_to_shared(p)), "r"(v) : "memory");
}
__device__ __forceinline__ int ld_acquire_cta(const int* p)
{
    int v;
    asm volatile("ld.acquire.cta.shared.s32 %0, [%1];" : "=r"(v) : "r"((unsigned)__cvta_generic_to_shared(p)) : "memory");
    return v;
}

struct Chol32State {
    double d, inv, l, lmine;
    bool ok;
};

// 1/sqrt(d) from the MUFU seed + two Newton steps, sqrt(d) = d/sqrt(d) + one Heron step
__device__ __forceinline__ void chol_pivot(double d, double& inv, double& l)
{
    double y = rsqrt_seed(d);
    double hy = 0.5 * y;
    double e = kfma(-d * y, hy, 0.5);
    y = kfma(y, e, y);
    hy = 0.5 * y;
    l = d * y;
    e = kfma(-l, hy, 0.5);
    y = kfma(y, e, y);
    l = kfma(kfma(-l, l, d), 0.5 * y, l);
    inv = y;
}

// Column broadcast buffer of chol32: two parity halves of CHOL_CB_STRIDE doubles.  The lanes
// read L(c+1..c+W, c) of the finished column as 16-byte pairs, so the column is stored with
// a shift of one element for even c (element c+1 then sits at an even index either way).
constexpr int CHOL_CB_STRIDE = 72;
constexpr int CHOL_COLBUF = 2 * CHOL_CB_STRIDE;

template <int W>
__device__ __forceinline__ void chol32_steps(double (&s)[32], Chol32State& st, int c_lo, int c_hi, int lane,
                                             int nvalid, double* dinv_out, double* __restrict__ Lout,
                                             size_t ld, double* LT, int lt_ld, double* colbuf,
                                             int* progress, int progress_base)
{
    constexpr int NP2 = (W + 1) / 2;
#pragma unroll 1
    for (int c = c_lo; c < c_hi; c++) {
        if (!(st.d > 0.0)) st.ok = false;  // NaN or non-positive pivot; the NaNs that follow are discarded
        if (lane == c) st.lmine = st.l;
        if (lane == 0) dinv_out[c] = st.inv;
        const double lrc = s[0] * st.inv;
        const double lcc = st.l;
        // The next pivot is known to lane c+1 before the column is broadcast:
        // s'[0] = s[1] - L(c+1,c)^2, the same FMA the update below performs.  Its
        // rsqrt chain (latency-bound) then overlaps the loads + FMAs (issue-bound).
        st.d = __shfl_sync(0xffffffffu, kfma(-lrc, lrc, s[1]), (c + 1) & 31);
        const double dn = st.d;
        const int shift = (c & 1) ^ 1;
        double* cb = colbuf + (c & 1) * CHOL_CB_STRIDE + shift;
        cb[lane] = lrc;
        cb[lane + 32] = lrc;
        if (lane >= c && lane < nvalid) Lout[(size_t)c * ld + lane] = (lane == c) ? lcc : lrc;
        if (LT) LT[c * lt_ld + lane] = (lane > c) ? lrc : 0.0;
        __syncwarp();
        if (progress && lane == 0) st_release_cta(progress, progress_base + c + 1);
        // col2[j] = { L(c+1+2j, c), L(c+2+2j, c) }: c + 1 + shift is even
        const double2* col2 = reinterpret_cast<const double2*>(cb + c + 1);
        double2 cv[NP2];
#pragma unroll
        for (int j = 0; j < NP2; j++) cv[j] = col2[j];
        // A warp issues in order: the pivot chain (chol_pivot, ~10 dependent FP64 operations)
        // is written out here stage by stage with a group of the W independent update FMAs
        // after every stage, so that its latency is covered by them instead of preceding them.
        constexpr int G = (W + 7) / 8;
#define CHOL_FMAS(g)                                                                            \
        _Pragma("unroll") for (int k = 1 + (g) * G; k <= W && k <= ((g) + 1) * G; k++) {        \
            const double v = ((k - 1) & 1) ? cv[(k - 1) >> 1].y : cv[(k - 1) >> 1].x;           \
            s[k - 1] = kfma(-lrc, v, s[k]);                                                     \
        }                                                                                       \
        asm volatile("" : "+d"(y), "+d"(e), "+d"(l), "+d"(s[(g) * G]));
        double y = rsqrt_seed(dn), e = 0.0, l = 0.0;
        CHOL_FMAS(0)
        double hy = 0.5 * y;
        e = kfma(-dn * y, hy, 0.5);
        CHOL_FMAS(1)
        y = kfma(y, e, y);
        CHOL_FMAS(2)
        hy = 0.5 * y;
        l = dn * y;
        CHOL_FMAS(3)
        e = kfma(-l, hy, 0.5);
        CHOL_FMAS(4)
        y = kfma(y, e, y);
        CHOL_FMAS(5)
        e = kfma(-l, l, dn);
        CHOL_FMAS(6)
        l = kfma(e, 0.5 * y, l);
        CHOL_FMAS(7)
#undef CHOL_FMAS
        st.inv = y;
        st.l = l;
        s[W] = 0.0;
    }
}

__device__ __forceinline__ bool chol32_rows_in_lanes(double (&s)[32], int lane, int nvalid,
                                                     double& logdet, double* dinv_out,
                                                     double* __restrict__ Lout, size_t ld,
                                                     double* LT, int lt_ld, double* colbuf,
                                                     int* progress = nullptr, int progress_base = 0)
{
    // colbuf: CHOL_COLBUF doubles of 16-byte-aligned shared memory; the finished column is
    // broadcast through it (written twice, 32 apart, so that L(c+k, c) is read without
    // wrap-around; double-buffered by the parity of c: one __syncwarp per column)
    Chol32State st;
    st.ok = true;
    st.lmine = 1.0;  // lane c keeps L(c,c) for the log-determinant
    st.d = __shfl_sync(0xffffffffu, s[0], 0);
    chol_pivot(st.d, st.inv, st.l);
    chol32_steps<31>(s, st, 0, 8, lane, nvalid, dinv_out, Lout, ld, LT, lt_ld, colbuf, progress, progress_base);
    chol32_steps<23>(s, st, 8, 16, lane, nvalid, dinv_out, Lout, ld, LT, lt_ld, colbuf, progress, progress_base);
    chol32_steps<15>(s, st, 16, 24, lane, nvalid, dinv_out, Lout, ld, LT, lt_ld, colbuf, progress, progress_base);
    chol32_steps<7>(s, st, 24, 32, lane, nvalid, dinv_out, Lout, ld, LT, lt_ld, colbuf, progress, progress_base);
    double lg = lane < nvalid ? log(st.lmine) : 0.0;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) lg += __shfl_xor_sync(0xffffffffu, lg, o);
    logdet += lg;
    return st.ok;
}

}  // namespace kc
