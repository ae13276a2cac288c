// Pieces shared by the two GP kernel files: the quasi-periodic kernel with the
// reference's roundings (/root/reference/src/RVmodel.cpp:193-194), the DMMA wrapper
// and the per-lane state of the blocked path.
#pragma once
#include "kc_model.cuh"

namespace kc {

// Polynomial coefficients in the constant bank: an FP64 instruction takes a constant-bank operand
// directly, whereas a 64-bit literal costs two uniform-register moves per use and a runtime choice
// between two literals two selects more (ncu, r02: 25 UMOV + 28 FSEL of the ~200 instructions per
// covariance entry).  The arithmetic is unchanged: same operations, same order, same roundings.
static __constant__ double KC_SINCOS_POLY[2][6] = {
    // [0]: sin on [-pi/4, pi/4], S6 .. S1;  [1]: cos, C6 .. C1 (fdlibm)
    {1.58969099521155010221e-10, -2.50507602534068634195e-08, 2.75573137070700676789e-06,
     -1.98412698298579493134e-04, 8.33333333332248946124e-03, -1.66666666666666324348e-01},
    {-1.13596475577881948265e-11, 2.08757232129817482790e-09, -2.75573143513906633035e-07,
     2.48015872894767294178e-05, -1.38888888888741095749e-03, 4.16666666666666019037e-02}};
static __constant__ double KC_EXP_TAYLOR[12] = {
    1.0 / 6227020800.0, 1.0 / 479001600.0, 1.0 / 39916800.0, 1.0 / 3628800.0, 1.0 / 362880.0, 1.0 / 40320.0,
    1.0 / 5040.0, 1.0 / 720.0, 1.0 / 120.0, 1.0 / 24.0, 1.0 / 6.0, 0.5};
static __constant__ double KC_RED[6] = {TWO_OVER_PI, PIO2_HI, PIO2_MID, PIO2_LO, 1.44269504088896338700e+00 /* log2 e */,
                                        6.93147180559890330187e-01 /* ln2 hi, 42 bits */};

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
#if defined(__CUDA_ARCH__)
    double kd = rint_magic_i(x * KC_RED[0], k);
    double r = kfma(-kd, KC_RED[1], x);
    r = kfma(-kd, KC_RED[2], r);
    r = kfma(-kd, KC_RED[3], r);
    const bool odd = k & 1;
    double z = r * r;
    const double* cf = KC_SINCOS_POLY[k & 1];  // one indexed constant load per coefficient
    double p = kfma(z, cf[0], cf[1]);
    p = kfma(z, p, cf[2]);
    p = kfma(z, p, cf[3]);
    p = kfma(z, p, cf[4]);
    p = kfma(z, p, cf[5]);
    (void)S1; (void)S2; (void)S3; (void)S4; (void)S5; (void)S6;
    (void)C1; (void)C2; (void)C3; (void)C4; (void)C5; (void)C6;
#else
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
#endif
    double base = odd ? kfma(-0.5, z, 1.0) : r;
    double mult = (odd ? z : r) * z;
    double v = kfma(mult, p, base);
#if defined(__CUDA_ARCH__)
    // quadrants 2, 3: flip the sign bit (the same value as -v, one integer instruction)
    return __hiloint2double(__double2hiint(v) ^ ((k & 2) << 30), __double2loint(v));
#else
    return (k & 2) ? -v : v;
#endif
}

// exp_nonpos (kc_math.cuh) with its literals taken from the constant bank; bit-identical results.
__device__ __forceinline__ double exp_nonpos_cb(double x)
{
    double xc = x < -746.0 ? -746.0 : x;  // NaN falls through
    int k;
    double kd = rint_magic_i(xc * KC_RED[4], k);
    double rh = kfma(-kd, KC_RED[5], xc);          // exact
    double rl = -kd * 5.49792301870837115524e-14;   // ln2 lo
    double r = rh + rl;
    double p = KC_EXP_TAYLOR[0];
#pragma unroll
    for (int i = 1; i < 12; i++) p = kfma(p, r, KC_EXP_TAYLOR[i]);
    double r2 = r * r;
    double t1 = kadd(1.0, rh);
    double lo = kadd(kadd(rh, -kadd(t1, -1.0)), rl);
    p = kadd(t1, kfma(p, r2, lo));
    // 2^(k+64) as a high word: no 64-bit shift
    double s1 = __hiloint2double((k + 64 + 1023) << 20, 0);
    // exp(-746) = 9.9e-325 rounds to zero by itself: the clamp above already makes x < -746 return 0
    return (p * s1) * 5.42101086242752217004e-20;  // 2^-64
}

// ---- mbarrier + bulk-copy plumbing (sm_100a: SASS SYNCS / UBLKCP), shared by the GP kernels -------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* b, int count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(b)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_fence_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void mbar_expect_tx(uint64_t* b, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(b)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* b)
{
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(b)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* b, uint32_t parity)
{
    asm volatile(
        "{\n"
        ".reg .pred P1;\n"
        "LAB_WAIT:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n"
        "@P1 bra DONE;\n"
        "bra LAB_WAIT;\n"
        "DONE:\n"
        "}" ::"r"(smem_u32(b)), "r"(parity) : "memory");
}
__device__ __forceinline__ void bulk_load(void* dst, const void* src, uint32_t bytes, uint64_t* bar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}

// shared memory -> global, asynchronously: fence the generic-proxy writes first (after the CTA barrier that
// orders them), commit, and wait before the kernel ends
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void bulk_store(void* dst, const void* src, uint32_t bytes)
{
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst), "r"(smem_u32(src)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_store_commit_wait()
{
    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
    asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}

struct GpConsts {
    double eta1sq;            // eta1*eta1
    double eta2, r_eta2;      // eta and its correctly rounded reciprocal
    double eta3, r_eta3;
    double eta4, r_eta4;
    double g1, g2;            // -0.5 / eta2^2 and -2 / eta4^2: the two coefficients of the exponent (qp_entry)
};

// ---- the covariance generator of the product path ---------------------------------------------
// C(i,j) = eta1^2 exp(-tau^2 / (2 eta2^2) - 2 sin^2(pi tau / eta3) / eta4^2), tau = t_i - t_j (:193-194).
// Evaluated as written that is ~75 FP64 operations per entry (three divisions, a sine with its range
// reduction, a 12-term exponential), and the N^2/2 entries are the second largest consumer of the FP64
// pipe after the factorisation itself.  The product path uses two facts instead:
//   * sin(pi (t_i - t_j) / eta3) = s_i c_j - c_i s_j with s_i = sin(pi t_i / eta3), c_i = cos(pi t_i / eta3):
//     2N sines per particle instead of N^2/2.  The phases are formed from a double-double quotient
//     t_i / eta3 (gp_phase), so s_i, c_i carry <= 2 ulp whatever the magnitude of t_i (JD 2.45e6 included)
//     and the difference formula is as accurate as the direct one for every pair, near or far;
//   * exp(x), x <= 0, from a 64-entry table of 2^(j/64) and a degree-5 polynomial on |r| <= ln2/128.
// 19 FP64 operations per entry.  The entries agree with the reference's to a few 1e-16 eta1^2 -- the
// same distance the reference's own entries have from an 80-bit evaluation of the kernel
// (tools/gp_generator_error_budget.py; tests/test_gpu_parity.py pins the bound through kc_gp_cov).
static __device__ const double KC_EXP2_TAB[64] = {
    1.00000000000000000e+00, 1.01088928605170048e+00, 1.02189714865411663e+00, 1.03302487902122841e+00,
    1.04427378242741375e+00, 1.05564517836055716e+00, 1.06714040067682370e+00, 1.07876079775711986e+00,
    1.09050773266525769e+00, 1.10238258330784089e+00, 1.11438674259589243e+00, 1.12652161860824185e+00,
    1.13878863475669156e+00, 1.15118922995298267e+00, 1.16372485877757748e+00, 1.17639699165028122e+00,
    1.18920711500272103e+00, 1.20215673145270308e+00, 1.21524735998046896e+00, 1.22848053610687002e+00,
    1.24185781207348400e+00, 1.25538075702469110e+00, 1.26905095719173322e+00, 1.28287001607877826e+00,
    1.29683955465100964e+00, 1.31096121152476441e+00, 1.32523664315974132e+00, 1.33966752405330292e+00,
    1.35425554693689265e+00, 1.36900242297459052e+00, 1.38390988196383202e+00, 1.39897967253831124e+00,
    1.41421356237309515e+00, 1.42961333839197002e+00, 1.44518080697704665e+00, 1.46091779418064704e+00,
    1.47682614593949935e+00, 1.49290772829126484e+00, 1.50916442759342284e+00, 1.52559815074453842e+00,
    1.54221082540794074e+00, 1.55900440023783693e+00, 1.57598084510788650e+00, 1.59314215134226700e+00,
    1.61049033194925428e+00, 1.62802742185734783e+00, 1.64575547815396495e+00, 1.66367658032673638e+00,
    1.68179283050742900e+00, 1.70010635371852348e+00, 1.71861929812247793e+00, 1.73733383527370622e+00,
    1.75625216037329945e+00, 1.77537649252652119e+00, 1.79470907500310717e+00, 1.81425217550039886e+00,
    1.83400808640934243e+00, 1.85397912508338547e+00, 1.87416763411029996e+00, 1.89457598158696561e+00,
    1.91520656139714740e+00, 1.93606179349229435e+00, 1.95714412417540018e+00, 1.97845602638795093e+00};
// 64 / ln 2, ln 2 / 64 split (high part: 32 significant bits, so k * hi is exact for |k| < 2^20), 1/120 .. 1/2
static __constant__ double KC_EXPT[7] = {9.23324826168936567683e+01, 1.08304246932675596327e-02, 2.98158582698529328128e-12,
                                         1.0 / 120.0, 1.0 / 24.0, 1.0 / 6.0, 0.5};

// tab[64] in shared memory <- 2^(j/64); the caller synchronises before the first qp_entry
__device__ __forceinline__ void exp_tab_load(double* tab, int tid)
{
    if (tid < 64) tab[tid] = KC_EXP2_TAB[tid];
}

// exp(x) for x <= 0 (NaN propagates; x < -708 returns exp(-708) = 3.3e-308, i.e. nothing), < 1 ulp + the
// table's half ulp.  11 FP64 operations and one shared-memory load.
__device__ __forceinline__ double exp_tab(double x, const double* __restrict__ tab)
{
    double xc = x < -708.0 ? -708.0 : x;
    int k;
    double kd = rint_magic_i(xc * KC_EXPT[0], k);
    double r = kfma(-kd, KC_EXPT[1], xc);  // exact
    r = kfma(-kd, KC_EXPT[2], r);
    double q = kfma(r, KC_EXPT[3], KC_EXPT[4]);
    q = kfma(q, r, KC_EXPT[5]);
    q = kfma(q, r, KC_EXPT[6]);
    double pl = kfma(q, r * r, r);         // e^r - 1
    const double tj = tab[k & 63];
    double v = kfma(tj, pl, tj);
    // 2^(k >> 6), k >> 6 in [-1022, 0]: a normal number, built as a high word
    return v * __hiloint2double(((k >> 6) + 1023) << 20, 0);
}

// s = sin(pi t / eta3), c = cos(pi t / eta3) to <= 2 ulp for any finite t: the quotient in double-double
// (q correctly rounded, the remainder t - q eta3 is exact), its integer part dropped exactly.
__device__ __forceinline__ void gp_phase(const GpConsts& g, double t, double& s, double& c)
{
    double q = div_by(t, g.eta3, g.r_eta3);
    double ql = kfma(-q, g.eta3, t) * g.r_eta3;
    double n = rint(q);
    double w = (q - n) + ql;
    sincospi(w, &s, &c);
    // an odd integer part flips both (|q| >= 2^53 has no odd values; NaN stays NaN)
    if (fabs(n) < 9.0e15 && (((long long)n) & 1)) {
        s = -s;
        c = -c;
    }
}

// One off-diagonal entry from the phases of its two epochs.
__device__ __forceinline__ double qp_entry(const GpConsts& g, const double* __restrict__ tab, double tau,
                                           double si, double ci, double sj, double cj)
{
    double sd = kfma(-ci, sj, si * cj);           // sin(pi tau / eta3)
    double e = kfma(sd * sd, g.g2, (tau * tau) * g.g1);
    return g.eta1sq * exp_tab(e, tab);
}

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

__device__ __forceinline__ GpConsts gp_consts(double eta1, double eta2, double eta3, double eta4)
{
    GpConsts g;
    g.eta1sq = kmul(eta1, eta1);
    g.eta2 = eta2; g.r_eta2 = kdiv(1.0, eta2);
    g.eta3 = eta3; g.r_eta3 = kdiv(1.0, eta3);
    g.eta4 = eta4; g.r_eta4 = kdiv(1.0, eta4);
    g.g1 = kdiv(-0.5, kmul(eta2, eta2));
    g.g2 = kdiv(-2.0, kmul(eta4, eta4));
    return g;
}
__device__ __forceinline__ GpConsts gp_consts(const ParticleScalars& sc)
{
    return gp_consts(sc.eta1, sc.eta2, sc.eta3, sc.eta4);
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
    return kmul(g.eta1sq, exp_nonpos_cb(kadd(a1, -a2)));
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
    c0 = kmul(g.eta1sq, exp_nonpos_cb(e0));
    c1 = kmul(g.eta1sq, exp_nonpos_cb(e1));
}

__device__ __forceinline__ void dmma(double& c0, double& c1, double a, double b)
{
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                 : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}

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

// Cholesky of a 32x32 diagonal block (LLT.h:313-337) by ONE warp, the block resident in shared
// memory, column-major: element (r, c) at S[c * ldS + r], lower triangle valid, ldS = 4 mod 16.
// Blocked by panels of 8 columns:
//   panel   rows in lanes, the 8 columns of row `lane` in registers; per column the loop-carried
//           chain is scale -> square -> shuffle -> rsqrt (MUFU seed + 2 Newton steps), and the
//           rest of the step is <= 7 shuffle + FMA pairs: the warp is bound by that chain
//           (about 120 cycles per column), not by its issue slots (the unblocked variant of
//           round 1 updated a 31-wide register window per column: 380 cycles, 230 of them issue)
//   update  the 8-column panel goes back to shared memory and the trailing part of the block is
//           updated on the FP64 tensor cores (two DMMA.8x8x4 per 8x8 fragment, read-modify-write
//           of the fragment in shared memory); the columns of the next panel first.
// L overwrites S (rows < nvalid); dinv_out[c] = 1 / L(c, c); L also goes to Lout (column-major,
// leading dimension ld; HBM) when Lout is not null.  Rows / columns >= nvalid of a partial block
// are treated as identity.  MASK: rows >= nvalid of S belong to other warps (the shared-memory
// resident path keeps the residual row and the rows other warps substitute there): they are
// neither read as data nor written.  `progress` (shared memory, may be null) is set to
// progress_base + c + 8 once the panel that starts at column c and its 1/L(c,c) are visible, so
// that other warps can follow the factorisation panel by panel.
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

constexpr int CHOL_PANEL = 8;

// The same with the panel loop rolled (a quarter of the code; ~16 % slower alone: the trailing update's block
// structure is then run-time control flow).
template <bool MASK>
__device__ __forceinline__ bool chol32_panels_rolled(double* S, int ldS, int lane, int nvalid, double& logdet,
                                              double* dinv_out, double* __restrict__ Lout, size_t ld,
                                              int* progress = nullptr, int progress_base = 0)
{
    constexpr unsigned FULL = 0xffffffffu;
    const int fr = lane >> 2, fk = lane & 3;
    bool ok = true;
    double lmine = 1.0;  // lane c keeps L(c,c) for the log-determinant
    double d = 1.0, inv = 1.0, l = 1.0;
#pragma unroll 1
    for (int p = 0; p < 4; p++) {
        const int c0 = CHOL_PANEL * p;
        // ---- the panel: 8 columns of row `lane` ----
        double a[CHOL_PANEL];
#pragma unroll
        for (int k = 0; k < CHOL_PANEL; k++) {
            const int c = c0 + k;
            double v = (lane == c) ? 1.0 : 0.0;  // identity padding of a partial block; zeros above the diagonal
            if (lane < nvalid && c <= lane) v = S[c * ldS + lane];
            a[k] = v;
        }
        d = __shfl_sync(FULL, a[0], c0);
        chol_pivot(d, inv, l);
#pragma unroll
        for (int k = 0; k < CHOL_PANEL; k++) {
            const int c = c0 + k;
            if (!(d > 0.0)) ok = false;  // NaN or non-positive pivot; the NaNs that follow are discarded
            if (lane == c) lmine = l;
            if (lane == k) dinv_out[c] = inv;
            const double lrc = a[k] * inv;  // L(lane, c) below the diagonal
            const double lcc = l;
            if (k + 1 < CHOL_PANEL) {
                // The next pivot is known to lane c+1 before the column is passed round:
                // S(c+1,c+1) - L(c+1,c)^2, the same FMA the update below performs.  Its rsqrt
                // chain (latency-bound) overlaps the shuffles + FMAs of the rest of the panel.
                d = __shfl_sync(FULL, kfma(-lrc, lrc, a[k + 1]), c + 1);
                chol_pivot(d, inv, l);
#pragma unroll
                for (int k2 = k + 1; k2 < CHOL_PANEL; k2++)
                    a[k2] = kfma(-lrc, __shfl_sync(FULL, lrc, c0 + k2), a[k2]);
            }
            a[k] = (lane == c) ? lcc : lrc;
        }
#pragma unroll
        for (int k = 0; k < CHOL_PANEL; k++) {
            const int c = c0 + k;
            if (lane >= c && lane < nvalid) {
                S[c * ldS + lane] = a[k];
                if (Lout) Lout[(size_t)c * ld + lane] = a[k];
            }
        }
        __syncwarp();
        if (progress && lane == 0) st_release_cta(progress, progress_base + c0 + CHOL_PANEL);
        // ---- trailing update of the block on the tensor cores: S(i,j) -= sum_k L(i,k) L(j,k) ----
        if (p < 3) {
            double af[4][2];  // fragments of L[8I .. 8I+7, panel]: A operand of row block I, B operand of column block I
#pragma unroll
            for (int I = 1; I < 4; I++) {
                af[I][0] = af[I][1] = 0.0;
                const int row = 8 * I + fr;
                if (I > p && (!MASK || row < nvalid)) {
                    af[I][0] = S[(c0 + fk) * ldS + row];
                    af[I][1] = S[(c0 + 4 + fk) * ldS + row];
                }
            }
#pragma unroll
            for (int J = 1; J < 4; J++) {
#pragma unroll
                for (int I = J; I < 4; I++) {
                    if (J > p) {  // warp-uniform
                        double u0 = 0.0, u1 = 0.0;
                        dmma(u0, u1, af[I][0], af[J][0]);
                        dmma(u0, u1, af[I][1], af[J][1]);
                        const int row = 8 * I + fr;
                        double* q = S + (8 * J + 2 * fk) * ldS + row;
                        if (!MASK || row < nvalid) {
                            q[0] -= u0;
                            q[ldS] -= u1;
                        }
                    }
                }
            }
            __syncwarp();
        }
    }
    double lg = lane < nvalid ? log(lmine) : 0.0;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) lg += __shfl_xor_sync(FULL, lg, o);
    logdet += lg;
    return ok;
}

// One 8-column panel (P = 0 .. 3, compile time: the trailing update's block structure is then straight-line code).
template <bool MASK, int P>
__device__ __forceinline__ void chol32_panel(double* S, int ldS, int lane, int nvalid, bool& ok, double& lmine,
                                             double* dinv_out, double* __restrict__ Lout, size_t ld, int* progress,
                                             int progress_base)
{
    constexpr unsigned FULL = 0xffffffffu;
    constexpr int c0 = CHOL_PANEL * P;
    const int fr = lane >> 2, fk = lane & 3;
    // ---- the panel: 8 columns of row `lane` ----
    double a[CHOL_PANEL];
#pragma unroll
    for (int k = 0; k < CHOL_PANEL; k++) {
        const int c = c0 + k;
        double v = (lane == c) ? 1.0 : 0.0;  // identity padding of a partial block; zeros above the diagonal
        if (lane < nvalid && c <= lane) v = S[c * ldS + lane];
        a[k] = v;
    }
    double d = __shfl_sync(FULL, a[0], c0), inv, l;
    chol_pivot(d, inv, l);
#pragma unroll
    for (int k = 0; k < CHOL_PANEL; k++) {
        const int c = c0 + k;
        if (!(d > 0.0)) ok = false;  // NaN or non-positive pivot; the NaNs that follow are discarded
        if (lane == c) lmine = l;
        if (lane == k) dinv_out[c] = inv;
        const double lrc = a[k] * inv;  // L(lane, c) below the diagonal
        const double lcc = l;
        if (k + 1 < CHOL_PANEL) {
            // The next pivot is known to lane c+1 before the column is passed round:
            // S(c+1,c+1) - L(c+1,c)^2, the same FMA the update below performs.  Its rsqrt
            // chain (latency-bound) overlaps the shuffles + FMAs of the rest of the panel.
            d = __shfl_sync(FULL, kfma(-lrc, lrc, a[k + 1]), c + 1);
            chol_pivot(d, inv, l);
#pragma unroll
            for (int k2 = k + 1; k2 < CHOL_PANEL; k2++)
                a[k2] = kfma(-lrc, __shfl_sync(FULL, lrc, c0 + k2), a[k2]);
        }
        a[k] = (lane == c) ? lcc : lrc;
    }
#pragma unroll
    for (int k = 0; k < CHOL_PANEL; k++) {
        const int c = c0 + k;
        if (lane >= c && lane < nvalid) {
            S[c * ldS + lane] = a[k];
            if (Lout) Lout[(size_t)c * ld + lane] = a[k];
        }
    }
    __syncwarp();
    if (progress && lane == 0) st_release_cta(progress, progress_base + c0 + CHOL_PANEL);
    // ---- trailing update of the block on the tensor cores: S(i,j) -= sum_k L(i,k) L(j,k) ----
    if (P < 3) {
        double af[4][2];  // fragments of L[8I .. 8I+7, panel]: A operand of row block I, B operand of column block I
#pragma unroll
        for (int I = P + 1; I < 4; I++) {
            af[I][0] = af[I][1] = 0.0;
            const int row = 8 * I + fr;
            if (!MASK || row < nvalid) {
                af[I][0] = S[(c0 + fk) * ldS + row];
                af[I][1] = S[(c0 + 4 + fk) * ldS + row];
            }
        }
#pragma unroll
        for (int J = P + 1; J < 4; J++) {
#pragma unroll
            for (int I = J; I < 4; I++) {
                double u0 = 0.0, u1 = 0.0;
                dmma(u0, u1, af[I][0], af[J][0]);
                dmma(u0, u1, af[I][1], af[J][1]);
                const int row = 8 * I + fr;
                double* q = S + (8 * J + 2 * fk) * ldS + row;
                if (!MASK || row < nvalid) {
                    q[0] -= u0;
                    q[ldS] -= u1;
                }
            }
        }
        __syncwarp();
    }
}

template <bool MASK>
__device__ __forceinline__ bool chol32_panels(double* S, int ldS, int lane, int nvalid, double& logdet,
                                              double* dinv_out, double* __restrict__ Lout, size_t ld,
                                              int* progress = nullptr, int progress_base = 0)
{
    constexpr unsigned FULL = 0xffffffffu;
    bool ok = true;
    double lmine = 1.0;  // lane c keeps L(c,c) for the log-determinant
    chol32_panel<MASK, 0>(S, ldS, lane, nvalid, ok, lmine, dinv_out, Lout, ld, progress, progress_base);
    chol32_panel<MASK, 1>(S, ldS, lane, nvalid, ok, lmine, dinv_out, Lout, ld, progress, progress_base);
    chol32_panel<MASK, 2>(S, ldS, lane, nvalid, ok, lmine, dinv_out, Lout, ld, progress, progress_base);
    chol32_panel<MASK, 3>(S, ldS, lane, nvalid, ok, lmine, dinv_out, Lout, ld, progress, progress_base);
    double lg = lane < nvalid ? log(lmine) : 0.0;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) lg += __shfl_xor_sync(FULL, lg, o);
    logdet += lg;
    return ok;
}

}  // namespace kc
