// Device-side proposals and accept logic (include/kima_sampler.h).
//
// The arithmetic of a lane's step is kc_sampler_logic.cuh (one thread per lane: the step is
// branchy scalar code over a few dozen parameters, the lanes of a warp diverge by design);
// this file is the kernels around it, the HBM-resident sampler state and the C ABI.
//
// State in HBM (T lanes, P particles per lane, n = T*P particles):
//   S     [n_state][ldp]   particle columns: the likelihood's kc_layout rows, then the RJObject's
//                          u_components, then the conditional prior's hyperparameters
//   np, logl, tb, level [n]
//   Q     [n_state][ld]    the lanes' proposals; rows [0, n_rows) are exactly the parameter block
//                          the likelihood kernels read (same ld as the context's own block), so a
//                          step scores them in place: no gather, no host
//   cnt   [4][Lcap][T] u32 each lane's increments of accepts / tries / visits / exceeds since the
//                          barrier (its private copy of the level counters, SamplerImpl.h:283-284)
//   above [T][max_steps][2], n_above [T]
// One MCMC step of all lanes = smp_propose (or the fused smp_accept_propose of the previous
// step) -> the likelihood kernels of the context over Q -> smp_accept.  kc_sampler_run enqueues
// `steps` of them back to back on the context stream; nothing returns to the host before the
// barrier (kc_sampler_collect).
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>
#include <cub/device/device_radix_sort.cuh>
#include "kc_ctx.h"
#include "kc_sampler_logic.cuh"

using kc::fail;
using kc::fail_cuda;
using kc::smp::SamplerParams;

#define CU(call)                                                  \
    do {                                                          \
        cudaError_t e_ = (call);                                  \
        if (e_ != cudaSuccess) return fail_cuda(e_, #call);       \
    } while (0)

namespace kc {

struct SamplerState {
    SamplerParams sp{};
    kc_sampler_config cfg{};
    kc_prior_table* d_pri = nullptr;
    double *d_lv_logl = nullptr, *d_lv_tb = nullptr, *d_lv_logX = nullptr;
    uint64_t* d_lv_tries = nullptr;
    // barrier outputs
    uint64_t* d_cnt_sum = nullptr;   // [4 * Lcap], level-major: 4*j + c
    int32_t* d_offsets = nullptr;    // [T + 1]
    double* d_above_compact = nullptr;
    uint64_t* h_cnt_sum = nullptr;   // pinned
    int32_t* h_total = nullptr;      // pinned
    double* h_above = nullptr;       // pinned, [T * max_steps * 2], allocated by the first kc_sampler_collect that returns pairs
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;  // around the last kc_sampler_run
    // `all_above` in HBM: values and tiebreakers (SoA), ping-pong buffers for the ordered cut
    double *pool_v[2] = {nullptr, nullptr}, *pool_t[2] = {nullptr, nullptr};
    uint64_t *pool_key[2] = {nullptr, nullptr};
    uint32_t *pool_idx[2] = {nullptr, nullptr};
    void* pool_tmp = nullptr;
    size_t pool_tmp_bytes = 0, pool_cap = 0, pool_n = 0;
    int max_levels = 0;               // of kc_sampler_create; sp.Lcap = rows of cnt allocated so far
    int pool_cur = 0;
    int pool_ranks = 1;                // processes whose lists meet in the pool (the communicator's size at creation)
    int32_t* d_rank_counts = nullptr;  // [pool_ranks]
    int32_t* h_rank_counts = nullptr;  // pinned
    double* h_pivot = nullptr;  // pinned, 2 doubles
    bool timed = false;
    std::vector<void*> allocs;
};

namespace {

constexpr int LANE_THREADS = 32;  // lanes per CTA: small CTAs spread a few thousand lanes over all SMs
// -DKC_SMP_MIN_CTAS=n compiles the step kernels for n resident CTAs per SM (register cap 64K / 32 / n);
// measured on B200 (profiles/smp_regcap_r01e.txt): 96 / 80 / 64 registers change nothing at 65 536 lanes, where
// every warp is resident anyway, and gain up to 10 % at 262 144 lanes; the default stays ptxas' own choice.
#ifdef KC_SMP_MIN_CTAS
#define KC_SMP_BOUNDS __launch_bounds__(LANE_THREADS, KC_SMP_MIN_CTAS)
#else
#define KC_SMP_BOUNDS __launch_bounds__(LANE_THREADS)
#endif

__global__ void KC_SMP_BOUNDS smp_propose_kernel(SamplerParams sp)
{
    const int lane = blockIdx.x * LANE_THREADS + threadIdx.x;
    if (lane < sp.lanes) smp::step_propose(sp, lane);
}
__global__ void KC_SMP_BOUNDS smp_accept_kernel(SamplerParams sp)
{
    const int lane = blockIdx.x * LANE_THREADS + threadIdx.x;
    if (lane < sp.lanes) smp::step_accept(sp, lane);
}
// the second half of step i and the first half of step i+1 of a lane in one launch: both touch
// only the lane's own particles, proposal column, counters and stream
__global__ void KC_SMP_BOUNDS smp_accept_propose_kernel(SamplerParams sp)
{
    const int lane = blockIdx.x * LANE_THREADS + threadIdx.x;
    if (lane < sp.lanes) {
        smp::step_accept(sp, lane);
        smp::step_propose(sp, lane);
    }
}

// Many lanes: CTAs of B lanes that sort their lanes by the move they are about to make (smp::Move, decided
// by the first draws of the proposal) between the two halves of the proposal, so that the threads of a warp
// run one branch of RVmodel::perturb instead of up to 32.  A lane's state is addressed by its index, so any
// thread of the CTA can finish it; which thread does is the only thing the (unordered) counting sort decides.
template <int B, bool ACCEPT>
__global__ void __launch_bounds__(B) smp_sorted_propose_kernel(SamplerParams sp)
{
    __shared__ int count[smp::MV_COUNT + 1], cursor[smp::MV_COUNT + 1];
    __shared__ unsigned short perm[B];
    __shared__ unsigned char moves[B];
    const int tid = threadIdx.x, lane = blockIdx.x * B + tid;
    if (tid <= smp::MV_COUNT) count[tid] = 0;
    __syncthreads();
    int move = smp::MV_COUNT;  // past the end: sorts behind the live lanes
    if (lane < sp.lanes) {
        if (ACCEPT) smp::step_accept(sp, lane);
        move = smp::step_propose_begin(sp, lane);
    }
    moves[tid] = (unsigned char)move;
    atomicAdd(&count[move], 1);
    __syncthreads();
    if (tid == 0) {
        int o = 0;
        for (int k = 0; k <= smp::MV_COUNT; k++) { cursor[k] = o; o += count[k]; }
    }
    __syncthreads();
    perm[atomicAdd(&cursor[move], 1)] = (unsigned short)tid;
    __syncthreads();  // also orders the lanes' global writes of the first half before the second half reads them
    const int t2 = perm[tid], lane2 = blockIdx.x * B + t2;
    if (lane2 < sp.lanes) smp::step_propose_finish(sp, lane2, moves[t2]);
}

// CTA size of the move-sorted kernels for this many lanes; KC_SMP_BLOCK (0, 64, 128, 256) overrides
int sort_block(const SamplerParams& sp)
{
    if (const char* e = std::getenv("KC_SMP_BLOCK")) {
        const int b = std::atoi(e);
        return (b == 64 || b == 128 || b == 256) ? b : 0;
    }
    // measured (profiles/smp_sorted_r01e.txt, 51Peg shape): the step kernel 0.158 -> 0.070 ms at 65 536 lanes and
    // 0.65 -> 0.30 ms at 262 144 with B = 256 (128: 0.090, 64: 0.092); 0.078 -> 0.061 ms at 16 384
    // the step is latency-bound, not SM-bound: even 2048 lanes (8 CTAs) gain, 0.101 -> 0.085 ms per step
    return sp.lanes >= 2048 ? 256 : 0;
}
template <bool ACCEPT>
void launch_propose_t(const SamplerParams& sp, int B, int grid32, cudaStream_t st)
{
    const int g = B ? (sp.lanes + B - 1) / B : grid32;
    if (B == 256) smp_sorted_propose_kernel<256, ACCEPT><<<g, 256, 0, st>>>(sp);
    else if (B == 128) smp_sorted_propose_kernel<128, ACCEPT><<<g, 128, 0, st>>>(sp);
    else if (B == 64) smp_sorted_propose_kernel<64, ACCEPT><<<g, 64, 0, st>>>(sp);
    else if (ACCEPT) smp_accept_propose_kernel<<<grid32, LANE_THREADS, 0, st>>>(sp);
    else smp_propose_kernel<<<grid32, LANE_THREADS, 0, st>>>(sp);
}
void launch_propose(const SamplerParams& sp, int B, bool accept, int grid32, cudaStream_t st)
{
    if (accept) launch_propose_t<true>(sp, B, grid32, st);
    else launch_propose_t<false>(sp, B, grid32, st);
}

// sum over lanes of one (counter, level) row of cnt: one CTA per row, coalesced
__global__ void __launch_bounds__(256) smp_reduce_counters_kernel(SamplerParams sp, uint64_t* out)
{
    const int j = blockIdx.x, c = blockIdx.y;
    const uint32_t* row = sp.cnt + ((size_t)c * sp.Lcap + j) * sp.ldl;
    uint64_t s = 0;
    for (int k = threadIdx.x; k < sp.lanes; k += 256) s += row[k];
    __shared__ uint64_t sh[8];
    for (int o = 16; o > 0; o >>= 1) s += __shfl_down_sync(0xffffffffu, s, o);
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = s;
    __syncthreads();
    if (threadIdx.x == 0) {
        uint64_t t = 0;
        for (int w = 0; w < 8; w++) t += sh[w];
        out[4 * j + c] = t;
    }
}

// exclusive scan of n_above over lanes (one CTA; T is at most a few 10^5)
__global__ void __launch_bounds__(1024) smp_scan_above_kernel(const int32_t* n_above, int T, int32_t* offsets)
{
    __shared__ int32_t sh[1024];
    __shared__ int32_t carry;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    for (int base = 0; base < T; base += 1024) {
        const int k = base + threadIdx.x;
        const int32_t v = k < T ? n_above[k] : 0;
        sh[threadIdx.x] = v;
        __syncthreads();
        for (int o = 1; o < 1024; o <<= 1) {
            const int32_t add = threadIdx.x >= o ? sh[threadIdx.x - o] : 0;
            __syncthreads();
            sh[threadIdx.x] += add;
            __syncthreads();
        }
        if (k < T) offsets[k] = carry + sh[threadIdx.x] - v;
        __syncthreads();
        if (threadIdx.x == 1023) carry += sh[1023];
        __syncthreads();
    }
    if (threadIdx.x == 0) offsets[T] = carry;
}
__global__ void smp_gather_above_kernel(SamplerParams sp, const int32_t* offsets, double* out)
{
    const int lane = blockIdx.x * blockDim.x + threadIdx.x;
    if (lane >= sp.lanes) return;
    const int n = sp.n_above[lane];
    const double* a = sp.above + (size_t)lane * sp.max_steps * 2;
    double* o = out + (size_t)offsets[lane] * 2;
    for (int i = 0; i < 2 * n; i++) o[i] = a[i];
}

// the lanes' `above` lists appended to the pool (SoA) at `base`, in lane order
__global__ void smp_gather_pool_kernel(SamplerParams sp, const int32_t* offsets, double* pv, double* pt, size_t base)
{
    const int lane = blockIdx.x * blockDim.x + threadIdx.x;
    if (lane >= sp.lanes) return;
    const int n = sp.n_above[lane];
    const double* a = sp.above + (size_t)lane * sp.max_steps * 2;
    const size_t o = base + (size_t)offsets[lane];
    for (int i = 0; i < n; i++) {
        pv[o + i] = a[2 * i];
        pt[o + i] = a[2 * i + 1];
    }
}
// the ranks' staged lists (rank r: values at recv[r*2M .. +counts[r]), tiebreakers M further) appended to
// the pool in rank order = the order in which one process would have concatenated its threads' lists
__global__ void smp_pool_append_ranks_kernel(const double* recv, const int32_t* counts, int R, size_t M, double* pv,
                                             double* pt, size_t base)
{
    const int r = blockIdx.y;
    size_t o = base;
    for (int k = 0; k < r; k++) o += (size_t)counts[k];
    const size_t n = (size_t)counts[r];
    const double* src = recv + (size_t)r * 2 * M;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        pv[o + i] = src[i];
        pt[o + i] = src[M + i];
    }
}
// order-preserving map double -> uint64 (radix sort key)
__device__ __forceinline__ uint64_t ordered_key(double x)
{
    const uint64_t b = (uint64_t)__double_as_longlong(x);
    return (b >> 63) ? ~b : (b | 0x8000000000000000ull);
}
__global__ void smp_pool_keys_kernel(const double* src, const uint32_t* idx, size_t n, uint64_t* key, uint32_t* iota)
{
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    key[i] = ordered_key(idx ? src[idx[i]] : src[i]);
    if (iota) iota[i] = (uint32_t)i;
}
// entries of sorted rank > index, in order, into the other buffers; the pivot to `pivot`
__global__ void smp_pool_cut_kernel(const double* v, const double* t, const uint32_t* perm, size_t n, size_t index,
                                    double* vo, double* to, double* pivot)
{
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint32_t src = perm[i];
    if (i == index) {
        pivot[0] = v[src];
        pivot[1] = t[src];
    } else if (i > index) {
        vo[i - index - 1] = v[src];
        to[i - index - 1] = t[src];
    }
}
__global__ void smp_pool_pairs_kernel(const double* v, const double* t, size_t n, double* out)
{
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    out[2 * i] = v[i];
    out[2 * i + 1] = t[i];
}

__global__ void smp_copy_particles_kernel(SamplerParams sp, int n, const int32_t* dst, const int32_t* src)
{
    const int i = blockIdx.x;
    if (i >= n) return;
    const int d = dst[i], s = src[i];
    for (int r = threadIdx.x; r < sp.SL.n_state; r += blockDim.x) sp.S[(size_t)r * sp.ldp + d] = sp.S[(size_t)r * sp.ldp + s];
    if (threadIdx.x == 0) {
        sp.np[d] = sp.np[s];
        sp.logl[d] = sp.logl[s];
        sp.tb[d] = sp.tb[s];
        sp.level[d] = sp.level[s];
    }
}

// ---- test hooks ----
__global__ void smp_prior_eval_kernel(kc_prior q, int what, int n, const double* x, double* out)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    out[i] = what == 0 ? smp::prior_cdf(q, x[i]) : (what == 1 ? smp::prior_cdf_inverse(q, x[i]) : smp::prior_log_pdf(q, x[i]));
}
__global__ void smp_rng_draws_kernel(uint64_t key, uint64_t ctr0, int what, int n, double* out, uint64_t* ctr_end)
{
    smp::Rng r{key, ctr0};
    for (int i = 0; i < n; i++)
        out[i] = what == 0 ? r.rand() : (what == 1 ? r.randn() : (what == 2 ? r.randt2() : r.randh()));
    *ctr_end = r.ctr;
}
__global__ void __launch_bounds__(LANE_THREADS) smp_test_perturb_kernel(SamplerParams sp, double* logH)
{
    const int lane = blockIdx.x * LANE_THREADS + threadIdx.x;
    if (lane >= sp.lanes) return;
    smp::Rng rng{sp.key0 + (uint64_t)lane, sp.ctr[lane]};
    const smp::Col q{sp.Q + lane, sp.ldq};
    for (int r = 0; r < sp.SL.n_state; r++) q(r) = sp.S[(size_t)r * sp.ldp + lane];
    int n = sp.np[lane];
    logH[lane] = smp::perturb_model(sp, q, n, rng);
    sp.npq[lane] = n;
    sp.ctr[lane] = rng.ctr;
}

template <class T>
cudaError_t dalloc(SamplerState* s, T** p, size_t count)
{
    void* v = nullptr;
    cudaError_t e = cudaMalloc(&v, sizeof(T) * (count ? count : 1));
    if (e != cudaSuccess) return e;
    e = cudaMemset(v, 0, sizeof(T) * (count ? count : 1));
    s->allocs.push_back(v);
    *p = (T*)v;
    return e;
}

int lane_grid(const SamplerParams& sp) { return (sp.lanes + LANE_THREADS - 1) / LANE_THREADS; }

int need_sampler(kc_ctx* c)
{
    if (!c) return fail(KC_ERR_ARG, "null context");
    if (!c->sampler) return fail(KC_ERR_ARG, "kc_sampler_create has not been called");
    return KC_OK;
}

}  // namespace

void sampler_free(kc_ctx* c)
{
    SamplerState* s = c->sampler;
    if (!s) return;
    cudaStreamSynchronize(c->stream);
    for (void* p : s->allocs) cudaFree(p);
    if (s->h_cnt_sum) cudaFreeHost(s->h_cnt_sum);
    if (s->h_total) cudaFreeHost(s->h_total);
    if (s->h_rank_counts) cudaFreeHost(s->h_rank_counts);
    if (s->h_above) cudaFreeHost(s->h_above);
    if (s->h_pivot) cudaFreeHost(s->h_pivot);
    if (s->ev0) cudaEventDestroy(s->ev0);
    if (s->ev1) cudaEventDestroy(s->ev1);
    delete s;
    c->sampler = nullptr;
}

}  // namespace kc

using kc::SamplerState;

extern "C" {

int kc_prior_make(int type, const double* p, int n_p, kc_prior* out)
{
    if (!p || !out) return fail(KC_ERR_ARG, "null argument");
    if (kc::smp::prior_make(type, p, n_p, out))
        return fail(KC_ERR_ARG, "invalid prior: unknown type, wrong argument count, or arguments the reference's constructor rejects");
    return KC_OK;
}

static int sampler_create_impl(kc_ctx* c, const kc_prior_table* priors, const kc_sampler_config* cfg);

int kc_sampler_create(kc_ctx* c, const kc_prior_table* priors, const kc_sampler_config* cfg)
{
    const int rc = sampler_create_impl(c, priors, cfg);
    if (rc != KC_OK && c) kc::sampler_free(c);  // no half-built sampler behind a failed call
    if (rc == KC_OK && c->gp_cache_opt) {
        // one factor-cache slot per particle (kima_cuda.h: kc_gp_cache_enable); a cache that does not fit the
        // device memory, or a model it does not apply to, leaves the sampler as it is
        (void)kc_gp_cache_enable(c, cfg->lanes * cfg->num_particles);
    }
    return rc;
}

static int sampler_create_impl(kc_ctx* c, const kc_prior_table* priors, const kc_sampler_config* cfg)
{
    if (!c || !priors || !cfg) return fail(KC_ERR_ARG, "null argument");
    int rc = kc::check_ready(c, 0);
    if (rc) return rc;
    if (cfg->lanes < 1 || cfg->lanes > c->max_batch) return fail(KC_ERR_ARG, "lanes outside [1, max_batch of kc_configure]");
    if (cfg->num_particles < 1) return fail(KC_ERR_ARG, "num_particles < 1");
    if (cfg->max_steps < 1) return fail(KC_ERR_ARG, "max_steps < 1");
    if (cfg->max_levels < 1) return fail(KC_ERR_ARG, "max_levels < 1");
    if (!(cfg->lambda > 0.0)) return fail(KC_ERR_ARG, "lambda must be positive");
    // every prior the flags make perturb() touch must be there
    {
        const kc_prior_table& T = *priors;
        const kc_model_flags& F = c->F;
        auto have = [](const kc_prior& q) { return q.type >= 0 && q.type < KC_PRIOR_N_TYPES; };
        bool ok = have(T.Cprior) && have(T.Jprior);
        if (F.trend) ok = ok && have(T.slope_prior) && (F.degree < 2 || have(T.quadr_prior)) && (F.degree < 3 || have(T.cubic_prior));
        if (c->L.n_off) ok = ok && have(T.offsets_prior);
        if (c->L.n_beta) ok = ok && have(T.betaprior);
        if (F.MA) ok = ok && have(T.sigmaMA_prior) && have(T.tauMA_prior);
        if (F.GP) ok = ok && have(T.log_eta1_prior) && have(T.eta2_prior) && have(T.eta3_prior) && (F.kernel != 0 || have(T.log_eta4_prior));
        if (F.studentt) ok = ok && have(T.nu_prior);
        for (int i = 0; i < c->L.n_ko; i++)
            ok = ok && have(T.KO_Pprior[i]) && have(T.KO_Kprior[i]) && have(T.KO_eprior[i]) && have(T.KO_phiprior[i]) && have(T.KO_wprior[i]);
        if (F.npmax > 0) {
            ok = ok && have(T.eprior) && have(T.phiprior) && have(T.wprior);
            if (F.hyperpriors) ok = ok && have(T.log_muP_prior) && have(T.wP_prior) && have(T.log_muK_prior);
            else ok = ok && have(T.Pprior) && have(T.Kprior);
        }
        if (!ok) return fail(KC_ERR_ARG, "prior table: a prior the model flags need is missing (type KC_PRIOR_NONE)");
    }
    CU(cudaSetDevice(c->device));
    kc::sampler_free(c);
    SamplerState* s = new SamplerState();
    c->sampler = s;
    s->cfg = *cfg;
    SamplerParams& sp = s->sp;
    const int T = cfg->lanes, P = cfg->num_particles;
    const size_t n = (size_t)T * P;
    sp.F = c->F;
    sp.L = c->L;
    sp.SL.row_u = c->L.n_rows;
    sp.SL.row_hyper = c->L.n_rows + 5 * c->L.npmax;
    sp.SL.n_state = sp.SL.row_hyper + 3;
    sp.SL.n_particles = (int)n;
    sp.n_ind = c->n_ind;
    sp.ldp = (n + 3) & ~(size_t)3;
    sp.ldq = (size_t)c->ld;
    sp.ldl = ((size_t)T + 3) & ~(size_t)3;
    sp.lanes = T;
    sp.P = P;
    sp.max_steps = cfg->max_steps;
    // the per-lane counter rows grow with the level table (kc_sampler_set_levels): 4 x levels x lanes u32 for
    // max_levels = 1024 and 65 536 lanes would be 1 GB that a typical run (a few dozen levels) never touches
    s->max_levels = cfg->max_levels;
    sp.Lcap = cfg->max_levels < 64 ? cfg->max_levels : 64;
    sp.lambda = cfg->lambda;
    sp.beta = cfg->beta;
    sp.key0 = cfg->seed + (uint64_t)cfg->lane0;
    sp.n_levels = 0;
#define DA(ptr, count)                                                           \
    do {                                                                         \
        cudaError_t e_ = kc::dalloc(s, &(ptr), (count));                         \
        if (e_ != cudaSuccess) { kc::sampler_free(c); return fail_cuda(e_, "sampler allocation"); } \
    } while (0)
    const auto t_0 = std::chrono::steady_clock::now();
    DA(s->d_pri, 1);
    DA(sp.S, (size_t)sp.SL.n_state * sp.ldp);
    DA(sp.np, sp.ldp);
    DA(sp.logl, sp.ldp);
    DA(sp.tb, sp.ldp);
    DA(sp.level, sp.ldp);
    DA(sp.Q, (size_t)sp.SL.n_state * sp.ldq);
    DA(sp.npq, sp.ldq);
    double* lq = nullptr;
    DA(lq, sp.ldq);
    sp.logl_q = lq;
    DA(sp.ctr, (size_t)T);
    DA(sp.which, (size_t)T);
    DA(sp.flags, (size_t)T);
    DA(sp.above, (size_t)T * sp.max_steps * 2);
    DA(sp.n_above, (size_t)T);
    DA(s->d_lv_logl, (size_t)s->max_levels);
    DA(s->d_lv_tb, (size_t)s->max_levels);
    DA(s->d_lv_logX, (size_t)s->max_levels);
    DA(s->d_lv_tries, (size_t)s->max_levels);
    DA(sp.cnt, (size_t)4 * sp.Lcap * sp.ldl);
    DA(s->d_cnt_sum, (size_t)4 * s->max_levels);
    DA(s->d_offsets, (size_t)T + 1);
    DA(s->d_above_compact, (size_t)T * sp.max_steps * 2);
#undef DA
    sp.pri = s->d_pri;
    sp.lv_logl = s->d_lv_logl;
    sp.lv_tb = s->d_lv_tb;
    sp.lv_logX = s->d_lv_logX;
    sp.lv_tries = s->d_lv_tries;
    CU(cudaMemcpy(s->d_pri, priors, sizeof(kc_prior_table), cudaMemcpyHostToDevice));
    CU(cudaMallocHost(&s->h_cnt_sum, sizeof(uint64_t) * 4 * s->max_levels));
    CU(cudaMallocHost(&s->h_total, sizeof(int32_t)));
    CU(cudaEventCreate(&s->ev0));
    CU(cudaEventCreate(&s->ev1));
    const auto t_1 = std::chrono::steady_clock::now();
    if (cfg->pool_reserve > 0) {
        // leftovers of a cut are at most 1/compression of the pool; without a cut the pool stops
        // growing at new_level_interval + one barrier's worth
        // with a communicator every rank keeps the whole run's pool (the lists of all ranks, in rank order)
        s->pool_ranks = kc::comm_ranks(c);
        s->pool_cap = (size_t)3 * T * sp.max_steps * s->pool_ranks + (size_t)cfg->pool_reserve;
        if (kc::dalloc(s, &s->d_rank_counts, (size_t)s->pool_ranks) != cudaSuccess) {
            kc::sampler_free(c);
            return fail(KC_ERR_CUDA, "sampler allocation (rank counts)");
        }
        CU(cudaMallocHost(&s->h_rank_counts, sizeof(int32_t) * s->pool_ranks));
        for (int k = 0; k < 2; k++) {
            cudaError_t e1 = kc::dalloc(s, &s->pool_v[k], s->pool_cap), e2 = kc::dalloc(s, &s->pool_t[k], s->pool_cap);
            cudaError_t e3 = kc::dalloc(s, &s->pool_key[k], s->pool_cap), e4 = kc::dalloc(s, &s->pool_idx[k], s->pool_cap);
            if (e1 != cudaSuccess || e2 != cudaSuccess || e3 != cudaSuccess || e4 != cudaSuccess) {
                kc::sampler_free(c);
                return fail(KC_ERR_CUDA, "sampler allocation (above pool)");
            }
        }
        size_t bytes = 0;
        CU(cub::DeviceRadixSort::SortPairs(nullptr, bytes, s->pool_key[0], s->pool_key[1], s->pool_idx[0], s->pool_idx[1],
                                           (int64_t)s->pool_cap, 0, 64, c->stream));
        s->pool_tmp_bytes = bytes;
        cudaError_t e = kc::dalloc(s, (char**)&s->pool_tmp, bytes);
        if (e != cudaSuccess) { kc::sampler_free(c); return fail_cuda(e, "sampler allocation (sort scratch)"); }
        CU(cudaMallocHost(&s->h_pivot, 2 * sizeof(double)));
    }
    // the perturb call tree keeps its component vectors on the local stack
    size_t stack = 0;
    CU(cudaDeviceGetLimit(&stack, cudaLimitStackSize));
    const auto t_2 = std::chrono::steady_clock::now();
    if (stack < 4096) CU(cudaDeviceSetLimit(cudaLimitStackSize, 4096));
    if (std::getenv("KIMA_PROFILE")) {
        CU(cudaDeviceSynchronize());
        const auto t_3 = std::chrono::steady_clock::now();
        auto sec = [](std::chrono::steady_clock::time_point a, std::chrono::steady_clock::time_point b) {
            return std::chrono::duration<double>(b - a).count();
        };
        fprintf(stderr, "# kc_sampler_create: state arrays %.4f s, above pool + sort scratch %.4f s, stack limit %.4f s\n",
                sec(t_0, t_1), sec(t_1, t_2), sec(t_2, t_3));
    }
    return KC_OK;
}

int kc_sampler_destroy(kc_ctx* c)
{
    if (!c) return fail(KC_ERR_ARG, "null context");
    CU(cudaSetDevice(c->device));
    kc::sampler_free(c);
    return KC_OK;
}

int kc_sampler_get_layout(kc_ctx* c, kc_sampler_layout* out)
{
    int rc = kc::need_sampler(c);
    if (rc) return rc;
    if (!out) return fail(KC_ERR_ARG, "out is null");
    *out = c->sampler->sp.SL;
    return KC_OK;
}

static int state_copy(kc_ctx* c, bool up, double* state, int32_t* np, double* logl, double* tb, int32_t* level)
{
    int rc = kc::need_sampler(c);
    if (rc) return rc;
    CU(cudaSetDevice(c->device));
    const SamplerParams& sp = c->sampler->sp;
    const size_t n = (size_t)sp.SL.n_particles;
    const cudaMemcpyKind kind = up ? cudaMemcpyHostToDevice : cudaMemcpyDeviceToHost;
    if (state) {
        if (up) CU(cudaMemcpy2DAsync(sp.S, sizeof(double) * sp.ldp, state, sizeof(double) * n, sizeof(double) * n, sp.SL.n_state, kind, c->stream));
        else CU(cudaMemcpy2DAsync(state, sizeof(double) * n, sp.S, sizeof(double) * sp.ldp, sizeof(double) * n, sp.SL.n_state, kind, c->stream));
    }
    auto cp = [&](void* h, void* d, size_t bytes) { return up ? cudaMemcpyAsync(d, h, bytes, kind, c->stream) : cudaMemcpyAsync(h, d, bytes, kind, c->stream); };
    if (np) CU(cp(np, sp.np, sizeof(int32_t) * n));
    if (logl) CU(cp(logl, sp.logl, sizeof(double) * n));
    if (tb) CU(cp(tb, sp.tb, sizeof(double) * n));
    if (level) CU(cp(level, sp.level, sizeof(int32_t) * n));
    CU(cudaStreamSynchronize(c->stream));
    return KC_OK;
}

int kc_sampler_upload(kc_ctx* c, const double* state, const int32_t* np, const double* logl,
                      const double* tiebreaker, const int32_t* level)
{
    return state_copy(c, true, const_cast<double*>(state), const_cast<int32_t*>(np), const_cast<double*>(logl),
                      const_cast<double*>(tiebreaker), const_cast<int32_t*>(level));
}
int kc_sampler_download(kc_ctx* c, double* state, int32_t* np, double* logl, double* tiebreaker, int32_t* level)
{
    return state_copy(c, false, state, np, logl, tiebreaker, level);
}

int kc_sampler_get_particle(kc_ctx* c, int which, double* state_col, int32_t* np)
{
    int rc = kc::need_sampler(c);
    if (rc) return rc;
    const SamplerParams& sp = c->sampler->sp;
    if (which < 0 || which >= sp.SL.n_particles) return fail(KC_ERR_ARG, "particle index out of range");
    CU(cudaSetDevice(c->device));
    if (state_col)
        CU(cudaMemcpy2DAsync(state_col, sizeof(double), sp.S + which, sizeof(double) * sp.ldp, sizeof(double),
                             sp.SL.n_state, cudaMemcpyDeviceToHost, c->stream));
    if (np) CU(cudaMemcpyAsync(np, sp.np + which, sizeof(int32_t), cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    return KC_OK;
}

int kc_sampler_set_particle(kc_ctx* c, int which, const double* state_col, int32_t np, double logl,
                            double tiebreaker, int32_t level)
{
    int rc = kc::need_sampler(c);
    if (rc) return rc;
    const SamplerParams& sp = c->sampler->sp;
    if (which < 0 || which >= sp.SL.n_particles) return fail(KC_ERR_ARG, "particle index out of range");
    if (!state_col) return fail(KC_ERR_ARG, "state_col is null");
    CU(cudaSetDevice(c->device));
    CU(cudaMemcpy2DAsync(sp.S + which, sizeof(double) * sp.ldp, state_col, sizeof(double), sizeof(double),
                         sp.SL.n_state, cudaMemcpyHostToDevice, c->stream));
    CU(cudaMemcpyAsync(sp.np + which, &np, sizeof(int32_t), cudaMemcpyHostToDevice, c->stream));
    CU(cudaMemcpyAsync(sp.logl + which, &logl, sizeof(double), cudaMemcpyHostToDevice, c->stream));
    CU(cudaMemcpyAsync(sp.tb + which, &tiebreaker, sizeof(double), cudaMemcpyHostToDevice, c->stream));
    CU(cudaMemcpyAsync(sp.level + which, &level, sizeof(int32_t), cudaMemcpyHostToDevice, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    return KC_OK;
}

int kc_sampler_copy_particles(kc_ctx* c, int n, const int32_t* dst, const int32_t* src)
{
    int rc = kc::need_sampler(c);
    if (rc) return rc;
    if (n == 0) return KC_OK;
    if (n < 0 || !dst || !src) return fail(KC_ERR_ARG, "bad copy list");
    const SamplerParams& sp = c->sampler->sp;
    for (int i = 0; i < n; i++)
        if (dst[i] < 0 || dst[i] >= sp.SL.n_particles || src[i] < 0 || src[i] >= sp.SL.n_particles)
            return fail(KC_ERR_ARG, "particle index out of range");
    CU(cudaSetDevice(c->device));
    int32_t* d = nullptr;
    CU(cudaMalloc(&d, sizeof(int32_t) * 2 * n));
    CU(cudaMemcpyAsync(d, dst, sizeof(int32_t) * n, cudaMemcpyHostToDevice, c->stream));
    CU(cudaMemcpyAsync(d + n, src, sizeof(int32_t) * n, cudaMemcpyHostToDevice, c->stream));
    kc::smp_copy_particles_kernel<<<n, 64, 0, c->stream>>>(sp, n, d, d + n);
    c->launches += 1;
    CU(cudaGetLastError());
    CU(cudaStreamSynchronize(c->stream));
    CU(cudaFree(d));
    return KC_OK;
}

int kc_sampler_set_levels(kc_ctx* c, int L, const double* logl, const double* tiebreaker, const double* log_X,
                          const uint64_t* tries, int enough, int at_max)
{
    int rc = kc::need_sampler(c);
    if (rc) return rc;
    SamplerState* s = c->sampler;
    SamplerParams& sp = s->sp;
    if (L < 1 || L > s->max_levels) return fail(KC_ERR_ARG, "number of levels outside [1, max_levels of kc_sampler_create]");
    if (!logl || !tiebreaker || !log_X || !tries) return fail(KC_ERR_ARG, "null level array");
    CU(cudaSetDevice(c->device));
    if (L > sp.Lcap) {
        // the rows only hold the increments since the last barrier and are cleared below: nothing to carry over
        int cap = 2 * sp.Lcap > L ? 2 * sp.Lcap : L;
        if (cap > s->max_levels) cap = s->max_levels;
        CU(cudaStreamSynchronize(c->stream));
        uint32_t* grown = nullptr;
        CU(cudaMalloc(&grown, sizeof(uint32_t) * 4 * (size_t)cap * sp.ldl));
        for (void*& p : s->allocs)
            if (p == (void*)sp.cnt) p = grown;
        CU(cudaFree(sp.cnt));
        sp.cnt = grown;
        sp.Lcap = cap;
    }
    // pageable sources: the copies are staged before the call returns
    CU(cudaMemcpyAsync(s->d_lv_logl, logl, sizeof(double) * L, cudaMemcpyHostToDevice, c->stream));
    CU(cudaMemcpyAsync(s->d_lv_tb, tiebreaker, sizeof(double) * L, cudaMemcpyHostToDevice, c->stream));
    CU(cudaMemcpyAsync(s->d_lv_logX, log_X, sizeof(double) * L, cudaMemcpyHostToDevice, c->stream));
    CU(cudaMemcpyAsync(s->d_lv_tries, tries, sizeof(uint64_t) * L, cudaMemcpyHostToDevice, c->stream));
    for (int k = 0; k < 4; k++)
        CU(cudaMemsetAsync(sp.cnt + (size_t)k * sp.Lcap * sp.ldl, 0, sizeof(uint32_t) * (size_t)L * sp.ldl, c->stream));
    CU(cudaMemsetAsync(sp.n_above, 0, sizeof(int32_t) * sp.lanes, c->stream));
    sp.n_levels = L;
    sp.enough = enough ? 1 : 0;
    sp.at_max = at_max ? 1 : 0;
    return KC_OK;
}

int kc_sampler_run(kc_ctx* c, int steps)
{
    int rc = kc::need_sampler(c);
    if (rc) return rc;
    SamplerState* s = c->sampler;
    const SamplerParams& sp = s->sp;
    if (steps < 0 || steps > sp.max_steps) return fail(KC_ERR_ARG, "steps outside [0, max_steps of kc_sampler_create]");
    if (sp.n_levels < 1) return fail(KC_ERR_ARG, "kc_sampler_set_levels has not been called");
    if (c->pending_out) return fail(KC_ERR_ARG, "an asynchronous evaluation is in flight");
    if (steps == 0) return KC_OK;
    CU(cudaSetDevice(c->device));
    const int grid = kc::lane_grid(sp);
    CU(cudaEventRecord(s->ev0, c->stream));
    const int B = kc::sort_block(sp);  // 0: too few lanes to fill the SMs with larger CTAs, one warp per CTA
    kc::launch_propose(sp, B, false, grid, c->stream);
    c->launches += 1;
    for (int i = 0; i < steps; i++) {
        // factor cache (kc_gp_cache_enable with >= n_particles slots): the slot of a lane is the particle it moves
        rc = kc::run_kernels(c, sp.lanes, const_cast<double*>(sp.logl_q), sp.Q, sp.npq, true,
                             c->gpc_nslots >= sp.SL.n_particles ? sp.which : nullptr);
        if (rc) return rc;
        if (i + 1 < steps) kc::launch_propose(sp, B, true, grid, c->stream);
        else kc::smp_accept_kernel<<<grid, kc::LANE_THREADS, 0, c->stream>>>(sp);
        c->launches += 1;
    }
    CU(cudaEventRecord(s->ev1, c->stream));
    s->timed = true;
    CU(cudaGetLastError());
    return KC_OK;
}

int kc_sampler_last_run_ms(kc_ctx* c, float* ms)
{
    int rc = kc::need_sampler(c);
    if (rc) return rc;
    if (!ms) return fail(KC_ERR_ARG, "ms is null");
    if (!c->sampler->timed) return fail(KC_ERR_ARG, "no kc_sampler_run has been timed yet");
    CU(cudaSetDevice(c->device));
    CU(cudaEventSynchronize(c->sampler->ev1));
    CU(cudaEventElapsedTime(ms, c->sampler->ev0, c->sampler->ev1));
    return KC_OK;
}

int kc_sampler_collect(kc_ctx* c, uint64_t* counters, double* above_pairs, int max_pairs, int* n_above)
{
    int rc = kc::need_sampler(c);
    if (rc) return rc;
    SamplerState* s = c->sampler;
    const SamplerParams& sp = s->sp;
    if (sp.n_levels < 1) return fail(KC_ERR_ARG, "kc_sampler_set_levels has not been called");
    CU(cudaSetDevice(c->device));
    const int L = sp.n_levels, T = sp.lanes;
    if (counters) {
        kc::smp_reduce_counters_kernel<<<dim3(L, 4), 256, 0, c->stream>>>(sp, s->d_cnt_sum);
        CU(cudaMemcpyAsync(s->h_cnt_sum, s->d_cnt_sum, sizeof(uint64_t) * 4 * L, cudaMemcpyDeviceToHost, c->stream));
        c->launches += 1;
    }
    int total = 0;
    if (above_pairs || n_above) {
        kc::smp_scan_above_kernel<<<1, 1024, 0, c->stream>>>(sp.n_above, T, s->d_offsets);
        kc::smp_gather_above_kernel<<<(T + 127) / 128, 128, 0, c->stream>>>(sp, s->d_offsets, s->d_above_compact);
        c->launches += 2;
        CU(cudaMemcpyAsync(s->h_total, s->d_offsets + T, sizeof(int32_t), cudaMemcpyDeviceToHost, c->stream));
        CU(cudaStreamSynchronize(c->stream));
        total = *s->h_total;
        const int take = total < max_pairs ? total : max_pairs;
        if (above_pairs && take > 0) {
            // pinned staging for the pairs: only runs that keep all_above on the host pay for it
            if (!s->h_above) CU(cudaMallocHost(&s->h_above, sizeof(double) * (size_t)T * sp.max_steps * 2));
            CU(cudaMemcpyAsync(s->h_above, s->d_above_compact, sizeof(double) * 2 * take, cudaMemcpyDeviceToHost, c->stream));
            CU(cudaStreamSynchronize(c->stream));
            std::memcpy(above_pairs, s->h_above, sizeof(double) * 2 * take);
        }
    } else {
        CU(cudaStreamSynchronize(c->stream));
    }
    CU(cudaGetLastError());
    if (counters) std::memcpy(counters, s->h_cnt_sum, sizeof(uint64_t) * 4 * L);
    if (n_above) *n_above = total;
    return KC_OK;
}

int kc_sampler_collect_pool(kc_ctx* c, uint64_t* counters, int64_t* n_new, int64_t* n_pool)
{
    int rc = kc::need_sampler(c);
    if (rc) return rc;
    SamplerState* s = c->sampler;
    const SamplerParams& sp = s->sp;
    if (!s->pool_cap) return fail(KC_ERR_ARG, "the sampler was created without an above pool (pool_reserve = 0)");
    if (sp.n_levels < 1) return fail(KC_ERR_ARG, "kc_sampler_set_levels has not been called");
    CU(cudaSetDevice(c->device));
    const int L = sp.n_levels, T = sp.lanes;
    if (counters) {
        kc::smp_reduce_counters_kernel<<<dim3(L, 4), 256, 0, c->stream>>>(sp, s->d_cnt_sum);
        c->launches += 1;
    }
    const int R = s->pool_ranks;
    const size_t M = (size_t)T * sp.max_steps;
    // capacity is checked before the append lands: one barrier adds at most R * T * max_steps
    if (s->pool_n + M * R > s->pool_cap)
        return fail(KC_ERR_ARG, "above pool full: place a level (kc_sampler_pool_select) or clear it");
    if (counters && R > 1) {
        int e = kc::comm_allreduce_u64_dev(c, s->d_cnt_sum, (size_t)4 * L);  // every rank folds the same sums in
        if (e) return e;
    }
    if (counters)
        CU(cudaMemcpyAsync(s->h_cnt_sum, s->d_cnt_sum, sizeof(uint64_t) * 4 * L, cudaMemcpyDeviceToHost, c->stream));
    kc::smp_scan_above_kernel<<<1, 1024, 0, c->stream>>>(sp.n_above, T, s->d_offsets);
    size_t added = 0;
    if (R == 1) {
        CU(cudaMemcpyAsync(s->h_total, s->d_offsets + T, sizeof(int32_t), cudaMemcpyDeviceToHost, c->stream));
        kc::smp_gather_pool_kernel<<<(T + 127) / 128, 128, 0, c->stream>>>(sp, s->d_offsets, s->pool_v[s->pool_cur],
                                                                          s->pool_t[s->pool_cur], s->pool_n);
        c->launches += 2;
        CU(cudaStreamSynchronize(c->stream));
        CU(cudaGetLastError());
        added = (size_t)*s->h_total;
    } else {
        // own lists compacted into a staging block (the sort's key arrays are idle between level creations),
        // every rank's block and count to every GPU over NVLink, appended in rank order: the pools stay equal
        double* stage = reinterpret_cast<double*>(s->pool_key[1]);  // [2M]: values, then tiebreakers
        double* recv = reinterpret_cast<double*>(s->pool_key[0]);   // [R][2M]
        kc::smp_gather_pool_kernel<<<(T + 127) / 128, 128, 0, c->stream>>>(sp, s->d_offsets, stage, stage + M, 0);
        int e = kc::comm_allgather_dev(c, s->d_offsets + T, s->d_rank_counts, sizeof(int32_t));
        if (e) return e;
        e = kc::comm_allgather_dev(c, stage, recv, sizeof(double) * 2 * M);
        if (e) return e;
        CU(cudaMemcpyAsync(s->h_rank_counts, s->d_rank_counts, sizeof(int32_t) * R, cudaMemcpyDeviceToHost, c->stream));
        kc::smp_pool_append_ranks_kernel<<<dim3(296, R), 256, 0, c->stream>>>(recv, s->d_rank_counts, R, M,
                                                                            s->pool_v[s->pool_cur], s->pool_t[s->pool_cur],
                                                                            s->pool_n);
        c->launches += 3;
        CU(cudaStreamSynchronize(c->stream));
        CU(cudaGetLastError());
        for (int r = 0; r < R; r++) added += (size_t)s->h_rank_counts[r];
    }
    s->pool_n += added;
    if (counters) std::memcpy(counters, s->h_cnt_sum, sizeof(uint64_t) * 4 * L);
    if (n_new) *n_new = (int64_t)added;
    if (n_pool) *n_pool = (int64_t)s->pool_n;
    return KC_OK;
}

int kc_sampler_pool_select(kc_ctx* c, int64_t index, double* value, double* tiebreaker, int64_t* n_pool)
{
    int rc = kc::need_sampler(c);
    if (rc) return rc;
    SamplerState* s = c->sampler;
    if (!s->pool_cap) return fail(KC_ERR_ARG, "the sampler was created without an above pool (pool_reserve = 0)");
    const size_t n = s->pool_n;
    if (index < 0 || (size_t)index >= n) return fail(KC_ERR_ARG, "index outside the pool");
    CU(cudaSetDevice(c->device));
    const int cur = s->pool_cur, alt = cur ^ 1;
    const unsigned grid = (unsigned)((n + 255) / 256);
    // lexicographic (value, tiebreaker) with a stable LSD sort: by tiebreaker first, then by value
    kc::smp_pool_keys_kernel<<<grid, 256, 0, c->stream>>>(s->pool_t[cur], nullptr, n, s->pool_key[0], s->pool_idx[0]);
    size_t bytes = s->pool_tmp_bytes;
    CU(cub::DeviceRadixSort::SortPairs(s->pool_tmp, bytes, s->pool_key[0], s->pool_key[1], s->pool_idx[0], s->pool_idx[1],
                                       (int64_t)n, 0, 64, c->stream));
    kc::smp_pool_keys_kernel<<<grid, 256, 0, c->stream>>>(s->pool_v[cur], s->pool_idx[1], n, s->pool_key[0], nullptr);
    bytes = s->pool_tmp_bytes;
    CU(cub::DeviceRadixSort::SortPairs(s->pool_tmp, bytes, s->pool_key[0], s->pool_key[1], s->pool_idx[1], s->pool_idx[0],
                                       (int64_t)n, 0, 64, c->stream));
    double* d_pivot = reinterpret_cast<double*>(s->pool_key[0]);  // keys are dead after the second sort
    kc::smp_pool_cut_kernel<<<grid, 256, 0, c->stream>>>(s->pool_v[cur], s->pool_t[cur], s->pool_idx[0], n, (size_t)index,
                                                        s->pool_v[alt], s->pool_t[alt], d_pivot);
    CU(cudaMemcpyAsync(s->h_pivot, d_pivot, 2 * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
    c->launches += 3;
    CU(cudaStreamSynchronize(c->stream));
    CU(cudaGetLastError());
    s->pool_cur = alt;
    s->pool_n = n - (size_t)index - 1;
    if (value) *value = s->h_pivot[0];
    if (tiebreaker) *tiebreaker = s->h_pivot[1];
    if (n_pool) *n_pool = (int64_t)s->pool_n;
    return KC_OK;
}

int kc_sampler_pool_clear(kc_ctx* c)
{
    int rc = kc::need_sampler(c);
    if (rc) return rc;
    c->sampler->pool_n = 0;
    return KC_OK;
}

int kc_sampler_pool_fetch(kc_ctx* c, double* pairs, int64_t max_pairs, int64_t* n)
{
    int rc = kc::need_sampler(c);
    if (rc) return rc;
    SamplerState* s = c->sampler;
    if (!s->pool_cap) return fail(KC_ERR_ARG, "the sampler was created without an above pool (pool_reserve = 0)");
    CU(cudaSetDevice(c->device));
    const size_t take = s->pool_n < (size_t)max_pairs ? s->pool_n : (size_t)max_pairs;
    if (take && pairs) {
        double* d = nullptr;
        CU(cudaMalloc(&d, sizeof(double) * 2 * take));
        kc::smp_pool_pairs_kernel<<<(unsigned)((take + 255) / 256), 256, 0, c->stream>>>(s->pool_v[s->pool_cur],
                                                                                       s->pool_t[s->pool_cur], take, d);
        CU(cudaMemcpyAsync(pairs, d, sizeof(double) * 2 * take, cudaMemcpyDeviceToHost, c->stream));
        CU(cudaStreamSynchronize(c->stream));
        CU(cudaFree(d));
    }
    if (n) *n = (int64_t)s->pool_n;
    return KC_OK;
}

int kc_sampler_get_counters(kc_ctx* c, uint64_t* ctr)
{
    int rc = kc::need_sampler(c);
    if (rc) return rc;
    if (!ctr) return fail(KC_ERR_ARG, "ctr is null");
    CU(cudaSetDevice(c->device));
    CU(cudaMemcpyAsync(ctr, c->sampler->sp.ctr, sizeof(uint64_t) * c->sampler->sp.lanes, cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    return KC_OK;
}
int kc_sampler_set_counters(kc_ctx* c, const uint64_t* ctr)
{
    int rc = kc::need_sampler(c);
    if (rc) return rc;
    if (!ctr) return fail(KC_ERR_ARG, "ctr is null");
    CU(cudaSetDevice(c->device));
    CU(cudaMemcpyAsync(c->sampler->sp.ctr, ctr, sizeof(uint64_t) * c->sampler->sp.lanes, cudaMemcpyHostToDevice, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    return KC_OK;
}

// ---- test hooks ------------------------------------------------------------------------------
int kc_prior_eval(kc_ctx* c, const kc_prior* prior, int what, int n, const double* x, double* out)
{
    if (!c || !prior || !x || !out) return fail(KC_ERR_ARG, "null argument");
    if (prior->type < 0 || prior->type >= KC_PRIOR_N_TYPES) return fail(KC_ERR_ARG, "unknown prior type");
    if (what < 0 || what > 2) return fail(KC_ERR_ARG, "what must be 0, 1 or 2");
    if (n <= 0) return KC_OK;
    CU(cudaSetDevice(c->device));
    double* d = nullptr;
    CU(cudaMalloc(&d, sizeof(double) * 2 * n));
    CU(cudaMemcpyAsync(d, x, sizeof(double) * n, cudaMemcpyHostToDevice, c->stream));
    kc::smp_prior_eval_kernel<<<(n + 127) / 128, 128, 0, c->stream>>>(*prior, what, n, d, d + n);
    c->launches += 1;
    CU(cudaMemcpyAsync(out, d + n, sizeof(double) * n, cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    CU(cudaFree(d));
    return KC_OK;
}

int kc_rng_draws(kc_ctx* c, uint64_t key, uint64_t ctr0, int what, int n, double* out, uint64_t* ctr_end)
{
    if (!c || !out) return fail(KC_ERR_ARG, "null argument");
    if (what < 0 || what > 3) return fail(KC_ERR_ARG, "what must be 0..3");
    if (n <= 0) return KC_OK;
    CU(cudaSetDevice(c->device));
    double* d = nullptr;
    CU(cudaMalloc(&d, sizeof(double) * n + sizeof(uint64_t)));
    uint64_t* de = reinterpret_cast<uint64_t*>(d + n);
    kc::smp_rng_draws_kernel<<<1, 1, 0, c->stream>>>(key, ctr0, what, n, d, de);
    c->launches += 1;
    CU(cudaMemcpyAsync(out, d, sizeof(double) * n, cudaMemcpyDeviceToHost, c->stream));
    uint64_t end = 0;
    CU(cudaMemcpyAsync(&end, de, sizeof(uint64_t), cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    CU(cudaFree(d));
    if (ctr_end) *ctr_end = end;
    return KC_OK;
}

int kc_sampler_test_perturb(kc_ctx* c, double* state, int32_t* np, double* logH)
{
    int rc = kc::need_sampler(c);
    if (rc) return rc;
    SamplerState* s = c->sampler;
    const SamplerParams& sp = s->sp;
    if (sp.P != 1) return fail(KC_ERR_ARG, "kc_sampler_test_perturb needs num_particles == 1");
    if (!state || !np || !logH) return fail(KC_ERR_ARG, "null argument");
    CU(cudaSetDevice(c->device));
    const int T = sp.lanes;
    double* d = nullptr;
    CU(cudaMalloc(&d, sizeof(double) * T));
    kc::smp_test_perturb_kernel<<<kc::lane_grid(sp), kc::LANE_THREADS, 0, c->stream>>>(sp, d);
    c->launches += 1;
    CU(cudaMemcpy2DAsync(state, sizeof(double) * T, sp.Q, sizeof(double) * sp.ldq, sizeof(double) * T, sp.SL.n_state,
                         cudaMemcpyDeviceToHost, c->stream));
    CU(cudaMemcpyAsync(np, sp.npq, sizeof(int32_t) * T, cudaMemcpyDeviceToHost, c->stream));
    CU(cudaMemcpyAsync(logH, d, sizeof(double) * T, cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    CU(cudaFree(d));
    CU(cudaGetLastError());
    return KC_OK;
}

}  // extern "C"
