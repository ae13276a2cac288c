/*
 * libkima_cuda — C ABI of the B200-native batched likelihood engine for kima.
 *
 * This is the drop-in boundary of the hot path: the reference evaluates one
 * particle at a time through three private RVmodel methods,
 *
 *     RVmodel::calculate_mu()    /root/reference/src/RVmodel.cpp:286-395
 *     RVmodel::calculate_C()     /root/reference/src/RVmodel.cpp:170-216 (standard kernel)
 *     RVmodel::log_likelihood()  /root/reference/src/RVmodel.cpp:827-943
 *     kepler::kepler / true_anomaly                 src/RVmodel.cpp:1422-1475
 *
 * reading its inputs from the Data singleton (src/Data.h:52-108).  Here the
 * same three steps run for a whole batch of particles ("lanes") per call.
 * Plain C: pointers and sizes only, int status codes, no exceptions.
 *
 * Threading: one kc_ctx per GPU.  A context is single-caller (the sampler's
 * step loop); different contexts may be used from different threads.
 */
#ifndef KIMA_CUDA_H
#define KIMA_CUDA_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define KC_OK 0
#define KC_ERR_ARG 1      /* bad argument / call order                        */
#define KC_ERR_CUDA 2     /* CUDA runtime failure (see kc_last_error)         */
#define KC_ERR_UNSUPPORTED 3 /* a flag combination the library does not implement */
#define KC_ERR_COMM 4     /* NCCL failure                                     */

#define KC_MAX_NPMAX 32   /* planets per particle held in shared memory       */
#define KC_MAX_KO 8
#define KC_MAX_INST 16
#define KC_MAX_IND 8

typedef struct kc_ctx kc_ctx;

/* The compile-time model flags of a kima_setup.cpp (src/RVmodel.h:16-34,
 * src/RVConditionalPrior.h:11) plus npmax (RVmodel ctor).  They are link-time
 * constants in the reference; the host shim reads them once and passes them
 * here. */
typedef struct kc_model_flags {
    int32_t GP;               /* extern const bool GP                         */
    int32_t MA;               /* MA(1) term of calculate_mu (:378-386)        */
    int32_t hyperpriors;      /* P row holds log(P): P = exp(c[0]) (:357-360) */
    int32_t trend;
    int32_t degree;           /* 0..3                                         */
    int32_t multi_instrument;
    int32_t known_object;
    int32_t n_known_object;
    int32_t studentt;
    int32_t npmax;            /* max planets per particle                     */
    int32_t kernel;           /* 0 = standard QP (only one supported)         */
} kc_model_flags;

/* Row map of the SoA parameter block params[n_rows][ld] (ld >= B, column b
 * is lane b).  A row index of -1 means "group absent for these flags". */
typedef struct kc_layout {
    int32_t n_rows;
    int32_t npmax;
    int32_t row_P;     /* npmax rows: components[j][0] (P, or log P if hyperpriors) */
    int32_t row_K;     /* npmax rows: components[j][1]                       */
    int32_t row_phi;   /* npmax rows: components[j][2]                       */
    int32_t row_ecc;   /* npmax rows: components[j][3]                       */
    int32_t row_w;     /* npmax rows: components[j][4]                       */
    int32_t row_vsys;  /* background                                         */
    int32_t n_jit;     /* 1 (extra_sigma) or number_instruments (jitters)    */
    int32_t row_jit;
    int32_t n_off;     /* number_instruments-1 if multi_instrument else 0    */
    int32_t row_off;
    int32_t row_trend; /* 3 rows slope,quadr,cubic if trend else -1          */
    int32_t n_beta;    /* number_indicators                                  */
    int32_t row_beta;
    int32_t row_eta;   /* 4 rows eta1..eta4 if GP else -1                    */
    int32_t row_nu;    /* 1 row if studentt else -1                          */
    int32_t n_ko;
    int32_t row_KO_P, row_KO_K, row_KO_e, row_KO_phi, row_KO_w; /* n_ko rows each */
    int32_t row_MA;    /* 2 rows sigmaMA, tauMA if MA else -1                */
} kc_layout;

/* Pure function of the flags and the data shape; shared by the library, the
 * host mirror, the oracle and the reference harness so they cannot drift. */
static inline void kc_layout_from_flags(const kc_model_flags* f, int n_inst,
                                        int n_ind, kc_layout* L)
{
    int r = 0;
    L->npmax = f->npmax;
    L->row_P = r;   r += f->npmax;
    L->row_K = r;   r += f->npmax;
    L->row_phi = r; r += f->npmax;
    L->row_ecc = r; r += f->npmax;
    L->row_w = r;   r += f->npmax;
    L->row_vsys = r; r += 1;
    L->n_jit = f->multi_instrument ? n_inst : 1;
    L->row_jit = r; r += L->n_jit;
    L->n_off = f->multi_instrument ? (n_inst - 1) : 0;
    L->row_off = L->n_off ? r : -1; r += L->n_off;
    L->row_trend = f->trend ? r : -1; r += f->trend ? 3 : 0;
    L->n_beta = n_ind;
    L->row_beta = n_ind ? r : -1; r += n_ind;
    L->row_eta = f->GP ? r : -1; r += f->GP ? 4 : 0;
    L->row_nu = f->studentt ? r : -1; r += f->studentt ? 1 : 0;
    L->n_ko = f->known_object ? f->n_known_object : 0;
    L->row_KO_P = L->n_ko ? r : -1;   r += L->n_ko;
    L->row_KO_K = L->n_ko ? r : -1;   r += L->n_ko;
    L->row_KO_e = L->n_ko ? r : -1;   r += L->n_ko;
    L->row_KO_phi = L->n_ko ? r : -1; r += L->n_ko;
    L->row_KO_w = L->n_ko ? r : -1;   r += L->n_ko;
    L->row_MA = f->MA ? r : -1; r += f->MA ? 2 : 0;
    L->n_rows = r;
}

/* ---- lifetime ---------------------------------------------------------- */
int kc_create(int device, kc_ctx** out);
int kc_destroy(kc_ctx* ctx);
/* Thread-local message of the last failing call. */
const char* kc_last_error(void);
/* Library / build identification, e.g. "kima_cuda 0.1 sm_100a". */
const char* kc_version(void);

/* ---- data: replaces the per-call Data::get_instance() copies ------------
 * (src/RVmodel.cpp:173,288,829; getters src/Data.h:52-108).  Called once per
 * data set; the library copies, the caller keeps ownership.
 * obsi is 1-based (src/Data.cpp:263-269) and may be NULL when n_inst == 1.
 * actind is [n_ind][N] row-major, already mean-subtracted (Data.cpp:166-173),
 * NULL if n_ind == 0.  M0_epoch = t[0] in the reference (Data.cpp:162);
 * t_middle = Data::get_t_middle() (Data.h:66). */
int kc_upload_data(kc_ctx* ctx, int N, const double* t, const double* y,
                   const double* sig, const int32_t* obsi, int n_inst,
                   int n_ind, const double* actind, double M0_epoch,
                   double t_middle);

/* ---- model flags + batch capacity --------------------------------------- */
int kc_configure(kc_ctx* ctx, const kc_model_flags* flags, int max_batch);
int kc_get_layout(kc_ctx* ctx, kc_layout* out);

/* Pinned host staging block params[n_rows][max_batch] and np[max_batch]
 * (num_components of each lane's RJObject).  The sampler shim writes lane b's
 * proposal into column b, then calls kc_eval. */
int kc_host_params(kc_ctx* ctx, double** params, int32_t** np, int* ld);
/* The device-resident copies (same shape), for producers that already live on
 * the GPU and for kc_eval_device. */
int kc_device_params(kc_ctx* ctx, double** d_params, int32_t** d_np, int* ld);

/* ---- the hot path --------------------------------------------------------
 * Scores lanes [0,B): replaces calculate_mu -> [calculate_C] -> log_likelihood
 * for each.  logL follows RVmodel::log_likelihood exactly, including the
 * non-finite -> +inf rule (src/RVmodel.cpp:938-941); DNest4's LikelihoodType
 * then maps that to -1e300 (LikelihoodType.cpp:20-21).
 * kc_eval: H2D of the staged block, kernels, D2H into logL_host, synchronous.
 * kc_eval_async / kc_wait: same, split so the host can perturb the next
 * half-batch meanwhile (logL_host must be pinned or will be staged).
 * kc_eval_device: parameters already in the device block; result stays on the
 * device (d_logL, B doubles); asynchronous on the context stream. */
int kc_eval(kc_ctx* ctx, int B, double* logL_host);
int kc_eval_async(kc_ctx* ctx, int B, double* logL_host);
int kc_wait(kc_ctx* ctx);
int kc_eval_device(kc_ctx* ctx, int B, double* d_logL);
/* Allocation-free variant of the kernel=celerite likelihood.  NO COUNTERPART IN THE REFERENCE:
 * added for callers that cannot call cudaMalloc (code
 * under CUDA graph capture, tight serving loops, harnesses that forbid it).  Same two kernels
 * and the same results as kc_eval_device on a context configured with GP=true, kernel=1, but
 * the [N][roundup(B,32)] residual scratch is a file-scope `static __device__` array of
 * KC_NOALLOC_MAX_DOUBLES doubles (4 MiB, held by every GPU that loads the library until the
 * process exits) instead of the block kc_configure allocates.  The call makes no cudaMalloc /
 * cudaFree / allocating library call; the context (kc_create, kc_upload_data, kc_configure)
 * is set up beforehand.  The caller passes the inputs and the output as device pointers:
 *   d_params  [n_rows][ld] SoA parameter block (kc_get_layout), column b = lane b
 *   d_np      [B] planets per lane;   d_logL  [B] result
 * (kc_device_params hands out the context's own block, but any device memory will do).
 * Asynchronous on the context stream (kc_sync / kc_stream).  Limit: N * roundup(B,32) <=
 * KC_NOALLOC_MAX_DOUBLES, else KC_ERR_ARG with a message naming the limit.  The first call
 * on each device resolves the array's address (cudaGetSymbolAddress; B = 0 does only that,
 * e.g. before a graph capture starts).  All calls on a GPU share the one array: calls that
 * overlap on that GPU (two contexts, two streams) race; nothing in it needs clearing between
 * calls (every entry read is written first by the same call). */
#define KC_NOALLOC_MAX_DOUBLES (1 << 19)
int kc_eval_celerite_noalloc(kc_ctx* ctx, int B, const double* d_params,
                             const int32_t* d_np, int ld, double* d_logL);
int kc_sync(kc_ctx* ctx);
/* cudaStream_t of the context, as void*. */
int kc_stream(kc_ctx* ctx, void** stream);

/* ---- tuning (all have safe defaults) -------------------------------------
 * kepler_mode: 0 = algebraic true-anomaly form (default, no acos),
 *              1 = the reference's operation chain (acos + cos(f+w)).
 * gp_small:    1 = shared-memory-resident GP kernel whenever N fits (default), 0 = always the blocked path.
 * gp_streams:  lane groups of the blocked GP path, 1..4 (default 2); gp_slots: lanes in flight (0 = by memory);
 *              both take effect at the next kc_configure.
 * eval_chunk:  kc_eval batches of >= 2 chunks overlap H2D with the kernels (default 32768 lanes, 0 = off).
 * gp_order:    1 = the shared-memory-resident GP kernel takes the lanes of a batch (>= 1024) in order of planet
 *              count, which keeps the SMs in step (default; results do not depend on it), 0 = in lane order.
 * gp_cache:    1 = kc_sampler_create attaches a factor cache, one slot per particle (default), 0 = it does not. */
int kc_set_option(kc_ctx* ctx, const char* key, int value);

/* Factor cache of the shared-memory-resident GP path (GP=true, standard kernel, N small enough for it, no
 * MA term).  The covariance matrix of RVmodel::calculate_C (src/RVmodel.cpp:170-216) depends on eta1..eta4
 * and the jitters only, and most of RVmodel::perturb's proposals (src/RVmodel.cpp:448-544) leave them alone;
 * the reference factors C anyway on every log_likelihood() call (src/RVmodel.cpp:841-873).  With a cache of
 * n_slots slots, a lane that names a slot (kc_set_slots; the device-side sampler names the particle it moves)
 * finds the factor of its covariance matrix in one of the slot's two entries whenever that matrix was one of
 * the last two factored there -- the particle's state or its latest proposal -- and then only recomputes the
 * residuals and the solve.  Results do not depend on the cache: a hit returns the bits a miss would.
 * n_slots = 0 frees the cache.  Costs 2 x n_slots x ~N^2/2 doubles of device memory (KC_ERR_ARG when that
 * does not fit); succeeds and stays inactive on contexts the cache does not apply to (kc_gp_cache_slots
 * then reports 0).  kc_upload_data / kc_configure drop the cache.  No counterpart in the reference. */
int kc_gp_cache_enable(kc_ctx* ctx, int n_slots);
int kc_gp_cache_slots(kc_ctx* ctx, int* n_slots);
/* The slots of lanes 0 .. B-1 of the NEXT kc_eval / kc_eval_async / kc_eval_device call of exactly B lanes
 * (consumed by it): slots[i] < 0 = lane i is evaluated without the cache; no slot twice in one batch. */
int kc_set_slots(kc_ctx* ctx, int B, const int32_t* slots);

/* ---- instrumentation ---------------------------------------------------- */
/* Device time (CUDA events on the context stream) of the kernels of the last
 * evaluation call, and the number of kernel launches since kc_create. */
int kc_last_kernel_ms(kc_ctx* ctx, float* keplerian_ms, float* gp_ms);
int64_t kc_launch_count(kc_ctx* ctx);
/* Writes a 256 MiB scratch buffer on the context stream so that the next timed
 * evaluation starts with a cold L2 (bench hygiene; untimed by kc_last_kernel_ms). */
int kc_flush_l2(kc_ctx* ctx);
/* FP64 peak microbenchmarks run on the context's device: vector DFMA and
 * tensor DMMA TFLOP/s (best of `reps`). */
int kc_measure_fp64_peak(kc_ctx* ctx, int reps, double* dfma_tflops,
                         double* dmma_tflops);

/* ---- test hooks (parity tests call these; not on the hot path) ----------- */
/* E = kepler::kepler(M, ecc)  (src/RVmodel.cpp:1422-1445), n scalars. */
int kc_kepler_E(kc_ctx* ctx, int n, const double* M, const double* ecc,
                double* E);
/* f = kepler::true_anomaly(t, P, ecc, Tp)  (src/RVmodel.cpp:1457-1475). */
int kc_true_anomaly(kc_ctx* ctx, int n, const double* t, const double* P,
                    const double* ecc, const double* Tp, double* f);
/* mu[N] of lane `lane` of the currently staged host block (calculate_mu). */
int kc_model_mu(kc_ctx* ctx, int lane, double* mu);
/* C[N][N] (column-major, full symmetric) of lane `lane` (calculate_C). */
int kc_gp_cov(kc_ctx* ctx, int lane, double* C);

/* ---- posterior-predictive curves --------------
 * Replaces the per-sample Python loops of pykima's post-processing,
 *   KimaResults.model()        pykima/results.py:723-797
 *   GP.predict(y - model, t)   pykima/GP.py:96-112, per sample at results.py:1313-1318
 * for the first B staged parameter vectors (kc_host_params) at once.
 *   t_pred   nt grid times (host), or NULL = the data epochs (then nt must be N)
 *   what     KC_PRED_MODEL: model_out[b*nt + k] = Keplerians (known objects included) +
 *            vsys + trend at t_pred[k]; at the data epochs also the instrument offsets and
 *            indicator terms, i.e. calculate_mu without the MA term
 *            KC_PRED_GP: gp_out[b*nt + k] = K(t_pred[k], t) C^-1 (y - mu), C and mu those of
 *            the C++ likelihood (src/RVmodel.cpp:186-216, 286-375); NaN where C is not
 *            positive definite.  Needs GP=true.  With kernel=celerite this is
 *            GP_celerite.predict (pykima/GP.py:349-355): the solver's own O(N + nt) recursions,
 *            CholeskySolver::solve / predict (celerite/cpp/include/celerite/solver/
 *            cholesky.h:217-262, 598-690) on the solver set up as in src/RVmodel.cpp:218-266;
 *            t_pred must then be sorted ascending (KC_ERR_ARG otherwise).
 * Outputs are host arrays of B*nt doubles (either may be NULL when not asked for). */
#define KC_PRED_MODEL 1
#define KC_PRED_GP 2
int kc_predict(kc_ctx* ctx, int B, int nt, const double* t_pred, int what,
               double* model_out, double* gp_out);
/* The predictive standard deviation of the same GP, GP.predict(..., return_std=True)
 * (pykima/GP.py:113-131, reached through predict_with_hyperpars(return_std=True), :134-175):
 *   gp_std_out[b*nt + k] = sqrt(max(0, k(t*,t*) - k(t*,t) C^-1 k(t,t*))),  k(t*,t*) = eta1^2,
 * by forward substitution of the nt right-hand sides against the Cholesky factor of the
 * likelihood (pykima forms C^-1 explicitly); NaN where C is not positive definite.
 * gp_out (the predictive mean, as KC_PRED_GP) may be NULL.  Needs GP=true, standard kernel. */
#define KC_PRED_GP_STD 4
int kc_predict_std(kc_ctx* ctx, int B, int nt, const double* t_pred, double* gp_out,
                   double* gp_std_out);

/* ---- multi-GPU: the level-creation barrier of DNest4 ----------------------
 * (DNest4/code/SamplerImpl.h:306-329).  One rank per GPU. */
/* Fills 128 bytes with an NCCL unique id (rank 0 calls, then broadcasts). */
int kc_comm_unique_id(void* id128);
int kc_comm_init(kc_ctx* ctx, int rank, int nranks, const void* id128);
/* Gather every rank's `above` list of (value, tiebreaker) pairs.
 * pairs: n*2 doubles from this rank; out: capacity max_total*2 doubles,
 * rank-major; counts[nranks] receives each rank's n. */
int kc_allgather_above(kc_ctx* ctx, const double* pairs, int n, double* out,
                       int max_total, int* counts);
/* Gather nbytes from every rank (the same nbytes on all): recv is nranks * nbytes, rank-major.
 * Used for the lanes' particle states, which kill_lagging_particles (SamplerImpl.h:534-587)
 * may copy between lanes of different ranks, and for the RNG that do_bookkeeping draws from. */
int kc_allgather_bytes(kc_ctx* ctx, const void* send, size_t nbytes, void* recv);
/* Sum the per-level accepts/tries/visits/exceeds deltas over ranks. */
int kc_allreduce_levels(kc_ctx* ctx, uint64_t* counters, int n);
int kc_comm_destroy(kc_ctx* ctx);

#ifdef __cplusplus
}
#endif
#endif /* KIMA_CUDA_H */
