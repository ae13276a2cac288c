/*
 * libkima_cuda — device-side proposals and accept logic.
 *
 * With kc_eval the host still runs, per lane and MCMC step, the reference's
 *     Sampler::mcmc_thread              DNest4/code/SamplerImpl.h:134-165
 *     Sampler::update_particle          DNest4/code/SamplerImpl.h:168-221
 *     Sampler::update_level_assignment  DNest4/code/SamplerImpl.h:224-265
 *     RVmodel::perturb                  src/RVmodel.cpp:435-820
 *     RJObject::perturb (+ add/remove/perturb_components/perturb_num_components)
 *                                       DNest4/code/RJObject/RJObjectImpl.h:62-231
 *     ConditionalPrior::perturb1/2      DNest4/code/RJObject/ConditionalPriors/ConditionalPrior.cpp:17-55
 *     RVConditionalPrior                src/RVConditionalPrior.cpp:57-158
 *     ContinuousDistribution::perturb   DNest4/code/Distributions/ContinuousDistribution.cpp:12-20
 *     LikelihoodType::perturb, operator< DNest4/code/LikelihoodType.cpp:26-43
 * around every likelihood call.  The entry points below keep the particles, their
 * likelihoods and level assignments, every lane's level counters and `above` list in HBM
 * and run whole blocks of `thread_steps` lock-step MCMC steps of all lanes without the
 * host: per step one proposal kernel, the likelihood kernels of kc_eval_device, one accept
 * kernel.  The host comes back at DNest4's barrier (SamplerImpl.h:296-331) for the
 * bookkeeping, which stays the reference's code.
 *
 * The algorithm is the reference's, operation for operation and draw for draw; the random
 * stream is not (include/kc_philox.h): chains are reproducible for a given seed and lane
 * count but differ from the reference's mt19937 chains.  Parity is pinned by running the
 * unmodified reference sampler on the same stream (oracle/philox_rng.cpp).
 *
 * Plain C, int status codes, kc_last_error() for the message; all calls are made by the
 * one thread that owns the context.
 */
#ifndef KIMA_SAMPLER_H
#define KIMA_SAMPLER_H

#include "kima_cuda.h"

#ifdef __cplusplus
extern "C" {
#endif

/* A ContinuousDistribution (DNest4/code/Distributions/, src/distributions/Fixed.h) as data:
 * type + constructor arguments p[] in constructor order + the constants k[] the reference's
 * constructor derives from them (filled by kc_prior_make with the host libm, as the
 * reference does). */
enum {
    KC_PRIOR_NONE = -1,
    KC_PRIOR_FIXED = 0,            /* Fixed(val)                                        */
    KC_PRIOR_UNIFORM = 1,          /* Uniform(lower, upper)                             */
    KC_PRIOR_LOGUNIFORM = 2,       /* LogUniform(lower, upper); k0 = log(upper/lower)   */
    KC_PRIOR_MODLOGUNIFORM = 3,    /* ModifiedLogUniform(knee, upper); k0 = 1/log(1+upper/knee) */
    KC_PRIOR_GAUSSIAN = 4,         /* Gaussian(center, width)                           */
    KC_PRIOR_TRUNC_GAUSSIAN = 5,   /* TruncatedGaussian(center, width, lower, upper); k0 = cdf(lower), k1 = c */
    KC_PRIOR_CAUCHY = 6,           /* Cauchy(center, width)                             */
    KC_PRIOR_TRUNC_CAUCHY = 7,     /* TruncatedCauchy(center, width, lower, upper); k0, k1 as above */
    KC_PRIOR_EXPONENTIAL = 8,      /* Exponential(scale)                                */
    KC_PRIOR_TRUNC_EXPONENTIAL = 9,/* TruncatedExponential(scale, lower, upper); k0, k1 */
    KC_PRIOR_LAPLACE = 10,         /* Laplace(center, width)                            */
    KC_PRIOR_KUMARASWAMY = 11,     /* Kumaraswamy(a, b)                                 */
    KC_PRIOR_RAYLEIGH = 12,        /* Rayleigh(scale)                                   */
    KC_PRIOR_TRUNC_RAYLEIGH = 13,  /* TruncatedRayleigh(scale, lower, upper); k0 = lcdf, k1 = tp, k2 = logtp */
    KC_PRIOR_TRIANGULAR = 14,      /* Triangular(lower, centre, upper)                  */
    KC_PRIOR_PARETO = 15,          /* Pareto(min, alpha)                                */
    KC_PRIOR_MIX_GAUSS_LOGUNIFORM = 16, /* mixGaussianLogUniform(mean, sigma, lower, upper), src/distributions/
                                      mixGaussianLogUniform.cpp; k0 = log(upper/lower); cdf_inverse by the
                                      reference's bisection.  Gaussian_from_Tc (src/distributions/
                                      Gaussian_from_Tc.cpp) is tabled as the Gaussian(m, s) it wraps and
                                      Jeffreys is an alias of LogUniform; Empirical has no table form.  */
    KC_PRIOR_N_TYPES = 17
};

typedef struct kc_prior {
    int32_t type;
    int32_t reserved;
    double p[4];
    double k[3];
} kc_prior;

/* Fills k[] from type and p[] (the reference's constructors); KC_ERR_ARG where the
 * reference's constructor would throw std::domain_error. */
int kc_prior_make(int type, const double* p, int n_p, kc_prior* out);

/* Every prior an RVmodel holds (src/RVmodel.h:108-145) and its RVConditionalPrior
 * (src/RVConditionalPrior.h:30-47), after setPriors() (src/RVmodel.cpp:23-85) has filled in
 * the defaults.  Slots the flags do not use have type KC_PRIOR_NONE. */
typedef struct kc_prior_table {
    kc_prior Cprior, Jprior, slope_prior, quadr_prior, cubic_prior, offsets_prior, betaprior;
    kc_prior sigmaMA_prior, tauMA_prior;
    kc_prior log_eta1_prior, eta2_prior, eta3_prior, log_eta4_prior;
    kc_prior nu_prior;
    kc_prior KO_Pprior[KC_MAX_KO], KO_Kprior[KC_MAX_KO], KO_eprior[KC_MAX_KO], KO_phiprior[KC_MAX_KO],
        KO_wprior[KC_MAX_KO];
    /* conditional prior of the planets; under hyperpriors Pprior / Kprior are
     * Laplace(center, width) / Exponential(muK) with the particle's own hyperparameters */
    kc_prior Pprior, Kprior, eprior, phiprior, wprior;
    kc_prior log_muP_prior, wP_prior, log_muK_prior;
    int32_t fix;         /* RVmodel::fix: number of planets fixed at npmax */
    int32_t reserved;
} kc_prior_table;

/* The options the MCMC steps read (DNest4/code/Options.h) */
typedef struct kc_sampler_config {
    int32_t lanes;           /* DNest4 threads (-t) stepped by this context             */
    int32_t num_particles;   /* particles per lane                                      */
    int32_t max_steps;       /* largest `steps` of one kc_sampler_run (thread_steps)    */
    int32_t max_levels;      /* capacity of the level table                             */
    double lambda;           /* backtracking scale length                               */
    double beta;             /* strength of the uniform-exploration term                */
    uint64_t seed;           /* lane k draws from the stream with key seed + lane0 + k  */
    int32_t lane0;           /* index of this context's first lane in the whole run     */
    int32_t pool_reserve;    /* extra capacity of the device-resident `all_above` pool:
                                options.new_level_interval (0 = the pool is not used)   */
} kc_sampler_config;

/* Rows of a particle's state column: the kc_layout rows (what the likelihood reads), then
 * the RJObject's u_components, 5 blocks of npmax rows in the order of the component rows
 * (RJObject.h:43), then the conditional prior's center, width, muK
 * (src/RVConditionalPrior.h:17-20). */
typedef struct kc_sampler_layout {
    int32_t n_state;   /* rows per particle */
    int32_t row_u;     /* 5 * npmax rows    */
    int32_t row_hyper; /* 3 rows            */
    int32_t n_particles; /* lanes * num_particles = columns */
} kc_sampler_layout;

/* Needs kc_upload_data + kc_configure(max_batch >= lanes).  One sampler per context.
 * Unless kc_set_option(ctx, "gp_cache", 0) was called before, the sampler attaches a factor cache with one
 * slot per particle (kima_cuda.h: kc_gp_cache_enable) where that applies and fits the device memory; a lane's
 * slot is the particle it moves.  Chains are the same bits with and without it. */
int kc_sampler_create(kc_ctx* ctx, const kc_prior_table* priors, const kc_sampler_config* cfg);
int kc_sampler_destroy(kc_ctx* ctx);
int kc_sampler_get_layout(kc_ctx* ctx, kc_sampler_layout* out);

/* Particle states, host <-> device.  state is [n_state][n_particles] row-major (column i =
 * particle i, particles of lane k are columns k*num_particles ...), np / logl / tiebreaker /
 * level have n_particles entries.  Any pointer may be NULL (skipped). */
int kc_sampler_upload(kc_ctx* ctx, const double* state, const int32_t* np, const double* logl,
                      const double* tiebreaker, const int32_t* level);
int kc_sampler_download(kc_ctx* ctx, double* state, int32_t* np, double* logl, double* tiebreaker,
                        int32_t* level);
/* One particle's column (save_particle, SamplerImpl.h:503-531). */
int kc_sampler_get_particle(kc_ctx* ctx, int which, double* state_col, int32_t* np);
/* The reverse: overwrite one particle (a replacement that kill_lagging_particles takes from a lane
 * another rank owns, SamplerImpl.h:571-573 across processes). */
int kc_sampler_set_particle(kc_ctx* ctx, int which, const double* state_col, int32_t np, double logl,
                            double tiebreaker, int32_t level);
/* particles[dst[i]] = particles[src[i]] with likelihood and level assignment
 * (kill_lagging_particles, SamplerImpl.h:566-577); no dst may also be a src. */
int kc_sampler_copy_particles(kc_ctx* ctx, int n, const int32_t* dst, const int32_t* src);

/* The level table every lane's copy starts from at the barrier (SamplerImpl.h:283-284):
 * L levels; tries[] are the master counters (read by the uniform-exploration term,
 * SamplerImpl.h:252-253); enough = enough_levels(levels); at_max = (L == options.max_num_levels).
 * Also zeroes the lanes' counter increments and `above` lists. */
int kc_sampler_set_levels(kc_ctx* ctx, int L, const double* logl, const double* tiebreaker,
                          const double* log_X, const uint64_t* tries, int enough, int at_max);

/* `steps` lock-step MCMC steps of all lanes (mcmc_thread for every lane); asynchronous on
 * the context stream. */
int kc_sampler_run(kc_ctx* ctx, int steps);

/* The barrier (SamplerImpl.h:296-329): waits for the steps, then returns the lanes' counter
 * increments summed over lanes, counters[4*j + {0,1,2,3}] = accepts, tries, visits, exceeds
 * of level j, and the `above` lists concatenated in lane order as (value, tiebreaker) pairs
 * (at most max_pairs; *n_above receives the number found).  Either output may be NULL. */
int kc_sampler_collect(kc_ctx* ctx, uint64_t* counters, double* above_pairs, int max_pairs,
                       int* n_above);

/* ---- `all_above` kept in HBM ----------------------------------------------------------------
 * Sampler::all_above (Sampler.h:59; SamplerImpl.h:324-329, 384-399) grows by up to lanes x
 * thread_steps (value, tiebreaker) pairs per barrier and is only ever (a) counted, (b) ordered to
 * place a new level at position `index` and cut at that position, (c) cleared.  With tens of
 * thousands of lanes, shipping and ordering it on the host is most of the barrier; these entry
 * points keep it on the device.
 * kc_sampler_collect_pool: kc_sampler_collect with the lanes' `above` lists appended to the pool
 * instead of returned; *n_new = pairs appended, *n_pool = pool size afterwards.
 * On a context with a communicator (kc_comm_init before kc_sampler_create) it is a collective:
 * the counters come back summed over the ranks and every rank's pool receives the lists of all
 * ranks in rank order (NCCL all-gather between the GPUs), so the pools stay equal and
 * kc_sampler_pool_select returns the same pair on every rank — the host-side
 * kc_allreduce_levels / kc_allgather_above are not needed then.
 * kc_sampler_pool_select: orders the pool by (value, tiebreaker) (LikelihoodType::operator<),
 * returns the element at position `index` and keeps the elements after it — do_bookkeeping's
 * sort + levels.push_back(all_above[index]) + erase(begin, begin + index + 1).
 * kc_sampler_pool_clear: all_above.clear().  kc_sampler_pool_fetch: test hook (pool order). */
int kc_sampler_collect_pool(kc_ctx* ctx, uint64_t* counters, int64_t* n_new, int64_t* n_pool);
int kc_sampler_pool_select(kc_ctx* ctx, int64_t index, double* value, double* tiebreaker, int64_t* n_pool);
int kc_sampler_pool_clear(kc_ctx* ctx);
int kc_sampler_pool_fetch(kc_ctx* ctx, double* pairs, int64_t max_pairs, int64_t* n);

/* Device time (CUDA events on the context stream) of the last kc_sampler_run, all kernels of
 * its steps included; waits for the run to finish. */
int kc_sampler_last_run_ms(kc_ctx* ctx, float* ms);

/* Stream positions of the lanes (the number of uniforms each has drawn); set for tests. */
int kc_sampler_get_counters(kc_ctx* ctx, uint64_t* ctr);
int kc_sampler_set_counters(kc_ctx* ctx, const uint64_t* ctr);

/* ---- test hooks ------------------------------------------------------------------------ */
/* what: 0 = cdf(x), 1 = cdf_inverse(x), 2 = log_pdf(x); n values. */
int kc_prior_eval(kc_ctx* ctx, const kc_prior* prior, int what, int n, const double* x, double* out);
/* n draws of lane-stream `key` starting at position ctr0: what 0 = rand, 1 = randn,
 * 2 = randt2, 3 = randh; out[n]; *ctr_end = position afterwards. */
int kc_rng_draws(kc_ctx* ctx, uint64_t key, uint64_t ctr0, int what, int n, double* out, uint64_t* ctr_end);
/* RVmodel::perturb alone: for every lane (num_particles must be 1) proposal = particle,
 * logH[lane] = proposal.perturb(rng of the lane); the proposals' state columns are
 * returned in state [n_state][lanes], np[lanes]. */
int kc_sampler_test_perturb(kc_ctx* ctx, double* state, int32_t* np, double* logH);

#ifdef __cplusplus
}
#endif
#endif /* KIMA_SAMPLER_H */
