// TEST INFRASTRUCTURE — not part of the product.  Only tests/ may load what this builds.
//
// Drives the UNMODIFIED reference sampler code (compiled where it lies under /root/reference
// by oracle/Makefile) on the counter-based random stream of include/kc_philox.h, so that the
// device-side sampler (kima_b200/csrc/kc_sampler_logic.cuh) can be
// compared with it draw for draw:
//     RVmodel::perturb                          src/RVmodel.cpp:435-820
//     Sampler::mcmc_thread / update_particle / update_level_assignment
//                                               DNest4/code/SamplerImpl.h:134-265
// and everything they call (RJObject, RVConditionalPrior, the distributions).  The only
// reference source left out of oracle/_ref/libkima_ref_philox.so is DNest4/code/RNG.cpp, whose
// stand-in is oracle/philox_rng.cpp.
//
// This TU includes ref_harness.cpp: the flags / data / parameter plumbing (ref_set_flags,
// ref_set_data, ref_load*, ref_eval_batch, ...) is the same code in both shared objects.
#include "ref_harness.cpp"

// the distributions of kima's own src/distributions need their private members too
#define private public
#define protected public
#include "distributions/Fixed.h"
#include "distributions/Gaussian_from_Tc.h"
#include "distributions/mixGaussianLogUniform.h"
#undef private
#undef protected

#include "../host/PriorTable.h"
#include "../kima_b200/csrc/kc_sampler_logic.cuh"  // prior_make only (host twin of the constructors)

extern "C" {
void philox_rng_set(DNest4::RNG* r, uint64_t key, uint64_t ctr);
void philox_rng_get(const DNest4::RNG* r, uint64_t* key, uint64_t* ctr);
}

namespace {

kc_prior_table g_pri;
bool g_have_pri = false;

int make_prior_host(int type, const double* p, int n_p, kc_prior* out)
{
    return kc::smp::prior_make(type, p, n_p, out);
}

std::shared_ptr<DNest4::ContinuousDistribution> make_dist(const kc_prior& q)
{
    using namespace DNest4;
    const double a = q.p[0], b = q.p[1], c = q.p[2], d = q.p[3];
    switch (q.type) {
    case KC_PRIOR_FIXED: return std::make_shared<Fixed>(a);
    case KC_PRIOR_UNIFORM: return std::make_shared<Uniform>(a, b);
    case KC_PRIOR_LOGUNIFORM: return std::make_shared<LogUniform>(a, b);
    case KC_PRIOR_MODLOGUNIFORM: return std::make_shared<ModifiedLogUniform>(a, b);
    case KC_PRIOR_GAUSSIAN: return std::make_shared<Gaussian>(a, b);
    case KC_PRIOR_TRUNC_GAUSSIAN: return std::make_shared<TruncatedGaussian>(a, b, c, d);
    case KC_PRIOR_CAUCHY: return std::make_shared<Cauchy>(a, b);
    case KC_PRIOR_TRUNC_CAUCHY: return std::make_shared<TruncatedCauchy>(a, b, c, d);
    case KC_PRIOR_EXPONENTIAL: return std::make_shared<Exponential>(a);
    case KC_PRIOR_TRUNC_EXPONENTIAL: return std::make_shared<TruncatedExponential>(a, b, c);
    case KC_PRIOR_LAPLACE: return std::make_shared<Laplace>(a, b);
    case KC_PRIOR_KUMARASWAMY: return std::make_shared<Kumaraswamy>(a, b);
    case KC_PRIOR_RAYLEIGH: return std::make_shared<Rayleigh>(a);
    case KC_PRIOR_TRUNC_RAYLEIGH: return std::make_shared<TruncatedRayleigh>(a, b, c);
    case KC_PRIOR_TRIANGULAR: return std::make_shared<Triangular>(a, b, c);
    case KC_PRIOR_PARETO: return std::make_shared<Pareto>(a, b);
    case KC_PRIOR_MIX_GAUSS_LOGUNIFORM: return std::make_shared<mixGaussianLogUniform>(a, b, c, d);
    }
    return nullptr;
}

// The table's priors into a model: the slots setPriors() only fills when empty are assigned
// before it runs, the ones it always overwrites (betaprior, sigmaMA_prior, tauMA_prior,
// nu_prior, src/RVmodel.cpp:27-31,82-83) afterwards.
void install_priors(RVmodel& m)
{
    if (g_have_pri) {
        auto put = [](std::shared_ptr<DNest4::ContinuousDistribution>& slot, const kc_prior& q) {
            if (q.type != KC_PRIOR_NONE) slot = make_dist(q);
        };
        put(m.Cprior, g_pri.Cprior);
        put(m.Jprior, g_pri.Jprior);
        put(m.slope_prior, g_pri.slope_prior);
        put(m.quadr_prior, g_pri.quadr_prior);
        put(m.cubic_prior, g_pri.cubic_prior);
        put(m.offsets_prior, g_pri.offsets_prior);
        put(m.log_eta1_prior, g_pri.log_eta1_prior);
        put(m.eta2_prior, g_pri.eta2_prior);
        put(m.eta3_prior, g_pri.eta3_prior);
        put(m.log_eta4_prior, g_pri.log_eta4_prior);
        for (int i = 0; i < g_flags.n_known_object && g_flags.known_object; i++) {
            put(m.KO_Pprior[i], g_pri.KO_Pprior[i]);
            put(m.KO_Kprior[i], g_pri.KO_Kprior[i]);
            put(m.KO_eprior[i], g_pri.KO_eprior[i]);
            put(m.KO_phiprior[i], g_pri.KO_phiprior[i]);
            put(m.KO_wprior[i], g_pri.KO_wprior[i]);
        }
        RVConditionalPrior& c = m.planets.conditional_prior;
        if (!g_flags.hyperpriors) {
            put(c.Pprior, g_pri.Pprior);
            put(c.Kprior, g_pri.Kprior);
        }
        put(c.eprior, g_pri.eprior);
        put(c.phiprior, g_pri.phiprior);
        put(c.wprior, g_pri.wprior);
        put(c.log_muP_prior, g_pri.log_muP_prior);
        put(c.wP_prior, g_pri.wP_prior);
        put(c.log_muK_prior, g_pri.log_muK_prior);
    }
    m.setPriors();
    if (g_have_pri) {
        auto put = [](std::shared_ptr<DNest4::ContinuousDistribution>& slot, const kc_prior& q) {
            if (q.type != KC_PRIOR_NONE) slot = make_dist(q);
        };
        put(m.betaprior, g_pri.betaprior);
        put(m.sigmaMA_prior, g_pri.sigmaMA_prior);
        put(m.tauMA_prior, g_pri.tauMA_prior);
        put(m.nu_prior, g_pri.nu_prior);
    }
}

struct ColView {
    double* base;
    size_t ld;
    double& operator()(int row) const { return base[(size_t)row * ld]; }
};

void load_model(RVmodel& m, const kc_sampler_layout& SL, double* state, size_t ld, int col, int np)
{
    kima_b200::column_to_model(m, g_layout, SL, ColView{state + col, ld}, np);
    // perturb() updates mu (and C) incrementally: they must hold the current model
    m.calculate_mu();
    if (g_flags.GP) m.calculate_C();
}

}  // namespace

extern "C" {

// The defaults of setPriors() (src/RVmodel.cpp:23-85) and of the RVConditionalPrior constructor
// (src/RVConditionalPrior.cpp:10-54) for the current flags and data, as a table.
int rs_default_priors(kc_prior_table* out)
{
    RVmodel m;
    // known objects have no default priors (setPriors throws without them, :71-76): placeholders
    // here, the caller fills those slots of the table in
    for (int i = 0; i < g_flags.n_known_object && g_flags.known_object; i++)
        for (auto* v : {&m.KO_Pprior, &m.KO_Kprior, &m.KO_eprior, &m.KO_phiprior, &m.KO_wprior})
            (*v)[i] = std::make_shared<DNest4::Uniform>(0., 1.);
    m.setPriors();
    const std::string bad = kima_b200::extract_prior_table(m, *out, make_prior_host);
    return bad.empty() ? 0 : 1;
}

int rs_set_priors(const kc_prior_table* t)
{
    g_have_pri = t != nullptr;
    if (t) g_pri = *t;
    return 0;
}

int rs_prior_make(int type, const double* p, int n_p, kc_prior* out) { return make_prior_host(type, p, n_p, out); }

// what: 0 = cdf, 1 = cdf_inverse, 2 = log_pdf of the reference's distribution object
int rs_prior_eval(const kc_prior* q, int what, int n, const double* x, double* out)
{
    auto d = make_dist(*q);
    if (!d) return 1;
    for (int i = 0; i < n; i++)
        out[i] = what == 0 ? d->cdf(x[i]) : (what == 1 ? d->cdf_inverse(x[i]) : d->log_pdf(x[i]));
    return 0;
}

// what: 0 = rand, 1 = randn, 2 = randt2, 3 = randh of the stand-in RNG
uint64_t rs_rng_draws(uint64_t key, uint64_t ctr0, int what, int n, double* out)
{
    DNest4::RNG r;
    philox_rng_set(&r, key, ctr0);
    for (int i = 0; i < n; i++)
        out[i] = what == 0 ? r.rand() : (what == 1 ? r.randn() : (what == 2 ? r.randt2() : r.randh()));
    uint64_t k, c;
    philox_rng_get(&r, &k, &c);
    return c;
}

int rs_sampler_layout(int n_particles, kc_sampler_layout* out)
{
    *out = kima_b200::sampler_layout(g_layout, n_particles);
    return 0;
}

// RVmodel::perturb on B models: column b of state [n_state][ld] in, the perturbed model out
// (in place), logH[b], stream of lane b = (key0 + b, ctr[b]) (ctr updated).
int rs_perturb(int B, int ld, double* state, int32_t* np, uint64_t key0, uint64_t* ctr, double* logH)
{
    const kc_sampler_layout SL = kima_b200::sampler_layout(g_layout, B);
    for (int b = 0; b < B; b++) {
        RVmodel m;
        install_priors(m);
        load_model(m, SL, state, (size_t)ld, b, np[b]);
        DNest4::RNG rng;
        philox_rng_set(&rng, key0 + (uint64_t)b, ctr[b]);
        logH[b] = m.perturb(rng);
        uint64_t k;
        philox_rng_get(&rng, &k, &ctr[b]);
        kima_b200::model_to_column(m, g_layout, SL, ColView{state + b, (size_t)ld}, np[b]);
    }
    return 0;
}

// `steps` iterations of Sampler::mcmc_thread for each of T threads with P particles each, from
// the given sampler state; everything in place.  levels: L rows of
// (logl, tiebreaker, log_X) doubles and (accepts, tries, visits, exceeds) counters; the
// counters come back as the sum over threads of each thread's increments (what run_thread
// folds into `levels`, SamplerImpl.h:303-321).  above_pairs receives the threads' `above`
// lists in thread order, n_above_lane[T] their lengths.
int rs_mcmc(int T, int P, int steps, int ld, double* state, int32_t* np, double* logl, double* tb,
            int32_t* level, int L, const double* lv_logl, const double* lv_tb, const double* lv_logX,
            uint64_t* lv_counters, double lambda, double beta, int max_num_levels, uint64_t key0,
            uint64_t* ctr, double* above_pairs, int max_pairs, int32_t* n_above_lane)
{
    using namespace DNest4;
    const kc_sampler_layout SL = kima_b200::sampler_layout(g_layout, T * P);
    Options opt(P, 1000000, 1000000, steps, max_num_levels, lambda, beta, 1);
    Sampler<RVmodel> s(T, exp(1.0), opt, false);
    for (int i = 0; i < T * P; i++) {
        install_priors(s.particles[i]);
        load_model(s.particles[i], SL, state, (size_t)ld, i, np[i]);
        s.log_likelihoods[i].value = logl[i];
        s.log_likelihoods[i].tiebreaker = tb[i];
        s.level_assignments[i] = (unsigned int)level[i];
    }
    s.levels.clear();
    for (int j = 0; j < L; j++) {
        LikelihoodType lt;
        lt.value = lv_logl[j];
        lt.tiebreaker = lv_tb[j];
        Level lv(lt);
        lv.log_X = lv_logX[j];
        lv.accepts = lv_counters[4 * j + 0];
        lv.tries = lv_counters[4 * j + 1];
        lv.visits = lv_counters[4 * j + 2];
        lv.exceeds = lv_counters[4 * j + 3];
        s.levels.push_back(lv);
    }
    for (int k = 0; k < T; k++) {
        s.copies_of_levels[k] = s.levels;
        philox_rng_set(&s.rngs[k], key0 + (uint64_t)k, ctr[k]);
        s.above[k].clear();
    }
    for (int k = 0; k < T; k++) s.mcmc_thread(k);
    std::vector<uint64_t> sum(4 * (size_t)L, 0);
    int n_pairs = 0;
    for (int k = 0; k < T; k++) {
        const std::vector<Level>& lv = s.copies_of_levels[k];
        for (int j = 0; j < L; j++) {
            sum[4 * j + 0] += lv[j].accepts - s.levels[j].accepts;
            sum[4 * j + 1] += lv[j].tries - s.levels[j].tries;
            sum[4 * j + 2] += lv[j].visits - s.levels[j].visits;
            sum[4 * j + 3] += lv[j].exceeds - s.levels[j].exceeds;
        }
        n_above_lane[k] = (int)s.above[k].size();
        for (const LikelihoodType& a : s.above[k]) {
            if (n_pairs < max_pairs) {
                above_pairs[2 * n_pairs] = a.value;
                above_pairs[2 * n_pairs + 1] = a.tiebreaker;
            }
            n_pairs++;
        }
        uint64_t key;
        philox_rng_get(&s.rngs[k], &key, &ctr[k]);
    }
    for (size_t i = 0; i < sum.size(); i++) lv_counters[i] = sum[i];
    for (int i = 0; i < T * P; i++) {
        kima_b200::model_to_column(s.particles[i], g_layout, SL, ColView{state + i, (size_t)ld}, np[i]);
        logl[i] = s.log_likelihoods[i].value;
        tb[i] = s.log_likelihoods[i].tiebreaker;
        level[i] = (int32_t)s.level_assignments[i];
    }
    return n_pairs;
}

}  // extern "C"
