// C++ face of libkima_cuda's C ABI (include/kima_cuda.h): RAII context, exceptions instead
// of status codes.  Header-only; this is what the RVmodel / Sampler shim talks to.
//
// Replaces, for a whole batch of particles per call, the three private methods the reference
// runs per particle:  RVmodel::calculate_mu / calculate_C / log_likelihood
// (/root/reference/src/RVmodel.cpp:286-395, 170-216, 827-943).
#pragma once
#include <cstdint>
#include <stdexcept>
#include <string>
#include <vector>
#include "kima_cuda.h"
#include "kima_sampler.h"

namespace kimacuda {

struct Error : std::runtime_error {
    int code;
    Error(int c, const std::string& what) : std::runtime_error(what), code(c) {}
};

inline void check(int rc)
{
    if (rc != KC_OK) throw Error(rc, std::string("libkima_cuda: ") + kc_last_error());
}

// One context = one GPU = one data set + one set of model flags.  Not copyable.
class Context {
public:
    explicit Context(int device = 0) { check(kc_create(device, &ctx_)); }
    ~Context() { if (ctx_) kc_destroy(ctx_); }
    Context(const Context&) = delete;
    Context& operator=(const Context&) = delete;

    // Data::get_t/get_y/get_sig/get_obsi/get_actind, M0_epoch, get_t_middle (src/Data.h:52-108)
    void upload_data(const std::vector<double>& t, const std::vector<double>& y,
                     const std::vector<double>& sig, const std::vector<int>& obsi, int n_inst,
                     const std::vector<std::vector<double>>& actind, double M0_epoch,
                     double t_middle)
    {
        const int N = (int)t.size();
        std::vector<int32_t> ob(obsi.begin(), obsi.end());
        std::vector<double> act;
        for (const auto& a : actind) act.insert(act.end(), a.begin(), a.end());
        check(kc_upload_data(ctx_, N, t.data(), y.data(), sig.data(),
                             n_inst > 1 ? ob.data() : nullptr, n_inst, (int)actind.size(),
                             actind.empty() ? nullptr : act.data(), M0_epoch, t_middle));
    }

    void configure(const kc_model_flags& f, int max_batch)
    {
        check(kc_configure(ctx_, &f, max_batch));
        check(kc_get_layout(ctx_, &layout_));
        check(kc_host_params(ctx_, &params_, &np_, &ld_));
        max_batch_ = max_batch;
    }

    const kc_layout& layout() const { return layout_; }
    int ld() const { return ld_; }
    int max_batch() const { return max_batch_; }
    // pinned staging block: params()[row * ld() + lane], np()[lane]
    double* params() { return params_; }
    int32_t* np() { return np_; }

    // H2D of lanes [0,B), kernels, D2H: logL[b] as RVmodel::log_likelihood() returns it
    void eval(int B, double* logL) { check(kc_eval(ctx_, B, logL)); }
    void eval_async(int B, double* logL) { check(kc_eval_async(ctx_, B, logL)); }
    void wait() { check(kc_wait(ctx_)); }
    void set_option(const char* key, int value) { check(kc_set_option(ctx_, key, value)); }
    // factor cache of the shared-memory-resident GP path (kima_cuda.h: kc_gp_cache_enable): returns the
    // active slot count, 0 where the cache does not apply or does not fit
    int gp_cache_enable(int n_slots)
    {
        if (kc_gp_cache_enable(ctx_, n_slots) != KC_OK) return 0;  // does not fit: evaluate without it
        int n = 0;
        check(kc_gp_cache_slots(ctx_, &n));
        return n;
    }
    void set_slots(int B, const int32_t* slots) { check(kc_set_slots(ctx_, B, slots)); }
    long long launch_count() const { return (long long)kc_launch_count(ctx_); }

    // posterior-predictive curves of the first B staged parameter vectors
    // (pykima/results.py:723-855); t_pred == nullptr means the data epochs
    void predict(int B, int nt, const double* t_pred, int what, double* model_out, double* gp_out)
    {
        check(kc_predict(ctx_, B, nt, t_pred, what, model_out, gp_out));
    }

    // one process per GPU: the level-creation barrier (DNest4/code/SamplerImpl.h:306-329)
    static void unique_id(void* id128) { check(kc_comm_unique_id(id128)); }
    void comm_init(int rank, int nranks, const void* id128) { check(kc_comm_init(ctx_, rank, nranks, id128)); }
    void allgather_bytes(const void* send, size_t nbytes, void* recv) { check(kc_allgather_bytes(ctx_, send, nbytes, recv)); }
    void allgather_above(const double* pairs, int n, double* out, int max_total, int* counts)
    {
        check(kc_allgather_above(ctx_, pairs, n, out, max_total, counts));
    }
    void allreduce_levels(uint64_t* counters, int n) { check(kc_allreduce_levels(ctx_, counters, n)); }
    kc_ctx* raw() { return ctx_; }

    // ---- device-side proposals (include/kima_sampler.h) -------------------------------------
    void sampler_create(const kc_prior_table& priors, const kc_sampler_config& cfg)
    {
        check(kc_sampler_create(ctx_, &priors, &cfg));
        check(kc_sampler_get_layout(ctx_, &sampler_layout_));
    }
    const kc_sampler_layout& sampler_layout() const { return sampler_layout_; }
    void sampler_upload(const double* state, const int32_t* np, const double* logl, const double* tb, const int32_t* level)
    {
        check(kc_sampler_upload(ctx_, state, np, logl, tb, level));
    }
    void sampler_download(double* state, int32_t* np, double* logl, double* tb, int32_t* level)
    {
        check(kc_sampler_download(ctx_, state, np, logl, tb, level));
    }
    void sampler_get_particle(int which, double* col, int32_t* np) { check(kc_sampler_get_particle(ctx_, which, col, np)); }
    void sampler_set_particle(int which, const double* col, int32_t np, double logl, double tb, int32_t level)
    {
        check(kc_sampler_set_particle(ctx_, which, col, np, logl, tb, level));
    }
    void sampler_copy_particles(int n, const int32_t* dst, const int32_t* src) { check(kc_sampler_copy_particles(ctx_, n, dst, src)); }
    void sampler_set_levels(int L, const double* logl, const double* tb, const double* log_X, const uint64_t* tries,
                            bool enough, bool at_max)
    {
        check(kc_sampler_set_levels(ctx_, L, logl, tb, log_X, tries, enough, at_max));
    }
    void sampler_run(int steps) { check(kc_sampler_run(ctx_, steps)); }
    // `all_above` kept in HBM
    void sampler_collect_pool(uint64_t* counters, int64_t* n_new, int64_t* n_pool)
    {
        check(kc_sampler_collect_pool(ctx_, counters, n_new, n_pool));
    }
    void sampler_pool_select(int64_t index, double* value, double* tb, int64_t* n_pool)
    {
        check(kc_sampler_pool_select(ctx_, index, value, tb, n_pool));
    }
    void sampler_pool_clear() { check(kc_sampler_pool_clear(ctx_)); }
    int sampler_collect(uint64_t* counters, double* above_pairs, int max_pairs)
    {
        int n = 0;
        check(kc_sampler_collect(ctx_, counters, above_pairs, max_pairs, &n));
        return n;
    }

private:
    kc_ctx* ctx_ = nullptr;
    kc_layout layout_{};
    double* params_ = nullptr;
    int32_t* np_ = nullptr;
    int ld_ = 0, max_batch_ = 0;
    kc_sampler_layout sampler_layout_{};
};

}  // namespace kimacuda
