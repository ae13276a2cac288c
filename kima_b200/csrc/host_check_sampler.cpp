// Host instantiation of kc_sampler_logic.cuh, for CPU-side unit tests only
// (tests/test_sampler_logic.py): the per-lane proposal / accept logic of the device-side sampler
// run on host arrays, with the likelihoods of the proposals SUPPLIED BY THE CALLER.  There is
// deliberately no likelihood in here: this cannot score a particle, it is not a CPU fallback.
#include <cstring>
#include <vector>
#include "kc_sampler_logic.cuh"

namespace {

using namespace kc::smp;

struct HostSampler {
    SamplerParams sp{};
    kc_prior_table pri{};
    std::vector<double> S, Q, logl, tb, logl_q, above, lv_logl, lv_tb, lv_logX;
    std::vector<int32_t> np, level, npq, which, flags, n_above;
    std::vector<uint64_t> ctr, lv_tries;
    std::vector<uint32_t> cnt;
    void bind()
    {
        sp.pri = &pri;
        sp.S = S.data(); sp.np = np.data(); sp.logl = logl.data(); sp.tb = tb.data(); sp.level = level.data();
        sp.Q = Q.data(); sp.npq = npq.data(); sp.logl_q = logl_q.data();
        sp.ctr = ctr.data(); sp.which = which.data(); sp.flags = flags.data();
        sp.above = above.data(); sp.n_above = n_above.data();
        sp.lv_logl = lv_logl.data(); sp.lv_tb = lv_tb.data(); sp.lv_logX = lv_logX.data();
        sp.lv_tries = lv_tries.data(); sp.cnt = cnt.data();
    }
};

}  // namespace

extern "C" {

void* hc_smp_create(const kc_model_flags* f, int n_inst, int n_ind, const kc_prior_table* pri,
                    const kc_sampler_config* cfg)
{
    HostSampler* h = new HostSampler();
    SamplerParams& sp = h->sp;
    sp.F = *f;
    kc_layout_from_flags(f, n_inst, n_ind, &sp.L);
    const int T = cfg->lanes, P = cfg->num_particles, n = T * P;
    sp.SL.row_u = sp.L.n_rows;
    sp.SL.row_hyper = sp.L.n_rows + 5 * sp.L.npmax;
    sp.SL.n_state = sp.SL.row_hyper + 3;
    sp.SL.n_particles = n;
    sp.n_ind = n_ind;
    h->pri = *pri;
    sp.ldp = n; sp.ldq = T; sp.ldl = T;
    sp.lanes = T; sp.P = P;
    sp.max_steps = cfg->max_steps;
    sp.Lcap = cfg->max_levels;
    sp.lambda = cfg->lambda; sp.beta = cfg->beta;
    sp.key0 = cfg->seed + (uint64_t)cfg->lane0;
    h->S.assign((size_t)sp.SL.n_state * n, 0.0);
    h->Q.assign((size_t)sp.SL.n_state * T, 0.0);
    h->np.assign(n, 0); h->logl.assign(n, 0.0); h->tb.assign(n, 0.5); h->level.assign(n, 0);
    h->npq.assign(T, 0); h->logl_q.assign(T, 0.0); h->which.assign(T, 0); h->flags.assign(T, 0);
    h->ctr.assign(T, 0);
    h->above.assign((size_t)T * sp.max_steps * 2, 0.0);
    h->n_above.assign(T, 0);
    h->lv_logl.assign(sp.Lcap, 0.0); h->lv_tb.assign(sp.Lcap, 0.0); h->lv_logX.assign(sp.Lcap, 0.0);
    h->lv_tries.assign(sp.Lcap, 0);
    h->cnt.assign((size_t)4 * sp.Lcap * T, 0u);
    h->bind();
    return h;
}
void hc_smp_destroy(void* p) { delete (HostSampler*)p; }
void hc_smp_layout(void* p, kc_sampler_layout* out) { *out = ((HostSampler*)p)->sp.SL; }

void hc_smp_upload(void* p, const double* state, const int32_t* np, const double* logl, const double* tb,
                   const int32_t* level)
{
    HostSampler* h = (HostSampler*)p;
    const size_t n = h->np.size();
    if (state) std::memcpy(h->S.data(), state, sizeof(double) * h->S.size());
    if (np) std::memcpy(h->np.data(), np, sizeof(int32_t) * n);
    if (logl) std::memcpy(h->logl.data(), logl, sizeof(double) * n);
    if (tb) std::memcpy(h->tb.data(), tb, sizeof(double) * n);
    if (level) std::memcpy(h->level.data(), level, sizeof(int32_t) * n);
}
void hc_smp_download(void* p, double* state, int32_t* np, double* logl, double* tb, int32_t* level)
{
    HostSampler* h = (HostSampler*)p;
    const size_t n = h->np.size();
    if (state) std::memcpy(state, h->S.data(), sizeof(double) * h->S.size());
    if (np) std::memcpy(np, h->np.data(), sizeof(int32_t) * n);
    if (logl) std::memcpy(logl, h->logl.data(), sizeof(double) * n);
    if (tb) std::memcpy(tb, h->tb.data(), sizeof(double) * n);
    if (level) std::memcpy(level, h->level.data(), sizeof(int32_t) * n);
}
void hc_smp_set_levels(void* p, int L, const double* logl, const double* tb, const double* logX,
                       const uint64_t* tries, int enough, int at_max)
{
    HostSampler* h = (HostSampler*)p;
    h->sp.n_levels = L;
    for (int j = 0; j < L; j++) {
        h->lv_logl[j] = logl[j]; h->lv_tb[j] = tb[j]; h->lv_logX[j] = logX[j]; h->lv_tries[j] = tries[j];
    }
    h->sp.enough = enough;
    h->sp.at_max = at_max;
    std::fill(h->cnt.begin(), h->cnt.end(), 0u);
    std::fill(h->n_above.begin(), h->n_above.end(), 0);
}
void hc_smp_propose(void* p)
{
    HostSampler* h = (HostSampler*)p;
    for (int k = 0; k < h->sp.lanes; k++) step_propose(h->sp, k);
}
// the proposals' columns [n_state][lanes], their np and the step flags (bit 1 = scored)
void hc_smp_get_proposals(void* p, double* Q, int32_t* npq, int32_t* flags)
{
    HostSampler* h = (HostSampler*)p;
    if (Q) std::memcpy(Q, h->Q.data(), sizeof(double) * h->Q.size());
    if (npq) std::memcpy(npq, h->npq.data(), sizeof(int32_t) * h->npq.size());
    if (flags) std::memcpy(flags, h->flags.data(), sizeof(int32_t) * h->flags.size());
}
void hc_smp_accept(void* p, const double* logl_q)
{
    HostSampler* h = (HostSampler*)p;
    std::memcpy(h->logl_q.data(), logl_q, sizeof(double) * h->logl_q.size());
    for (int k = 0; k < h->sp.lanes; k++) step_accept(h->sp, k);
}
int hc_smp_collect(void* p, uint64_t* counters, double* above_pairs, int max_pairs, int32_t* n_above_lane)
{
    HostSampler* h = (HostSampler*)p;
    const SamplerParams& sp = h->sp;
    if (counters)
        for (int j = 0; j < sp.n_levels; j++)
            for (int c = 0; c < 4; c++) {
                uint64_t s = 0;
                for (int k = 0; k < sp.lanes; k++) s += counter(sp, c, j, k);
                counters[4 * j + c] = s;
            }
    int n = 0;
    for (int k = 0; k < sp.lanes; k++) {
        if (n_above_lane) n_above_lane[k] = h->n_above[k];
        for (int i = 0; i < h->n_above[k]; i++) {
            if (above_pairs && n < max_pairs) {
                above_pairs[2 * n] = sp.above[((size_t)k * sp.max_steps + i) * 2];
                above_pairs[2 * n + 1] = sp.above[((size_t)k * sp.max_steps + i) * 2 + 1];
            }
            n++;
        }
    }
    return n;
}
void hc_smp_counters(void* p, uint64_t* ctr, int set)
{
    HostSampler* h = (HostSampler*)p;
    if (set) std::memcpy(h->ctr.data(), ctr, sizeof(uint64_t) * h->ctr.size());
    else std::memcpy(ctr, h->ctr.data(), sizeof(uint64_t) * h->ctr.size());
}
// RVmodel::perturb alone (num_particles must be 1): proposals' columns, np and logH per lane
void hc_smp_perturb(void* p, double* state, int32_t* np, double* logH)
{
    HostSampler* h = (HostSampler*)p;
    const SamplerParams& sp = h->sp;
    for (int k = 0; k < sp.lanes; k++) {
        Rng rng{sp.key0 + (uint64_t)k, sp.ctr[k]};
        const Col q{sp.Q + k, sp.ldq};
        for (int r = 0; r < sp.SL.n_state; r++) q(r) = sp.S[(size_t)r * sp.ldp + k];
        int n = sp.np[k];
        logH[k] = perturb_model(sp, q, n, rng);
        sp.npq[k] = n;
        sp.ctr[k] = rng.ctr;
    }
    std::memcpy(state, h->Q.data(), sizeof(double) * h->Q.size());
    std::memcpy(np, h->npq.data(), sizeof(int32_t) * h->npq.size());
}
int hc_prior_make(int type, const double* p, int n_p, kc_prior* out) { return prior_make(type, p, n_p, out); }
void hc_prior_eval(const kc_prior* q, int what, int n, const double* x, double* out)
{
    for (int i = 0; i < n; i++)
        out[i] = what == 0 ? prior_cdf(*q, x[i]) : (what == 1 ? prior_cdf_inverse(*q, x[i]) : prior_log_pdf(*q, x[i]));
}
uint64_t hc_rng_draws(uint64_t key, uint64_t ctr0, int what, int n, double* out)
{
    Rng r{key, ctr0};
    for (int i = 0; i < n; i++)
        out[i] = what == 0 ? r.rand() : (what == 1 ? r.randn() : (what == 2 ? r.randt2() : r.randh()));
    return r.ctr;
}
}
