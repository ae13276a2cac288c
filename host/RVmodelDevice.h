// Device side of RVmodel: what replaces the inline calls to calculate_mu / calculate_C /
// log_likelihood (/root/reference/src/RVmodel.cpp:286-395, 170-216, 827-943) once the
// sampler scores its particles in batches.
//
// Included from host/overlay/kima.h AFTER the reference's own headers (Data.h, RVmodel.h,
// RVConditionalPrior.h) have been included with the access keywords opened, so this file can
// read RVmodel's parameter members (src/RVmodel.h:44-76).  It owns no model state: a particle
// is still an RVmodel object held by the Sampler; `stage` copies its parameters into one
// column of libkima_cuda's pinned SoA block (include/kima_cuda.h, kc_layout).
#pragma once
#include <sys/stat.h>
#include <ctime>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <functional>
#include <memory>
#include <thread>
#include <iostream>
#include <limits>
#include <type_traits>
#include <vector>
#include "kima_cuda.hpp"
#include "LanePool.h"
#include "LanePriors.h"
#include "ParticlePack.h"
#include "PriorTable.h"

// the nine model flags every kima_setup.cpp defines (src/RVmodel.h:16-34)
extern const bool GP, MA, hyperpriors, trend, multi_instrument, known_object, studentt;
extern const int degree, n_known_object;

namespace kima_b200 {

class RVmodelEvaluator {
public:
    using Model = RVmodel;

    // Two contexts on the same GPU (own stream, own pinned staging block each): while one
    // group of lanes is being scored, the host perturbs the other (BatchedSampler.h).
    static constexpr int GROUPS = 2;
    // BatchedSampler::run(): this evaluator can also run the proposals on the GPU (include/kima_sampler.h)
    typedef std::true_type has_device_sampler;

    explicit RVmodelEvaluator(int max_batch)
    {
        // one process per GPU: RANK / WORLD_SIZE / LOCAL_RANK as torchrun and mpirun-style
        // launchers export them; a single process is rank 0 of 1
        if (const char* e = std::getenv("WORLD_SIZE")) world_ = std::max(1, std::atoi(e));
        if (const char* e = std::getenv("RANK")) rank_ = std::atoi(e);
        int device = 0;
        if (const char* e = std::getenv("LOCAL_RANK")) device = std::atoi(e);
        if (const char* e = std::getenv("KIMA_CUDA_DEVICE")) device = std::atoi(e);
        const Data& d = Data::get_instance();
        std::vector<std::vector<double>> act(d.get_actind().begin(),
                                             d.get_actind().begin() + d.number_indicators);
        RVmodel probe;  // the user's constructor: fix, npmax, kernel (kima_setup.cpp)
        kc_model_flags f{};
        f.GP = GP; f.MA = MA; f.hyperpriors = hyperpriors; f.trend = trend; f.degree = degree;
        f.multi_instrument = multi_instrument; f.known_object = known_object;
        f.n_known_object = n_known_object; f.studentt = studentt;
        f.npmax = probe.npmax;
        f.kernel = (probe.kernel == RVmodel::celerite) ? 1 : 0;
        flags_ = f;
        for (int g = 0; g < GROUPS; g++) {
            ctx_[g].reset(new kimacuda::Context(device));
            // Data is uploaded once; the reference copies the whole singleton on every call
            // (src/RVmodel.cpp:173,288,829)
            ctx_[g]->upload_data(d.get_t(), d.get_y(), d.get_sig(), d.get_obsi(), d.number_instruments,
                                 act, d.M0_epoch, d.get_t_middle());
            if (const char* e = std::getenv("KIMA_CUDA_KEPLER_MODE"))
                ctx_[g]->set_option("kepler_mode", std::atoi(e));
            if (const char* e = std::getenv("KIMA_CUDA_GP_CACHE"))  // 0: no factor cache (device proposals included)
                ctx_[g]->set_option("gp_cache", std::atoi(e));
            // group 0 also scores the prior draws of initialise(), so it gets the full batch
            ctx_[g]->configure(f, g == 0 ? max_batch : (max_batch + 1) / 2);
        }
        if (verify_cache_) {
            // KIMA_CUDA_GP_CACHE_VERIFY=1: a third, cache-less context re-evaluates every cached batch
            verify_ctx_.reset(new kimacuda::Context(device));
            verify_ctx_->upload_data(d.get_t(), d.get_y(), d.get_sig(), d.get_obsi(), d.number_instruments, act, d.M0_epoch,
                                     d.get_t_middle());
            verify_ctx_->configure(f, max_batch);
        }
        if (world_ > 1) comm_bootstrap();
    }

    int rank() const { return rank_; }
    int world() const { return world_; }

    // ---- the barrier exchange between ranks (DNest4/code/SamplerImpl.h:306-329) ----------
    void allgather_bytes(const void* send, size_t nbytes, void* recv) { ctx_[0]->allgather_bytes(send, nbytes, recv); }
    void allgather_above(const double* pairs, int n, double* out, int max_total, int* counts)
    {
        ctx_[0]->allgather_above(pairs, n, out, max_total, counts);
    }
    void allreduce_levels(uint64_t* counters, int n) { ctx_[0]->allreduce_levels(counters, n); }

    // ---- a particle as a flat array of doubles, for the exchange between ranks (host/ParticlePack.h) ----
    size_t state_doubles() const { return particle_state_doubles(); }
    void pack(const RVmodel& m, double* out) const { particle_pack(m, out); }
    void unpack(RVmodel& m, const double* in) const { particle_unpack(m, in); }

    // ---- hyperpriors: one pair of planet-prior objects per lane (host/LanePriors.h says why) ----
    void reserve_lane_priors(size_t lanes) { lane_priors_.reserve(hyperpriors, lanes); }
    void own_lane_priors(RVmodel& m, size_t lane) const { lane_priors_.own(m, lane); }

    int max_batch() const { return ctx_[0]->max_batch(); }
    int max_batch(int g) const { return ctx_[g]->max_batch(); }

    // The host copies of a particle no longer need the model vector or the covariance
    // matrix (src/RVmodel.h:84,95: N long doubles and an N x N MatrixXd, both copied with
    // every proposal, SamplerImpl.h:184): nothing on the host reads them once calculate_mu /
    // calculate_C / log_likelihood run on the GPU, and perturb()'s incremental updates of mu
    // (src/RVmodel.cpp:548-804) loop over mu.size().
    void strip(RVmodel& m) const
    {
        std::vector<long double>().swap(m.mu);
        m.C.resize(0, 0);
    }

    // RVmodel::perturb() and print() copy the whole Data singleton by value on every call
    // (`auto data = Data::get_instance();`, src/RVmodel.cpp:441,444,975) although, with mu
    // gone from the host, all they still read of it is t_middle and the indicator flags: that
    // copy (three N-vectors, obsi, actind) is 3/4 of a perturb.  Once the data are on the GPU
    // and every particle has been drawn (setPriors reads the RV span), shrink the singleton's
    // arrays to what keeps those getters' values: {min, max} of t and y.
    void slim_host_data()
    {
        Data& D = Data::get_instance();
        if (D.t.size() <= 2) return;
        const double tmin = D.get_t_min(), tmax = D.get_t_max();
        const double ymin = D.get_RV_min(), ymax = D.get_RV_max();
        D.t = {tmin, tmax};
        D.y = {ymin, ymax};
        std::vector<double>().swap(D.sig);
        std::vector<int>().swap(D.obsi);
        for (auto& a : D.actind) std::vector<double>().swap(a);
    }

    // Parameters of particle `m` -> column `lane` of the staging block.
    void stage(int lane, const RVmodel& m) { stage(0, lane, m); }
    void stage(int g, int lane, const RVmodel& m)
    {
        kimacuda::Context* ctx_ = this->ctx_[g].get();
        stage_to(ctx_->layout(), ctx_->params(), (size_t)ctx_->ld(), ctx_->np(), lane, m);
    }
    // developer check (KIMA_VERIFY_STAGE=1, host/BatchedSampler.h): is column `lane` of the staging block what
    // staging particle `m` now would write?  Rows staging leaves alone compare equal by construction.
    bool check_staged(int g, int lane, const RVmodel& m, int tag)
    {
        kimacuda::Context* c = this->ctx_[g].get();
        const kc_layout& L = c->layout();
        const size_t ld = (size_t)c->ld();
        std::vector<double> col((size_t)L.n_rows);
        for (int r = 0; r < L.n_rows; r++) col[(size_t)r] = c->params()[(size_t)r * ld + lane];
        int32_t npv = c->np()[lane];
        const int32_t np_block = npv;
        std::vector<double> want = col;
        stage_to(L, want.data(), 1, &npv, 0, m);
        bool ok = npv == np_block;
        for (int r = 0; r < L.n_rows; r++)
            if (std::memcmp(&want[(size_t)r], &col[(size_t)r], sizeof(double)) != 0) {
                ok = false;
                std::fprintf(stderr, "# STAGE MISMATCH group %d lane-tag %d column %d row %d: block %.17g, particle %.17g, block now %.17g\n",
                             g, tag, lane, r, col[(size_t)r], want[(size_t)r], c->params()[(size_t)r * ld + lane]);
            }
        if (npv != np_block)
            std::fprintf(stderr, "# STAGE MISMATCH group %d lane-tag %d column %d: np block %d particle %d now %d\n", g, tag, lane,
                         (int)np_block, (int)npv, (int)c->np()[lane]);
        return ok;
    }
    void stage_to(const kc_layout& L, double* P, const size_t ld, int32_t* npv, int lane, const RVmodel& m)
    {
        auto put = [&](int row, double v) { P[(size_t)row * ld + lane] = v; };
        const auto& comps = m.planets.get_components();
        const int np = (int)comps.size();
        npv[lane] = np;
        for (int j = 0; j < np; j++) {
            put(L.row_P + j, comps[j][0]);   // P, or log P under hyperpriors (:357-360)
            put(L.row_K + j, comps[j][1]);
            put(L.row_phi + j, comps[j][2]);
            put(L.row_ecc + j, comps[j][3]);
            put(L.row_w + j, comps[j][4]);
        }
        put(L.row_vsys, m.background);
        if (multi_instrument) {
            for (int i = 0; i < L.n_jit; i++) put(L.row_jit + i, m.jitters[i]);
            for (int i = 0; i < L.n_off; i++) put(L.row_off + i, m.offsets[i]);
        } else {
            put(L.row_jit, m.extra_sigma);
        }
        if (trend) {
            put(L.row_trend + 0, m.slope);
            put(L.row_trend + 1, m.quadr);
            put(L.row_trend + 2, m.cubic);
        }
        for (int i = 0; i < L.n_beta; i++) put(L.row_beta + i, m.betas[i]);
        if (GP) {
            put(L.row_eta + 0, m.eta1);
            put(L.row_eta + 1, m.eta2);
            put(L.row_eta + 2, m.eta3);
            put(L.row_eta + 3, m.eta4);
        }
        if (studentt) put(L.row_nu, m.nu);
        if (MA) {
            put(L.row_MA + 0, m.sigmaMA);
            put(L.row_MA + 1, m.tauMA);
        }
        for (int i = 0; i < L.n_ko; i++) {
            put(L.row_KO_P + i, m.KO_P[i]);
            put(L.row_KO_K + i, m.KO_K[i]);
            put(L.row_KO_e + i, m.KO_e[i]);
            put(L.row_KO_phi + i, m.KO_phi[i]);
            put(L.row_KO_w + i, m.KO_w[i]);
        }
    }

    // logL[b] for the staged lanes [0,B): exactly what RVmodel::log_likelihood() returns,
    // +inf for a non-finite value included (src/RVmodel.cpp:938-941)
    void evaluate(int B, double* logL) { ctx_[0]->eval(B, logL); }
    // the same, split: submit returns at once, wait blocks until logL[0..B) is there
    void submit(int g, int B, double* logL)
    {
        if (trace_) trace_in_[g] = trace_hash_batch(g, B);
        if (cache_slots_[g] > 0) ctx_[g]->set_slots(B, slot_of_lane_[g].data());
        ctx_[g]->eval_async(B, logL);
        pending_B_[g] = B;
        pending_out_[g] = logL;
    }
    // Factor cache (kima_cuda.h: kc_gp_cache_enable): group g evaluates proposals of n_particles particles,
    // numbered 0 .. n_particles-1 by the caller; stage_slot names the particle behind a staged lane.  The
    // reference factors C on every log_likelihood() call (src/RVmodel.cpp:841-873) although most of
    // perturb()'s proposals leave it alone (:448-544); results are the same bits with or without the cache.
    void enable_factor_cache(int g, int n_particles)
    {
        // On unless KIMA_CUDA_GP_CACHE=0, as on the device-proposal path.
        const char* e = std::getenv("KIMA_CUDA_GP_CACHE");
        cache_slots_[g] = (e && std::atoi(e) == 0) ? 0 : ctx_[g]->gp_cache_enable(n_particles);
        if (cache_slots_[g] > 0) slot_of_lane_[g].assign((size_t)ctx_[g]->max_batch(), -1);
    }
    void stage_slot(int g, int lane, int particle)
    {
        if (cache_slots_[g] > 0) slot_of_lane_[g][(size_t)lane] = particle < cache_slots_[g] ? particle : -1;
    }
    void wait(int g)
    {
        ctx_[g]->wait();
        if (trace_ && pending_B_[g] > 0) {
            // developer trace (KIMA_TRACE_HASH=file): one line per batch, hashes of what went in and what came back
            uint64_t h = 1469598103934665603ull;
            trace_mix(h, pending_out_[g], sizeof(double) * (size_t)pending_B_[g]);
            std::fprintf(trace_, "%llu %d %d %016llx %016llx\n", (unsigned long long)trace_seq_++, g, pending_B_[g],
                         (unsigned long long)trace_in_[g], (unsigned long long)h);
        }
        if (verify_cache_ && cache_slots_[g] > 0 && pending_B_[g] > 0) {
            // KIMA_CUDA_GP_CACHE_VERIFY=1: the same staged batch once more without the cache, compared bit for bit
            // on a context of its own: nothing is added to the stream of the context under test
            const int B = pending_B_[g];
            const size_t n_rows = (size_t)ctx_[g]->layout().n_rows;
            for (size_t r = 0; r < n_rows; r++)
                std::memcpy(verify_ctx_->params() + r * (size_t)verify_ctx_->ld(), ctx_[g]->params() + r * (size_t)ctx_[g]->ld(),
                            sizeof(double) * (size_t)B);
            std::memcpy(verify_ctx_->np(), ctx_[g]->np(), sizeof(int32_t) * (size_t)B);
            std::vector<double> again((size_t)B);
            verify_ctx_->eval(B, again.data());
            for (int i = 0; i < pending_B_[g]; i++)
                if (std::memcmp(&again[(size_t)i], &pending_out_[g][i], sizeof(double)) != 0)
                    std::fprintf(stderr, "# factor cache MISMATCH group %d lane %d particle %d: cached %.17g uncached %.17g\n", g, i,
                                 slot_of_lane_[g][(size_t)i], pending_out_[g][i], again[(size_t)i]);
        }
        pending_B_[g] = 0;
    }

    long long launches() const { return ctx_[0]->launch_count() + ctx_[1]->launch_count(); }

    // ---- device-side proposals (include/kima_sampler.h) ------------
    // The particles move to the GPU (context 0) and stay there: perturb, prior-ratio test,
    // likelihood, accept, level assignment, level counters and `above` all run in HBM; the host
    // keeps DNest4's bookkeeping between blocks of thread_steps steps.
    //
    // "" when every prior of `proto` (a particle on which setPriors() has run) has a device
    // counterpart and the sampler was created; otherwise the name of the first one that has not
    // (Empirical, mixGaussianLogUniform, Gaussian_from_Tc, Jeffreys: host proposals only).
    std::string sampler_create(const RVmodel& proto, const kc_sampler_config& cfg)
    {
        kc_prior_table T;
        const std::string bad = extract_prior_table(proto, T, &kc_prior_make);
        if (!bad.empty()) return bad;
        ctx_[0]->sampler_create(T, cfg);
        return "";
    }
    struct ColRef {
        double* base;
        size_t ld;
        double& operator()(int row) const { return base[(size_t)row * ld]; }
    };
    // the particles [i0, i1) (this rank's lanes) with likelihoods and level assignments -> HBM;
    // the arrays are indexed like `particles`
    void sampler_upload(const std::vector<RVmodel>& particles, size_t i0, size_t i1, const std::vector<double>& logl,
                        const std::vector<double>& tb, const std::vector<int32_t>& level)
    {
        const kc_sampler_layout& SL = ctx_[0]->sampler_layout();
        const size_t n = i1 - i0;
        std::vector<double> state((size_t)SL.n_state * n, 0.0);
        std::vector<int32_t> np(n);
        for_range(n, [&](size_t a, size_t b) {
            for (size_t i = a; i < b; i++)
                model_to_column(particles[i0 + i], ctx_[0]->layout(), SL, ColRef{state.data() + i, n}, np[i]);
        });
        ctx_[0]->sampler_upload(state.data(), np.data(), logl.data() + i0, tb.data() + i0, level.data() + i0);
    }
    int sampler_state_rows() const { return ctx_[0]->sampler_layout().n_state; }
    // one column (n_state doubles) <-> a host RVmodel
    void column_into(RVmodel& m, const double* col, int32_t np) const
    {
        column_to_model(m, ctx_[0]->layout(), ctx_[0]->sampler_layout(), ColRef{const_cast<double*>(col), 1}, np);
    }
    // local particle `which` back into a host RVmodel (save_particle, the record holder)
    void sampler_get_particle(int which, RVmodel& m)
    {
        std::vector<double> col(sampler_state_rows());
        int32_t np = 0;
        ctx_[0]->sampler_get_particle(which, col.data(), &np);
        column_into(m, col.data(), np);
    }
    // all local particles: state [n_state][n_local] and np
    void sampler_local_state(std::vector<double>& state, std::vector<int32_t>& np)
    {
        const kc_sampler_layout& SL = ctx_[0]->sampler_layout();
        state.resize((size_t)SL.n_state * SL.n_particles);
        np.resize(SL.n_particles);
        ctx_[0]->sampler_download(state.data(), np.data(), nullptr, nullptr, nullptr);
    }
    void columns_into(std::vector<RVmodel>& particles, size_t i0, const std::vector<double>& state,
                      const std::vector<int32_t>& np)
    {
        const size_t n = np.size();
        for_range(n, [&](size_t a, size_t b) {
            for (size_t i = a; i < b; i++)
                column_to_model(particles[i0 + i], ctx_[0]->layout(), ctx_[0]->sampler_layout(),
                                ColRef{const_cast<double*>(state.data()) + i, n}, np[i]);
        });
    }
    // the per-particle conversions above are independent: the sampler lends its host threads
    void set_host_pool(std::shared_ptr<kima_b200::LanePool> p) { host_pool_ = std::move(p); }
    kimacuda::Context& sampler_ctx() { return *ctx_[0]; }

private:
    std::shared_ptr<kima_b200::LanePool> host_pool_;
    LanePriors lane_priors_;
    void for_range(size_t n, const std::function<void(size_t, size_t)>& f)
    {
        if (host_pool_ && n >= 1024)
            host_pool_->run(0, (int)n, [&](int a, int b) { f((size_t)a, (size_t)b); });
        else
            f(0, n);
    }
    // NCCL's unique id travels through a file in the run directory (all ranks of one node
    // share it): rank 0 writes it atomically, the others wait for it.  A file left behind by a
    // killed run must not be taken for this run's: the name carries the launcher's run id when
    // there is one (KIMA_RUN_NONCE, else TORCHELASTIC_RUN_ID), rank 0 unlinks before it writes,
    // and the other ranks only accept a file written after their own start (minus a margin for
    // ranks that start late).
    void comm_bootstrap()
    {
        const char* tag = std::getenv("KIMA_COMM_ID_FILE");
        std::string path = tag ? tag : ".kima_nccl_id";
        if (!tag) {
            if (const char* port = std::getenv("MASTER_PORT")) path += std::string(".") + port;
            const char* nonce = std::getenv("KIMA_RUN_NONCE");
            if (!nonce) nonce = std::getenv("TORCHELASTIC_RUN_ID");
            if (nonce) path += std::string(".") + nonce;
        }
        const time_t started = std::time(nullptr);
        unsigned char id[128];
        if (rank_ == 0) {
            std::remove(path.c_str());
            kimacuda::Context::unique_id(id);
            const std::string tmp = path + ".tmp";
            {
                std::ofstream f(tmp, std::ios::binary | std::ios::trunc);
                f.write((const char*)id, 128);
            }
            std::rename(tmp.c_str(), path.c_str());
        } else {
            for (int tries = 0;; tries++) {
                struct stat sb;
                if (::stat(path.c_str(), &sb) == 0 && sb.st_mtime + 300 >= started) {
                    std::ifstream f(path, std::ios::binary);
                    if (f && f.read((char*)id, 128)) break;
                }
                if (tries > 6000) throw std::runtime_error("timed out waiting for " + path);
                std::this_thread::sleep_for(std::chrono::milliseconds(10));
            }
        }
        ctx_[0]->comm_init(rank_, world_, id);
        // everybody has read the id once this completes: rank 0 can remove the file
        unsigned char token = 1, all[1024];
        ctx_[0]->allgather_bytes(&token, 1, all);
        if (rank_ == 0) std::remove(path.c_str());
    }

    int rank_ = 0, world_ = 1;
    std::unique_ptr<kimacuda::Context> ctx_[GROUPS];
    int cache_slots_[GROUPS] = {0, 0};               // active factor-cache slots per group
    std::unique_ptr<kimacuda::Context> verify_ctx_;
    // developer trace of every batch (KIMA_TRACE_HASH=file)
    FILE* trace_ = std::getenv("KIMA_TRACE_HASH") ? std::fopen(std::getenv("KIMA_TRACE_HASH"), "w") : nullptr;
    uint64_t trace_seq_ = 0, trace_in_[GROUPS] = {0, 0};
    static void trace_mix(uint64_t& h, const void* p, size_t n)
    {
        const unsigned char* b = static_cast<const unsigned char*>(p);
        size_t i = 0;
        for (; i + 8 <= n; i += 8) {
            uint64_t w;
            std::memcpy(&w, b + i, 8);
            h = (h ^ w) * 1099511628211ull;
            h ^= h >> 29;
        }
        for (; i < n; i++) h = (h ^ b[i]) * 1099511628211ull;
    }
    uint64_t trace_hash_batch(int g, int B)
    {
        uint64_t h = 1469598103934665603ull;
        const size_t n_rows = (size_t)ctx_[g]->layout().n_rows, ld = (size_t)ctx_[g]->ld();
        for (size_t r = 0; r < n_rows; r++) trace_mix(h, ctx_[g]->params() + r * ld, sizeof(double) * (size_t)B);
        trace_mix(h, ctx_[g]->np(), sizeof(int32_t) * (size_t)B);
        return h;
    }
    int pending_B_[GROUPS] = {0, 0};
    double* pending_out_[GROUPS] = {nullptr, nullptr};
    bool verify_cache_ = std::getenv("KIMA_CUDA_GP_CACHE_VERIFY") && std::atoi(std::getenv("KIMA_CUDA_GP_CACHE_VERIFY")) != 0;
    std::vector<int32_t> slot_of_lane_[GROUPS];      // particle behind every staged lane
    kc_model_flags flags_{};
};

}  // namespace kima_b200
