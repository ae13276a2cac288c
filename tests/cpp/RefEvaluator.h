// TEST INFRASTRUCTURE, not a product path: an Evaluator for host/BatchedSampler.h that scores the staged
// particles with the reference's own RVmodel::log_likelihood() (src/RVmodel.cpp:827-943) on the host cores.
// With it the lock-step sampler -- phase A / phase B around the likelihood call, the two lane groups, the lane
// pool, the level bookkeeping -- runs on a machine without a GPU, and must write the files of the reference's
// sampler byte for byte (tests/test_host_sampler_cpu.py).  Built only by that test, from the reference's
// UNWEAKENED objects: perturb() keeps mu and C up to date as in the reference, nothing is stripped.
//
// The Evaluator concept is described at the top of host/BatchedSampler.h.
#pragma once
#include <cstdio>
#include <cstdlib>
#include <memory>
#include <type_traits>
#include <vector>
#include "kima_cuda.hpp"   // declarations only: BatchedSampler::run_device names these types (never instantiated here)
#include "LanePool.h"
#include "LanePriors.h"

extern const bool hyperpriors;

namespace kima_b200 {

class RefEvaluator {
public:
    static constexpr int GROUPS = 2;
    typedef std::false_type has_device_sampler;   // host proposals only: run_device() is never instantiated

    explicit RefEvaluator(int max_batch) : cap_(max_batch)
    {
        for (auto& s : staged_) s.assign((size_t)max_batch, nullptr);
    }
    int max_batch() const { return cap_; }
    int max_batch(int) const { return cap_; }
    int rank() const { return 0; }
    int world() const { return 1; }
    long long launches() const { return 0; }
    void set_host_pool(std::shared_ptr<LanePool> p) { pool_ = std::move(p); }

    // the CPU likelihood reads mu, C and the data: nothing is dropped
    void strip(RVmodel&) const {}
    void slim_host_data() {}

    void stage(int lane, const RVmodel& m) { stage(0, lane, m); }
    void stage(int g, int lane, const RVmodel& m) { staged_[g][(size_t)lane] = &m; }
    void stage_slot(int, int, int) {}
    void enable_factor_cache(int, int) {}
    bool check_staged(int, int, const RVmodel&, int) { return true; }

    void evaluate(int B, double* logL) { score(0, B, logL); }
    void submit(int g, int B, double* logL) { score(g, B, logL); }   // blocking: the result is there on return
    void wait(int) {}

    void reserve_lane_priors(size_t lanes) { lane_priors_.reserve(hyperpriors, lanes); }
    void own_lane_priors(RVmodel& m, size_t lane) const
    {
        // KIMA_SHARED_PRIORS=1: leave the reference's shared objects in place (to show the race they carry)
        static const bool shared = std::getenv("KIMA_SHARED_PRIORS") != nullptr;
        if (!shared) lane_priors_.own(m, lane);
    }

    // one process only: the exchange between ranks is never reached
    size_t state_doubles() const { return 0; }
    void pack(const RVmodel&, double*) const { unreachable(); }
    void unpack(RVmodel&, const double*) const { unreachable(); }
    void allgather_bytes(const void*, size_t, void*) { unreachable(); }
    void allgather_above(const double*, int, double*, int, int*) { unreachable(); }
    void allreduce_levels(uint64_t*, int) { unreachable(); }

private:
    static void unreachable()
    {
        std::fprintf(stderr, "RefEvaluator: multi-process entry point reached\n");
        std::abort();
    }
    void score(int g, int B, double* logL)
    {
        const std::vector<const RVmodel*>& s = staged_[g];
        auto body = [&](int a, int b) {
            for (int i = a; i < b; ++i) logL[i] = s[(size_t)i]->log_likelihood();
        };
        if (pool_) pool_->run(0, B, body);
        else body(0, B);
    }
    int cap_;
    std::vector<const RVmodel*> staged_[GROUPS];
    std::shared_ptr<LanePool> pool_;
    LanePriors lane_priors_;
};

}  // namespace kima_b200
