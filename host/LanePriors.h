// One pair of planet-prior objects per lane, for models with hyperpriors.
//
// With hyperpriors the reference re-parameterises the period and amplitude priors before every use
// (Pprior->setpars(center, width); Kprior->setpars(muK): src/RVConditionalPrior.cpp:111-112, 135-136,
// 150-151) -- and those objects are shared: every copy of a particle (SamplerImpl.h:184, 208, and
// kill_lagging_particles' copies between threads, :487) aliases the shared_ptr of its source
// (RVConditionalPrior.h:30-32).  Two host threads that perturb descendants of one particle then race
// between one's setpars() and the other's cdf(): a period drawn with another particle's centre.  The
// reference's own -t threads have that race too; in the lock-step sampler it made one run in seven at
// 4096 lanes and 16 host threads differ from the next.  Each lane therefore gets objects of its own
// (default-constructed as the reference's are, src/RVConditionalPrior.cpp:43-44; their state is
// overwritten by setpars before every read), and a proposal is pointed at its lane's pair before
// perturb() runs: the values are those of a serial run.
//
// Include after the reference's headers (RVmodel.h with its private section opened).
#pragma once
#include <memory>
#include <vector>

namespace kima_b200 {

class LanePriors {
public:
    // call once, from one thread, before the lanes are stepped
    void reserve(bool hyperpriors, size_t lanes)
    {
        if (!hyperpriors || P_.size() == lanes) return;
        P_.resize(lanes);
        K_.resize(lanes);
        for (size_t k = 0; k < lanes; k++) {
            P_[k] = std::make_shared<DNest4::Laplace>();
            K_[k] = std::make_shared<DNest4::Exponential>();
        }
    }
    // lane `lane`'s thread only
    void own(RVmodel& m, size_t lane) const
    {
        if (lane >= P_.size()) return;
        m.planets.conditional_prior.Pprior = P_[lane];
        m.planets.conditional_prior.Kprior = K_[lane];
    }

private:
    std::vector<std::shared_ptr<DNest4::ContinuousDistribution>> P_, K_;
};

}  // namespace kima_b200
