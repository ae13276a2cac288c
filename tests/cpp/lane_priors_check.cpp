// Unit check of host/LanePriors.h against the reference's own classes: copies of a particle share the
// period / amplitude prior objects that the hyperprior code re-parameterises before every use
// (src/RVConditionalPrior.cpp:111-112, 135-136, 150-151; RVConditionalPrior.h:30-32), LanePriors::own gives
// each lane a pair of its own, and the values drawn through them are those of the shared objects used serially.
// Built by tests/test_host_sampler_cpu.py from the reference's unmodified objects.
// every header the reference's headers pull in, before the access keywords are opened (as host/overlay/kima.h)
#include <algorithm>
#include <atomic>
#include <chrono>
#include <cmath>
#include <condition_variable>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <functional>
#include <iomanip>
#include <iostream>
#include <limits>
#include <map>
#include <memory>
#include <mutex>
#include <numeric>
#include <random>
#include <set>
#include <sstream>
#include <stdexcept>
#include <string>
#include <thread>
#include <typeinfo>
#include <vector>
#include <Eigen/Core>
#include <Eigen/Dense>
#include <Eigen/Cholesky>
#include "celerite/celerite.h"
#define private public
#define protected public
#include "DNest4.h"
#include "Data.h"
#include "RVConditionalPrior.h"
#include "RVmodel.h"
#undef private
#undef protected
#include "LanePriors.h"

const bool GP = false, MA = false, hyperpriors = true, trend = false, multi_instrument = false, known_object = false,
           studentt = false;
const int degree = 0, n_known_object = 0;
RVmodel::RVmodel() : fix(false), npmax(3) {}

#define CHECK(c) do { if (!(c)) { std::printf("FAILED line %d: %s\n", __LINE__, #c); return 1; } } while (0)

int main(int argc, char** argv)
{
    // RVmodel sizes its arrays from the data: argv[1] is an RV file (tests/golden/e2e/BL2009_dataset2.kms.rv)
    if (argc < 2) { std::printf("usage: lane_priors_check <rv file>\n"); return 2; }
    Data::get_instance().load(argv[1], "kms", 2);
    RVmodel a;                       // default priors: Laplace / Exponential under hyperpriors
    RVmodel b = a;                   // what every proposal and every kill_lagging copy does
    auto& ca = a.planets.conditional_prior;
    auto& cb = b.planets.conditional_prior;
    CHECK(ca.Pprior.get() == cb.Pprior.get());   // the root of the race: one object behind both particles
    CHECK(ca.Kprior.get() == cb.Kprior.get());
    ca.center = 3.0; ca.width = 0.5; ca.muK = 2.0;
    cb.center = 5.5; cb.width = 1.5; cb.muK = 9.0;
    // serial use of the shared objects: the values to reproduce
    std::vector<double> ua = {0.3, 0.6, 0.1, 0.2, 0.4}, ub = ua, va = ua, vb = ua;
    ca.from_uniform(ua);
    cb.from_uniform(ub);
    // what a thread switch between setpars() and cdf_inverse() does with shared objects
    ca.Pprior->setpars(ca.center, ca.width);
    cb.Pprior->setpars(cb.center, cb.width);      // the other thread gets there first
    CHECK(ca.Pprior->cdf_inverse(0.3) != ua[0]);  // a's period drawn with b's centre and width
    kima_b200::LanePriors lp;
    lp.reserve(true, 2);
    lp.own(a, 0);
    lp.own(b, 1);
    CHECK(ca.Pprior.get() != cb.Pprior.get());
    CHECK(ca.Kprior.get() != cb.Kprior.get());
    ca.from_uniform(va);
    cb.from_uniform(vb);
    for (int i = 0; i < 5; i++) CHECK(va[i] == ua[i] && vb[i] == ub[i]);
    // the interleaving again: now harmless
    ca.Pprior->setpars(ca.center, ca.width);
    cb.Pprior->setpars(cb.center, cb.width);
    CHECK(ca.Pprior->cdf_inverse(0.3) == ua[0]);
    // a copy made from a lane's particle carries the lane's objects until its own lane claims it
    RVmodel c = a;
    CHECK(c.planets.conditional_prior.Pprior.get() == ca.Pprior.get());
    lp.own(c, 1);
    CHECK(c.planets.conditional_prior.Pprior.get() == cb.Pprior.get());
    // without hyperpriors nothing is re-parameterised and nothing is replaced
    kima_b200::LanePriors none;
    none.reserve(false, 2);
    RVmodel d = a;
    none.own(d, 0);
    CHECK(d.planets.conditional_prior.Pprior.get() == ca.Pprior.get());
    std::printf("ok\n");
    return 0;
}
