// Stand-ins for the four RVmodel methods whose work now happens on the GPU:
//     RVmodel::calculate_mu         /root/reference/src/RVmodel.cpp:286-395
//     RVmodel::calculate_C          /root/reference/src/RVmodel.cpp:170-280
//     RVmodel::add_known_object     /root/reference/src/RVmodel.cpp:416-432
//     RVmodel::remove_known_object  /root/reference/src/RVmodel.cpp:397-413
// The reference's perturb()/from_prior() keep calling them (src/RVmodel.cpp:160-162, 454,
// 514, 528-544, ...); host/Makefile weakens the reference's definitions in a COPY of its
// object file (objcopy -W), so these strong ones are linked instead and the per-proposal
// O(N Np) / O(N^2) CPU work disappears.  The one piece of state they still own is the
// `staleness` counter, because RVmodel::print writes it to sample.txt (src/RVmodel.cpp:1003):
// the rule below is the reference's (src/RVmodel.cpp:296-297, 306, 344-345).
#include <atomic>
#include <Eigen/Core>
#include <Eigen/Dense>
#include "celerite/celerite.h"
#define private public
#define protected public
#include "DNest4.h"
#include "Data.h"
#include "RVConditionalPrior.h"
#include "RVmodel.h"
#undef private
#undef protected

void RVmodel::calculate_mu()
{
    const bool update = (planets.get_added().size() < planets.get_components().size()) &&
                        (staleness <= 10);
    if (update) staleness++;
    else staleness = 0;
}

void RVmodel::calculate_C() {}
void RVmodel::add_known_object() {}
void RVmodel::remove_known_object() {}

// save_setup() writes kima_model_setup.txt from the priors, which are the same objects for
// every particle; from_prior() calls it for each particle (src/RVmodel.cpp:87-88).  Run the
// reference's own function (kept reachable as kima_ref_save_setup by host/Makefile) for the
// first particle only: same file, written once.
extern "C" void kima_ref_save_setup(RVmodel*);
void RVmodel::save_setup()
{
    static std::atomic<bool> done(false);
    if (!done.exchange(true)) kima_ref_save_setup(this);
}
