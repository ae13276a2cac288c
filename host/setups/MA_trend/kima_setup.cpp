// Test set-up (ours): MA(1) term + cubic trend, no planets.  With npmax = 0 the reference
// never takes the incremental-update branch of calculate_mu (src/RVmodel.cpp:296), so its
// MA term is the from-scratch one (:378-386) and the chains are comparable.
#include "kima.h"

const bool GP = false;
const bool MA = true;
const bool hyperpriors = false;
const bool trend = true;
const int degree = 3;
const bool multi_instrument = false;
const bool known_object = false;
const int n_known_object = 0;
const bool studentt = false;

RVmodel::RVmodel():fix(true),npmax(0)
{
}

int main(int argc, char** argv)
{
    datafile = "d1.txt";
    load(datafile, "ms", 2);
    Sampler<RVmodel> sampler = setup<RVmodel>(argc, argv);
    sampler.run(100);
    return 0;
}
