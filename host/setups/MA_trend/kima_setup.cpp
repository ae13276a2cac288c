// Test set-up (ours): MA(1) term + cubic trend, no planets.  The reference's perturb() still
// patches mu incrementally for vsys / trend moves without redoing the MA term
// (src/RVmodel.cpp:548-604), so its chain depends on each particle's history and is NOT expected
// to equal a stateless evaluator's: the test checks files, columns and the first saved particle.
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
