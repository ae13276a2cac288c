// Test set-up (ours, in the format of the reference's examples/*/kima_setup.cpp): the CoRoT-7
// data with the celerite kernel, the one GP branch no shipped example switches on
// (examples/arbitrary_units/kima_setup.cpp:15 has it commented out).
#include "kima.h"

const bool GP = true;
const bool MA = false;
const bool hyperpriors = true;
const bool trend = false;
const int degree = 0;
const bool multi_instrument = false;
const bool known_object = false;
const int n_known_object = 0;
const bool studentt = false;

RVmodel::RVmodel():fix(false),npmax(3)
{
    kernel = celerite;
}

int main(int argc, char** argv)
{
    datafile = "corot7.txt";
    load(datafile, "ms");
    Sampler<RVmodel> sampler = setup<RVmodel>(argc, argv);
    sampler.run(50);
    return 0;
}
