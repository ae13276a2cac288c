// Test set-up (ours): BASELINE.json's many_planets variant -- up to 10 planets under hyperpriors,
// Student-t likelihood, four instruments with offsets and per-instrument jitters.
#include "kima.h"

const bool GP = false;
const bool MA = false;
const bool hyperpriors = true;
const bool trend = false;
const int degree = 0;
const bool multi_instrument = true;
const bool known_object = false;
const int n_known_object = 0;
const bool studentt = true;

RVmodel::RVmodel():fix(false),npmax(10)
{
}

int main(int argc, char** argv)
{
    datafiles = {"HD106252_ELODIE.txt", "HD106252_HET.txt", "HD106252_HJS.txt", "HD106252_Lick.txt"};
    load_multi(datafiles, "ms", 2);
    Sampler<RVmodel> sampler = setup<RVmodel>(argc, argv);
    sampler.run(100);
    return 0;
}
