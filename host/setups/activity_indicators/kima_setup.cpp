// Test set-up (ours): linear correlations with two activity indicators (columns 4 and 6 of
// examples/activity_correlations/simulated2.txt), the fourth variant listed in that example.
#include "kima.h"

const bool GP = false;
const bool MA = false;
const bool hyperpriors = false;
const bool trend = false;
const int degree = 0;
const bool multi_instrument = false;
const bool known_object = false;
const int n_known_object = 0;
const bool studentt = false;

RVmodel::RVmodel():fix(false),npmax(2)
{
    auto cp = planets.get_conditional_prior();
    cp->Pprior = make_prior<LogUniform>(1, 10 * get_data().get_timespan());
    cp->Kprior = make_prior<ModifiedLogUniform>(1.0, 100);
    cp->eprior = make_prior<Kumaraswamy>(0.867, 3.03);
}

int main(int argc, char** argv)
{
    datafile = "simulated2.txt";
    indicators = {"b", "", "indc"};
    load(datafile, "ms", 4, indicators);
    Sampler<RVmodel> sampler = setup<RVmodel>(argc, argv);
    sampler.run(100);
    return 0;
}
