// Test set-up (ours): the BL2009 data with up to 5 planets (BASELINE.json: "variable Np up to 5").
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

RVmodel::RVmodel():fix(false),npmax(5)
{
    Cprior = make_prior<Uniform>(-2000, 2000);
    Jprior = make_prior<ModifiedLogUniform>(1.0, 2000.);
    auto cp = planets.get_conditional_prior();
    cp->Pprior = make_prior<LogUniform>(0.2, 15E3);
    cp->Kprior = make_prior<ModifiedLogUniform>(1.0, 2E3);
}

int main(int argc, char** argv)
{
    datafile = "BL2009_dataset2.kms.rv";
    load(datafile, "kms", 0);
    Sampler<RVmodel> sampler = setup<RVmodel>(argc, argv);
    sampler.run(100);
    return 0;
}
