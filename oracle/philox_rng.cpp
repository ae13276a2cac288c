// TEST INFRASTRUCTURE — not part of the product.  Only tests/ may load what this builds.
//
// Stand-in for DNest4/code/RNG.cpp (the ONE reference source file left out of
// oracle/_ref/libkima_ref_philox.so): the same class (the reference's own RNG.h is used
// unchanged), the same public methods, but the numbers come from the counter-based stream
// of include/kc_philox.h instead of std::mt19937_64.  Every other reference source —
// RVmodel::perturb, RJObject, the conditional prior, the distributions, the Sampler — is
// compiled unmodified and consumes this stream, which is what the device-side sampler
// (kima_b200/csrc/kc_sampler_logic.cuh) draws from as well: the two can then be compared
// draw for draw.
//
// State: {key, counter}, kept in the first 16 bytes of the (otherwise unused) twister
// member, so RNG objects stay copyable exactly as the reference's are.
#include <cmath>
#include <cstdint>
#include <cstring>
#include "RNG.h"
#include "../include/kc_philox.h"

namespace {
struct PhiloxState {
    uint64_t key, ctr;
};
}  // namespace

extern "C" {
// harness access to a stream's position (oracle/ref_sampler_harness.cpp)
void philox_rng_set(DNest4::RNG* r, uint64_t key, uint64_t ctr)
{
    PhiloxState s{key, ctr};
    std::memcpy(reinterpret_cast<void*>(r), &s, sizeof s);
}
void philox_rng_get(const DNest4::RNG* r, uint64_t* key, uint64_t* ctr)
{
    PhiloxState s;
    std::memcpy(&s, reinterpret_cast<const void*>(r), sizeof s);
    *key = s.key;
    *ctr = s.ctr;
}
}

namespace DNest4 {

static_assert(sizeof(RNG) >= sizeof(PhiloxState), "RNG too small to hold the stream position");

bool RNG::randh_is_randh2 = false;

RNG::RNG() : uniform(0., 1.), normal(0., 1.) { philox_rng_set(this, 0, 0); }

RNG::RNG(unsigned int seed) : uniform(0., 1.), normal(0., 1.) { set_seed(seed); }

void RNG::set_seed(unsigned int seed) { philox_rng_set(this, seed, 0); }

double RNG::rand()
{
    uint64_t key, ctr;
    philox_rng_get(this, &key, &ctr);
    const double u = kc_philox_uniform(key, ctr);
    philox_rng_set(this, key, ctr + 1);
    return u;
}

double RNG::randn()
{
    const double u1 = rand();
    const double u2 = rand();
    return sqrt(-2.0 * log(1.0 - u1)) * cos(2.0 * M_PI * u2);
}

double RNG::randt2()
{
    const double n = randn();
    const double u = rand();
    return n / sqrt(-log(u));
}

double RNG::randh()
{
    const double t = randt2();
    const double n = randn();
    return pow(10.0, 1.5 - 3 * std::abs(t)) * n;
}

double RNG::randh2()
{
    // Cauchy().generate(*this) = tan(pi (u - 0.5)); unused by kima (randh_is_randh2 stays false)
    const double c = tan(M_PI * (rand() - 0.5));
    const double n = randn();
    return pow(10.0, 0.5 - std::abs(c)) * n;
}

int RNG::rand_int(int N) { return static_cast<int>(floor(N * rand())); }

}  // namespace DNest4
