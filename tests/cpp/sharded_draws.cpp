// The property host/BatchedSampler.h::kill_lagging_device relies on when it shards the first loop of
// kill_lagging_particles (DNest4/code/SamplerImpl.h:543-558) over ranks: one RNG::rand() -- uniform(twister) with
// std::mt19937_64 and std::uniform_real_distribution<double> (DNest4/code/RNG.h:17-20, RNG.cpp) -- consumes a fixed
// number of engine outputs, so a rank that draws for particles [a, b) only and discards the engine over the others
// flags the same particles and leaves the engine in the same state as the serial loop.
//   usage: sharded_draws <n_particles> <ranks> <seed>     prints "ok <per_draw> <n_bad>" or "MISMATCH ..."
#include <cstdio>
#include <cstdlib>
#include <random>
#include <vector>

struct Rng {
    std::mt19937_64 twister;
    std::uniform_real_distribution<double> uniform{0.0, 1.0};
    double rand() { return uniform(twister); }
};

static int outputs_per_rand(const Rng& r)
{
    Rng probe = r;
    auto stepped = probe.twister;
    probe.rand();
    for (int t = 1; t <= 2; ++t) {
        stepped.discard(1);
        if (stepped == probe.twister) return t;
    }
    return 0;
}

int main(int argc, char** argv)
{
    const size_t n = argc > 1 ? (size_t)std::atol(argv[1]) : 100000;
    const int W = argc > 2 ? std::atoi(argv[2]) : 8;
    const unsigned seed = argc > 3 ? (unsigned)std::atol(argv[3]) : 43;
    Rng start;
    start.twister.seed(seed);
    for (int i = 0; i < 1000; ++i) start.rand();  // somewhere inside a run
    // a kill probability per particle, as kill[level_assignments[i]]
    std::vector<double> kill(n);
    {
        std::mt19937_64 g(seed + 1);
        std::uniform_real_distribution<double> u(0.0, 0.02);
        for (double& k : kill) k = u(g);
    }
    // the serial loop
    Rng serial = start;
    std::vector<bool> good_serial(n, true);
    size_t bad_serial = 0;
    for (size_t i = 0; i < n; ++i)
        if (serial.rand() <= kill[i]) { good_serial[i] = false; ++bad_serial; }
    // the ranks
    const int per = outputs_per_rand(start);
    if (per <= 0) { std::printf("MISMATCH outputs per rand not fixed\n"); return 1; }
    std::vector<bool> good(n, true);
    for (int r = 0; r < W; ++r) {
        const size_t a = n * (size_t)r / (size_t)W, b = n * (size_t)(r + 1) / (size_t)W;
        Rng mine = start;
        mine.twister.discard((unsigned long long)per * a);
        for (size_t i = a; i < b; ++i)
            if (mine.rand() <= kill[i]) good[i] = false;
        mine.twister.discard((unsigned long long)per * (n - b));
        if (!(mine.twister == serial.twister)) { std::printf("MISMATCH engine state on rank %d\n", r); return 1; }
        if (mine.rand() != Rng(serial).rand()) { std::printf("MISMATCH next draw on rank %d\n", r); return 1; }
    }
    if (good != good_serial) { std::printf("MISMATCH flags\n"); return 1; }
    std::printf("ok %d %zu\n", per, bad_serial);
    return 0;
}
