// Overlay of the reference's src/kima.h: put this directory FIRST on the include path and an
// unmodified examples/*/kima_setup.cpp builds against the batched GPU sampler instead of the
// per-thread CPU one.  It declares what src/kima.h declares (same globals, same using-
// directive), includes the reference's own headers untouched, and then redirects the two
// names a kima_setup.cpp uses to start sampling:
//
//     Sampler<RVmodel> sampler = setup<RVmodel>(argc, argv);     (examples/51Peg/kima_setup.cpp:38)
//
// `Sampler` -> DNest4::BatchedSampler<RVmodel, kima_b200::RVmodelEvaluator>, `setup` -> the
// same start-up sequence as DNest4/code/StartImpl.h:12-40 for that type.  The user's file,
// the OPTIONS file, the CLI flags (-t lanes, -s seed, ...) and every output file keep their
// meaning.  (A maintainer would not ship macros: INTEGRATION.md shows the three-hunk patch to
// Sampler/RVmodel that this overlay emulates without touching the reference tree.)
#pragma once

// every header the reference's headers pull in, BEFORE the access keywords are opened, so
// that only kima's / DNest4's own classes are affected
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
#include <istream>
#include <limits>
#include <malloc.h>
#include <map>
#include <memory>
#include <mutex>
#include <numeric>
#include <ostream>
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
#include "distributions/Fixed.h"
#include "distributions/Empirical.h"
#include "distributions/mixGaussianLogUniform.h"
#include "distributions/Gaussian_from_Tc.h"
#include "Data.h"
#include "RVmodel.h"
#include "RVConditionalPrior.h"
#undef private
#undef protected

// the globals of src/kima.h:13-22, which kima_setup.cpp files assign to
const double PI = M_PI;
std::string datafile;
std::vector<std::string> datafiles;
using ind = std::vector<char*>;
std::vector<std::string> indicators;
using namespace DNest4;

// the evaluator behind the batched sampler: the GPU library, unless a test names another one
// (tests/cpp/RefEvaluator.h: the reference's own log_likelihood(), to test the sampler without a GPU)
#ifndef KIMA_B200_EVALUATOR_HEADER
#define KIMA_B200_EVALUATOR_HEADER "RVmodelDevice.h"
#define KIMA_B200_EVALUATOR RVmodelEvaluator
#endif
#include KIMA_B200_EVALUATOR_HEADER
#include "BatchedSampler.h"

namespace kima_b200 {

template <class ModelType>
using GpuSampler = DNest4::BatchedSampler<ModelType, KIMA_B200_EVALUATOR>;

// DNest4::setup (StartImpl.h:12-40) for the batched sampler: same messages, same order
template <class ModelType>
GpuSampler<ModelType> gpu_setup(int argc, char** argv)
{
    DNest4::CommandLineOptions options(argc, argv);
    std::cout << "# Using " << options.get_num_threads() << " thread"
              << ((options.get_num_threads() == 1) ? (".") : ("s.")) << std::endl;
    std::cout << "# Target compression factor between levels = ";
    std::cout << options.get_compression() << std::endl;
    DNest4::Options sampler_options(options.get_options_file().c_str());
    // one process per GPU (RANK / WORLD_SIZE in the environment): every rank keeps the whole
    // sampler state, rank 0 alone writes the output files
    const char* rk = std::getenv("RANK");
    const bool root = !rk || std::atoi(rk) == 0;
    // Every default-constructed RVmodel allocates (never touches) an N x N matrix and an N vector
    // (src/RVmodel.h:84,95) until RVmodelEvaluator::strip() drops them; with tens of thousands
    // of lanes that is one mmap per particle, against vm.max_map_count = 65530.  Serve them from
    // the heap instead.
    mallopt(M_MMAP_THRESHOLD, 32 * 1024 * 1024);
    const auto t0 = std::chrono::steady_clock::now();
    GpuSampler<ModelType> sampler(options.get_num_threads(), options.get_compression_double(),
                                  sampler_options, root);
    const auto t1 = std::chrono::steady_clock::now();
    sampler.initialise(options.get_seed_uint());
    if (std::getenv("KIMA_PROFILE"))
        std::cerr << "# setup: sampler constructed in " << std::chrono::duration<double>(t1 - t0).count()
                  << " s, initialised in "
                  << std::chrono::duration<double>(std::chrono::steady_clock::now() - t1).count() << " s" << std::endl;
    return sampler;
}

}  // namespace kima_b200

#define Sampler kima_b200::GpuSampler
#define setup kima_b200::gpu_setup
