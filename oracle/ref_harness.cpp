// TEST INFRASTRUCTURE — not part of the product.  Only tests/, smoke() and
// bench.py's cpu_baseline / --impl reference legs may load what this builds.
//
// Harness that drives the UNMODIFIED reference sources (compiled where they lie
// under /root/reference by oracle/Makefile into oracle/_ref/libkima_ref.so) on
// chosen parameter vectors.  It reaches the private hot path
//     RVmodel::calculate_mu      /root/reference/src/RVmodel.cpp:286-395
//     RVmodel::calculate_C       /root/reference/src/RVmodel.cpp:170-216
//     RVmodel::log_likelihood    /root/reference/src/RVmodel.cpp:827-943
// with the `#define private public` device described in SURVEY.md §8(c): every
// std / Eigen / celerite header is included first, then the access keywords are
// redefined, then the reference headers are included.  Object layout is
// unchanged, so the normally-compiled reference objects link as they are.
//
// The nine model flags are `extern const` link-time constants in the reference
// (src/RVmodel.h:16-34, src/RVConditionalPrior.h:11).  oracle/ref_flags.c
// defines them as ordinary writable C globals of the same (unmangled) names, so
// one shared object serves every flag combination: flags are set before a
// model is constructed and never while one is alive.

#include <algorithm>
#include <atomic>
#include <chrono>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <iomanip>
#include <iostream>
#include <limits>
#include <memory>
#include <set>
#include <sstream>
#include <string>
#include <thread>
#include <vector>
#include <Eigen/Core>
#include <Eigen/Dense>
#include <Eigen/Cholesky>
#include "celerite/celerite.h"

#define private public
#define protected public
#include "DNest4.h"
#include "Data.h"
#include "RVConditionalPrior.h"
#include "RVmodel.h"
#undef private
#undef protected

#include "kima_cuda.h"  // kc_model_flags / kc_layout only (shared layout rule)

// writable aliases of the reference's flag symbols (defined in ref_flags.c)
extern "C" {
void ref_flags_store(int GP_, int MA_, int hyper_, int trend_, int degree_,
                     int multi_, int ko_, int nko_, int studentt_);
}

static int g_npmax = 1;
static bool g_fix = false;
static kc_model_flags g_flags;
static kc_layout g_layout;

// The user-supplied constructor of every kima_setup.cpp
// (e.g. examples/51Peg/kima_setup.cpp:13).  Default priors are left alone: the
// harness never draws from them.
RVmodel::RVmodel() : fix(g_fix), npmax(g_npmax)
{
    kernel = g_flags.kernel ? celerite : standard;  // examples/*/kima_setup.cpp set it here
}

namespace {

void set_lane(RVmodel& m, const double* params, int ld, int b, int np)
{
    const kc_layout& L = g_layout;
    auto at = [&](int row) { return params[(size_t)row * ld + b]; };

    std::vector<std::vector<double>> comps(np, std::vector<double>(5));
    for (int j = 0; j < np; j++) {
        comps[j][0] = at(L.row_P + j);
        comps[j][1] = at(L.row_K + j);
        comps[j][2] = at(L.row_phi + j);
        comps[j][3] = at(L.row_ecc + j);
        comps[j][4] = at(L.row_w + j);
    }
    m.planets.num_components = np;
    m.planets.components = comps;
    m.planets.added = comps;  // same size as components -> from-scratch branch (:296)
    m.planets.removed.clear();
    m.staleness = 0;

    m.background = at(L.row_vsys);
    if (g_flags.multi_instrument) {
        for (int i = 0; i < L.n_jit; i++) m.jitters[i] = at(L.row_jit + i);
        for (int i = 0; i < L.n_off; i++) m.offsets[i] = at(L.row_off + i);
    } else {
        m.extra_sigma = at(L.row_jit);
    }
    m.slope = 0.0; m.quadr = 0.0; m.cubic = 0.0;
    if (g_flags.trend) {
        m.slope = at(L.row_trend + 0);
        m.quadr = at(L.row_trend + 1);
        m.cubic = at(L.row_trend + 2);
    }
    for (int i = 0; i < L.n_beta; i++) m.betas[i] = at(L.row_beta + i);
    if (g_flags.GP) {
        m.eta1 = at(L.row_eta + 0);
        m.eta2 = at(L.row_eta + 1);
        m.eta3 = at(L.row_eta + 2);
        m.eta4 = at(L.row_eta + 3);
    }
    if (g_flags.studentt) m.nu = at(L.row_nu);
    if (g_flags.MA) {
        m.sigmaMA = at(L.row_MA);
        m.tauMA = at(L.row_MA + 1);
    }
    if (L.n_ko) {
        m.KO_P.resize(L.n_ko); m.KO_K.resize(L.n_ko); m.KO_e.resize(L.n_ko);
        m.KO_phi.resize(L.n_ko); m.KO_w.resize(L.n_ko);
        for (int i = 0; i < L.n_ko; i++) {
            m.KO_P[i] = at(L.row_KO_P + i);
            m.KO_K[i] = at(L.row_KO_K + i);
            m.KO_e[i] = at(L.row_KO_e + i);
            m.KO_phi[i] = at(L.row_KO_phi + i);
            m.KO_w[i] = at(L.row_KO_w + i);
        }
    }
}

double eval_lane(RVmodel& m, const double* params, int ld, int b, int np)
{
    set_lane(m, params, ld, b, np);
    m.calculate_mu();
    if (g_flags.GP) m.calculate_C();
    return m.log_likelihood();
}

}  // namespace

extern "C" {

// ---- flags -----------------------------------------------------------------
int ref_set_flags(const kc_model_flags* f, int fix)
{
    g_flags = *f;
    g_npmax = f->npmax;
    g_fix = fix != 0;
    ref_flags_store(f->GP, f->MA, f->hyperpriors, f->trend, f->degree,
                    f->multi_instrument, f->known_object, f->n_known_object,
                    f->studentt);
    Data& d = Data::get_instance();
    kc_layout_from_flags(&g_flags, d.number_instruments, d.number_indicators,
                         &g_layout);
    return 0;
}

// ---- data ------------------------------------------------------------------
// Through the reference's own parsers (src/Data.cpp:86, 210, 301).
int ref_load(const char* filename, const char* units, int skip)
{
    Data::get_instance().load(filename, units, skip);
    return Data::get_instance().N();
}
int ref_load_indicators(const char* filename, const char* units, int skip,
                        int n_cols, const char** names)
{
    std::vector<std::string> ind;
    for (int i = 0; i < n_cols; i++) ind.push_back(names[i]);
    Data::get_instance().load(filename, units, skip, ind);
    return Data::get_instance().N();
}
int ref_load_multi(const char* filename, const char* units, int skip)
{
    Data::get_instance().load_multi(std::string(filename), units, skip);
    return Data::get_instance().N();
}
int ref_load_multi_files(int nfiles, const char** filenames, const char* units,
                         int skip)
{
    std::vector<std::string> fs;
    for (int i = 0; i < nfiles; i++) fs.push_back(filenames[i]);
    Data::get_instance().load_multi(fs, units, skip);
    return Data::get_instance().N();
}
// Directly from arrays (synthetic data sets), mirroring what load* leaves
// behind: M0_epoch = t[0] (Data.cpp:162), indicators mean-subtracted by caller.
int ref_set_data(int N, const double* t, const double* y, const double* sig,
                 const int* obsi, int n_inst, int n_ind, const double* actind)
{
    Data& d = Data::get_instance();
    d.t.assign(t, t + N);
    d.y.assign(y, y + N);
    d.sig.assign(sig, sig + N);
    d.obsi.clear();
    if (obsi) d.obsi.assign(obsi, obsi + N);
    d.number_instruments = n_inst;
    d.datamulti = n_inst > 1;
    d.number_indicators = n_ind;
    d.indicator_correlations = n_ind > 0;
    d.actind.assign(n_ind, std::vector<double>());
    for (int j = 0; j < n_ind; j++)
        d.actind[j].assign(actind + (size_t)j * N, actind + (size_t)(j + 1) * N);
    d.M0_epoch = t[0];
    kc_layout_from_flags(&g_flags, d.number_instruments, d.number_indicators, &g_layout);
    return N;
}
int ref_data_shape(int* N, int* n_inst, int* n_ind)
{
    Data& d = Data::get_instance();
    *N = d.N(); *n_inst = d.number_instruments; *n_ind = d.number_indicators;
    return 0;
}
int ref_get_data(double* t, double* y, double* sig, int* obsi, double* actind,
                 double* M0_epoch, double* t_middle)
{
    Data& d = Data::get_instance();
    int N = d.N();
    std::copy(d.t.begin(), d.t.end(), t);
    std::copy(d.y.begin(), d.y.end(), y);
    std::copy(d.sig.begin(), d.sig.end(), sig);
    for (int i = 0; i < N; i++)
        obsi[i] = ((int)d.obsi.size() == N) ? d.obsi[i] : 1;
    for (int j = 0; j < d.number_indicators; j++)
        std::copy(d.actind[j].begin(), d.actind[j].end(), actind + (size_t)j * N);
    *M0_epoch = d.M0_epoch;
    *t_middle = d.get_t_middle();
    return N;
}

// ---- Kepler solver (ordinary external functions, src/RVmodel.h:189-196) -----
void ref_kepler(int n, const double* M, const double* ecc, double* E)
{
    for (int i = 0; i < n; i++) E[i] = kepler::kepler(M[i], ecc[i]);
}
void ref_true_anomaly(int n, const double* t, const double* P, const double* ecc,
                      const double* Tp, double* f)
{
    for (int i = 0; i < n; i++) f[i] = kepler::true_anomaly(t[i], P[i], ecc[i], Tp[i]);
}

// ---- the hot path ----------------------------------------------------------
// logL[b] for lanes [0,B) of the SoA block params[n_rows][ld]; nthreads
// std::threads, each owning one RVmodel, lanes dealt round-robin in chunks.
// Returns wall seconds spent inside the evaluation loop.
double ref_eval_batch(int B, int ld, const double* params, const int* np,
                      double* logL, int nthreads)
{
    if (nthreads < 1) nthreads = 1;
    std::atomic<int> next(0);
    auto t0 = std::chrono::steady_clock::now();
    auto work = [&]() {
        RVmodel m;
        for (;;) {
            int b = next.fetch_add(1);
            if (b >= B) break;
            logL[b] = eval_lane(m, params, ld, b, np[b]);
        }
    };
    if (nthreads == 1) {
        work();
    } else {
        std::vector<std::thread> th;
        for (int k = 0; k < nthreads; k++) th.emplace_back(work);
        for (auto& x : th) x.join();
    }
    auto t1 = std::chrono::steady_clock::now();
    return std::chrono::duration<double>(t1 - t0).count();
}

// mu[N] (long double in the reference, returned as double and as the
// hi/lo split of the 80-bit value) for one lane.
int ref_mu(int ld, const double* params, int b, int np, double* mu, double* mu_lo)
{
    RVmodel m;
    set_lane(m, params, ld, b, np);
    m.calculate_mu();
    for (size_t i = 0; i < m.mu.size(); i++) {
        double hi = (double)m.mu[i];
        mu[i] = hi;
        if (mu_lo) mu_lo[i] = (double)(m.mu[i] - (long double)hi);
    }
    return (int)m.mu.size();
}

// C[N][N] (Eigen column-major) for one lane.
int ref_cov(int ld, const double* params, int b, int np, double* C)
{
    RVmodel m;
    set_lane(m, params, ld, b, np);
    m.calculate_C();
    size_t N = m.C.rows();
    std::memcpy(C, m.C.data(), N * N * sizeof(double));
    return (int)N;
}

// kernel=celerite: the conditional mean K_* (K + sigma)^-1 (y - mu) of the reference's own solver,
// celerite::solver::CholeskySolver::predict (celerite/cpp/include/celerite/solver/cholesky.h:590-690),
// after calculate_mu / calculate_C have set it up as log_likelihood's celerite branch does
// (src/RVmodel.cpp:218-266, 875-877).  x must be sorted.  Returns 0, or 1 if the solver threw.
int ref_celerite_predict(int ld, const double* params, int b, int np, int nt, const double* x, double* out)
{
    RVmodel m;
    set_lane(m, params, ld, b, np);
    m.calculate_mu();
    m.calculate_C();
    const std::vector<double>& y = Data::get_instance().get_y();
    const int N = (int)y.size();
    Eigen::VectorXd r(N), xs(nt);
    for (int i = 0; i < N; i++) r(i) = y[i] - (double)m.mu[i];
    for (int k = 0; k < nt; k++) xs(k) = x[k];
    try {
        Eigen::VectorXd pr = m.solver.predict(r, xs);
        for (int k = 0; k < nt; k++) out[k] = pr(k);
    } catch (...) {
        return 1;
    }
    return 0;
}

int ref_get_layout(kc_layout* out) { *out = g_layout; return 0; }

// What DNest4 does with the value (LikelihoodType.cpp:17-24).
double ref_likelihood_type_value(double v)
{
    DNest4::LikelihoodType l(v, 0.5);
    return l.value;
}

}  // extern "C"
