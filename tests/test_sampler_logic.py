"""Device-side sampler logic against the UNMODIFIED reference sampler
code run on the same counter-based stream (oracle/_ref/libkima_ref_philox.so).

CPU tests: kc_sampler_logic.cuh compiled for the host (libkc_hostcheck.so) — the very code the
kernels run — compared draw for draw with RVmodel::perturb (src/RVmodel.cpp:435-820) and
Sampler::mcmc_thread (DNest4/code/SamplerImpl.h:134-265).  The GPU twins of these tests are in
tests/test_sampler_gpu.py.
"""
import ctypes

import numpy as np
import pytest

from kima_b200.flags import ModelFlags
from kima_b200.sampler_types import CPrior, CPriorTable, PRIOR_TYPES

from sampler_util import CASES, HostTwin, case, host_prior_eval, host_prior_make, make_levels, make_state

PRIOR_EXAMPLES = [
    ("Fixed", (0.37,)), ("Uniform", (-3.0, 7.5)), ("LogUniform", (1.0, 1e5)), ("ModifiedLogUniform", (1.0, 1e3)),
    ("Gaussian", (0.3, 2.5)), ("TruncatedGaussian", (0.0, 1.0, -1.0, 2.0)), ("Cauchy", (1.0, 0.5)),
    ("TruncatedCauchy", (np.log(365.0), 1.0, np.log(365.0) - 21, np.log(365.0) + 21)), ("Exponential", (3.0,)),
    ("TruncatedExponential", (2.0, 0.5, 9.0)), ("Laplace", (5.9, 1.3)), ("Kumaraswamy", (0.867, 3.03)),
    ("Rayleigh", (0.2,)), ("TruncatedRayleigh", (0.2, 0.0, 1.0)), ("Triangular", (-1.0, 0.5, 3.0)), ("Pareto", (1.5, 2.0)),
    ("mixGaussianLogUniform", (3.0, 0.7, 0.5, 40.0)),   # src/distributions/mixGaussianLogUniform.cpp
]


@pytest.fixture(scope="module")
def ref():
    from oracle.ref_sampler import RefSampler, have_philox_reference
    if not have_philox_reference():
        pytest.skip("oracle/_ref/libkima_ref_philox.so not built (no /root/reference here)")
    return RefSampler()


def _setup(ref, name):
    flags, data, fix = case(name)
    ref.oracle.set_flags(flags, fix)
    ref.oracle.set_rvdata(data)
    ref.lib.rs_set_priors(None)
    pri = ref.default_priors()
    if flags.known_object:  # no defaults: the user must set them (src/RVmodel.cpp:71-76)
        for i in range(flags.n_known_object):
            pri.KO_Pprior[i] = ref.make_prior("LogUniform", 10.0, 500.0)
            pri.KO_Kprior[i] = ref.make_prior("ModifiedLogUniform", 1.0, 50.0)
            pri.KO_eprior[i] = ref.make_prior("Kumaraswamy", 0.867, 3.03)
            pri.KO_phiprior[i] = ref.make_prior("Uniform", 0.0, 2 * np.pi)
            pri.KO_wprior[i] = ref.make_prior("Fixed", 1.1) if i else ref.make_prior("Uniform", 0.0, 2 * np.pi)
    ref.set_priors(pri)
    return flags, data, pri


def test_stream_draws(ref):
    """rand / randn / randt2 / randh of the lane stream == the stand-in RNG the reference consumes."""
    tw = HostTwin(ModelFlags(npmax=1), case("white")[1], CPriorTable(), 1)
    for what in range(4):
        a, ea = tw.rng_draws(43 + what, 17, what, 2000)
        b, eb = ref.rng_draws(43 + what, 17, what, 2000)
        assert ea == eb
        np.testing.assert_array_equal(a, b)
    u, _ = tw.rng_draws(1, 0, 0, 200000)
    assert 0.0 <= u.min() and u.max() < 1.0 and abs(u.mean() - 0.5) < 5e-3 and abs(u.var() - 1 / 12) < 2e-3
    n, _ = tw.rng_draws(2, 0, 1, 200000)
    assert abs(n.mean()) < 1e-2 and abs(n.var() - 1.0) < 2e-2


@pytest.mark.parametrize("name,pars", PRIOR_EXAMPLES)
def test_prior_functions(ref, name, pars):
    """cdf / cdf_inverse / log_pdf and the constructor constants of every distribution against
    the reference's classes (DNest4/code/Distributions/*.cpp, src/distributions/Fixed.cpp)."""
    q = host_prior_make(name, *pars)
    qr = ref.make_prior(name, *pars)
    assert list(q.p) == list(qr.p) and list(q.k) == list(qr.k)
    rng = np.random.default_rng(5)
    u = np.concatenate([rng.uniform(0, 1, 500), [0.0, 1e-300, 1e-12, 0.5, 1 - 1e-12]])
    x = host_prior_eval(q, 1, u)
    np.testing.assert_array_equal(x, ref.prior_eval(qr, 1, u))
    # arguments inside and outside the support
    xs = np.concatenate([x, x - 10.0, x + 10.0, [0.0, -1.0, 1.0, 1e-8]])
    with np.errstate(all="ignore"):
        for what in (0, 2):
            np.testing.assert_array_equal(host_prior_eval(q, what, xs), ref.prior_eval(qr, what, xs))


def test_prior_make_rejects_what_the_reference_throws_on():
    for name, pars in (("Uniform", (1.0, 1.0)), ("LogUniform", (0.0, 2.0)), ("Gaussian", (0.0, 0.0)),
                       ("Kumaraswamy", (-1.0, 1.0)), ("Triangular", (0.0, 2.0, 1.0)), ("TruncatedCauchy", (0, 1, 2, 1))):
        with pytest.raises(ValueError):
            host_prior_make(name, *pars)


@pytest.mark.parametrize("name", sorted(CASES))
def test_perturb_matches_reference(ref, name):
    """RVmodel::perturb: same stream position afterwards (= same branches taken), same logH,
    same parameters, for every flag combination; 3 chained moves per particle."""
    flags, data, pri = _setup(ref, name)
    B = 400
    S, nps, L = make_state(flags, data, pri, B)
    tw = HostTwin(flags, data, pri, lanes=B, num_particles=1)
    ctr = np.zeros(B, np.uint64)
    for it in range(3):
        tw.upload(S, nps, np.zeros(B), np.full(B, 0.5), np.zeros(B, np.int32))
        tw.set_counters(ctr)
        got_S, got_np, got_H = tw.test_perturb()
        want_S, want_np, want_H, want_ctr = ref.perturb(S, nps, 43, ctr)
        np.testing.assert_array_equal(tw.get_counters(), want_ctr)
        np.testing.assert_array_equal(got_np, want_np)
        np.testing.assert_array_equal(got_H, want_H)
        # rows of components beyond np are dead storage on the device
        for j in range(flags.npmax):
            dead = want_np <= j
            for d in range(5):
                got_S[L["row_P"] + d * flags.npmax + j, dead] = 0.0
                got_S[tw.row_u + d * flags.npmax + j, dead] = 0.0
        np.testing.assert_allclose(got_S, want_S, rtol=1e-13, atol=1e-300)
        S, nps, ctr = want_S, want_np, want_ctr
    assert (ctr > 0).all()


@pytest.mark.parametrize("name,P,L_levels,at_max", [
    ("white", 2, 6, False), ("white", 1, 5, True), ("hyper", 2, 4, False), ("studentt_multi", 3, 1, False),
    ("gp", 2, 5, False), ("known_object", 1, 8, True), ("gp_multi_trend", 2, 3, False), ("trend3_ind", 2, 40, False),
    ("gp_celerite", 1, 4, True), ("fixed_np", 2, 7, False)])
def test_mcmc_steps_match_reference(ref, name, P, L_levels, at_max):
    """Sampler::mcmc_thread: particles, likelihoods, tiebreakers, level assignments, level
    counters, `above` lists and stream positions after several lock-step steps.  (No MA case:
    the reference's perturb() leaves the MA term of mu stale, src/RVmodel.cpp:548-604, so its
    likelihoods depend on a particle's history — DESIGN.md section 1; the MA moves themselves
    are covered by test_perturb_matches_reference[ma].)"""
    flags, data, pri = _setup(ref, name)
    T, steps = 48, 6
    n = T * P
    S, nps, L = make_state(flags, data, pri, n, seed=11)
    logl, _ = ref.oracle.eval_batch(S[:L["n_rows"]], nps)
    logl = np.where(np.isfinite(logl), logl, -1e300)
    rng = np.random.default_rng(2)
    tb = rng.uniform(0.01, 0.99, n)
    lv_logl, lv_tb, lv_logX, cnt = make_levels(logl, L_levels)
    # every particle sits in a level whose threshold it exceeds
    level = np.array([rng.integers(0, 1 + np.searchsorted(lv_logl[1:], v)) for v in logl], np.int32)
    max_num_levels = L_levels if at_max else (0 if L_levels < 30 else 1000)
    enough = at_max
    ctr0 = rng.integers(0, 1000, T).astype(np.uint64)
    want = ref.mcmc(T, P, steps, S, nps, logl, tb, level, lv_logl, lv_tb, lv_logX, cnt, 10.0, 100.0,
                    max_num_levels, 43, ctr0)
    tw = HostTwin(flags, data, pri, lanes=T, num_particles=P, max_steps=steps)
    tw.upload(S, nps, logl, tb, level)
    tw.set_counters(ctr0)
    tw.set_levels(lv_logl, lv_tb, lv_logX, cnt[:, 1], enough, at_max)
    for _ in range(steps):
        Q, npq, fl = tw.propose()
        lq, _ = ref.oracle.eval_batch(Q[:L["n_rows"]], npq)
        tw.accept(lq)
    gS, gnp, glogl, gtb, glevel = tw.download()
    gcnt, gabove, gn_lane = tw.collect()
    np.testing.assert_array_equal(tw.get_counters(), want["ctr"])
    np.testing.assert_array_equal(gnp, want["np"])
    np.testing.assert_array_equal(glevel, want["level"])
    np.testing.assert_array_equal(gcnt, want["counters"])
    np.testing.assert_array_equal(gn_lane, want["n_above_lane"])
    np.testing.assert_allclose(gtb, want["tb"], rtol=1e-13)
    # the reference updates mu incrementally inside perturb, the twin's proposals are scored from scratch
    np.testing.assert_allclose(glogl, want["logl"], rtol=1e-9)
    np.testing.assert_allclose(gabove, want["above"], rtol=1e-9)
    for j in range(flags.npmax):
        dead = want["np"] <= j
        for d in range(5):
            gS[L["row_P"] + d * flags.npmax + j, dead] = 0.0
            gS[tw.row_u + d * flags.npmax + j, dead] = 0.0
    np.testing.assert_allclose(gS, want["state"], rtol=1e-12, atol=1e-300)
    assert gcnt[:, 1].sum() == T * steps  # one try per lane and step


@pytest.mark.parametrize("name", sorted(CASES))
def test_default_prior_table_equals_setPriors(ref, name):
    """kima_b200.sampler.default_prior_table (the host mirror's restatement of setPriors and of the
    RVConditionalPrior constructor) against the table extracted from the reference's own objects."""
    from kima_b200.sampler import default_prior_table
    flags, data, fix = case(name)
    ref.oracle.set_flags(flags, fix)
    ref.oracle.set_rvdata(data)
    ref.lib.rs_set_priors(None)
    want = ref.default_priors()
    got = default_prior_table(flags, data, fix)

    def same(a, b, what):
        assert a.type == b.type, what
        if a.type >= 0:
            assert list(a.p) == list(b.p) and list(a.k) == list(b.k), what

    for f, _ in CPriorTable._fields_:
        a, b = getattr(got, f), getattr(want, f)
        if isinstance(a, CPrior):
            # slots the flags never read may be filled on one side only (setPriors fills offsets_prior
            # and Jprior always, nu_prior only under studentt, ...): compare what perturb() can reach
            used = {"slope_prior": flags.trend, "quadr_prior": flags.trend and flags.degree >= 2,
                    "cubic_prior": flags.trend and flags.degree == 3, "log_eta1_prior": flags.GP, "eta2_prior": flags.GP,
                    "eta3_prior": flags.GP, "log_eta4_prior": flags.GP and flags.kernel == 0, "nu_prior": flags.studentt,
                    "log_muP_prior": flags.hyperpriors, "wP_prior": flags.hyperpriors, "log_muK_prior": flags.hyperpriors,
                    "Pprior": not flags.hyperpriors, "Kprior": not flags.hyperpriors}.get(f, True)
            if used:
                same(a, b, f)
    assert got.fix == want.fix
