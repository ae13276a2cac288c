"""GPU: the device-side sampler (include/kima_sampler.h) through the C ABI
against the UNMODIFIED reference sampler code on the same counter-based stream
(oracle/_ref/libkima_ref_philox.so, built in the container and shipped to the GPU box).

Discrete outcomes (branches taken = stream positions, numbers of planets, level assignments,
level counters, lengths of the `above` lists) must be equal; continuous ones agree to the
rounding of the device's libm against glibc's (and, for likelihoods, to the 1e-10 of the
likelihood kernels).
"""
import numpy as np
import pytest

from kima_b200 import LikelihoodEngine, ModelFlags
from kima_b200.sampler import DeviceSampler, SamplerOptions
from kima_b200.sampler_types import CPriorTable

from sampler_util import CASES, case, make_levels, make_state
from test_sampler_logic import PRIOR_EXAMPLES, _setup

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ref():
    from oracle.ref_sampler import RefSampler, have_philox_reference
    if not have_philox_reference():
        pytest.fail("oracle/_ref/libkima_ref_philox.so is missing: it is built in the container and must travel")
    return RefSampler()


def _zero_dead(S, nps, L, npmax, row_u):
    for j in range(npmax):
        dead = nps <= j
        for d in range(5):
            S[L["row_P"] + d * npmax + j, dead] = 0.0
            S[row_u + d * npmax + j, dead] = 0.0


def test_stream_and_priors_on_device(gpu, ref):
    flags, data, _ = case("white")
    ref.oracle.set_flags(flags)
    ref.oracle.set_rvdata(data)
    ref.lib.rs_set_priors(None)
    pri = ref.default_priors()
    with LikelihoodEngine(data, flags, 8) as eng:
        ds = DeviceSampler(eng, pri, SamplerOptions(lanes=8))
        for what, tol in ((0, 0.0), (1, 1e-14), (2, 1e-14), (3, 1e-13)):
            a, ea = ds.rng_draws(43 + what, 5, what, 3000)
            b, eb = ref.rng_draws(43 + what, 5, what, 3000)
            assert ea == eb
            np.testing.assert_allclose(a, b, rtol=tol, atol=0)
        rng = np.random.default_rng(5)
        u = np.concatenate([rng.uniform(0, 1, 500), [0.0, 1e-12, 0.5, 1 - 1e-12]])
        for name, pars in PRIOR_EXAMPLES:
            q = ref.make_prior(name, *pars)
            x = ref.prior_eval(q, 1, u)
            np.testing.assert_allclose(ds.prior_eval(q, 1, u), x, rtol=2e-13, atol=1e-300, err_msg=name)
            xs = np.concatenate([x, x - 10.0, x + 10.0, [0.0, -1.0, 1.0, 1e-8]])
            with np.errstate(all="ignore"):
                for what in (0, 2):
                    np.testing.assert_allclose(ds.prior_eval(q, what, xs), ref.prior_eval(q, what, xs), rtol=2e-13,
                                               atol=1e-15, err_msg="%s %d" % (name, what))
        ds.close()


@pytest.mark.parametrize("name", sorted(CASES))
def test_perturb_on_device_matches_reference(gpu, ref, name):
    """RVmodel::perturb, 3 chained moves of 2048 particles per flag combination."""
    flags, data, pri = _setup(ref, name)
    B = 2048
    S, nps, L = make_state(flags, data, pri, B)
    with LikelihoodEngine(data, flags, B) as eng:
        ds = DeviceSampler(eng, pri, SamplerOptions(lanes=B, num_particles=1))
        ctr = np.zeros(B, np.uint64)
        for it in range(3):
            ds.upload(S, nps, np.zeros(B), np.full(B, 0.5), np.zeros(B, np.int32))
            ds.set_counters(ctr)
            got_S, got_np, got_H = ds.test_perturb()
            want_S, want_np, want_H, want_ctr = ref.perturb(S, nps, 43, ctr)
            np.testing.assert_array_equal(ds.get_counters(), want_ctr)
            np.testing.assert_array_equal(got_np, want_np)
            np.testing.assert_allclose(got_H, want_H, rtol=1e-9, atol=1e-9)
            _zero_dead(got_S, want_np, L, flags.npmax, ds.row_u)
            np.testing.assert_allclose(got_S, want_S, rtol=1e-11, atol=1e-12)  # u's wrap to values near 0
            S, nps, ctr = want_S, want_np, want_ctr
        ds.close()


@pytest.mark.parametrize("name,P,L_levels,at_max", [
    ("white", 2, 6, False), ("white", 1, 5, True), ("hyper", 2, 4, False), ("studentt_multi", 3, 1, False),
    ("gp", 2, 5, False), ("known_object", 1, 8, True), ("gp_multi_trend", 2, 3, False), ("trend3_ind", 2, 40, False),
    ("gp_celerite", 1, 4, True), ("fixed_np", 2, 7, False),
    ("white", 2, 100, False)])  # more levels than the counter rows allocated at creation: they grow
def test_mcmc_on_device_matches_reference(gpu, ref, name, P, L_levels, at_max):
    """kc_sampler_run (propose -> likelihood kernels -> accept, all on the GPU) against
    Sampler::mcmc_thread of the reference with its own CPU likelihood."""
    _mcmc_case(ref, name, P, L_levels, at_max, 256)


@pytest.mark.parametrize("block", ["64", "256"])
@pytest.mark.parametrize("name,P,L_levels", [("white", 2, 6), ("hyper", 2, 4), ("gp", 1, 5), ("known_object", 1, 8)])
def test_move_sorted_proposal_kernels_match_reference(gpu, ref, monkeypatch, block, name, P, L_levels):
    """From 2048 lanes on the proposal kernels sort the lanes of a CTA by their move before making it
    (smp_sorted_propose_kernel); KC_SMP_BLOCK forces them at a size the reference finishes in seconds.
    300 lanes: the last CTA is ragged.  Same decisions, counters and stream positions as the reference."""
    monkeypatch.setenv("KC_SMP_BLOCK", block)
    _mcmc_case(ref, name, P, L_levels, False, 300)


def _mcmc_case(ref, name, P, L_levels, at_max, T):
    flags, data, pri = _setup(ref, name)
    steps = 8
    n = T * P
    S, nps, L = make_state(flags, data, pri, n, seed=11)
    logl, _ = ref.oracle.eval_batch(S[:L["n_rows"]], nps, nthreads=8)
    logl = np.where(np.isfinite(logl), logl, -1e300)
    rng = np.random.default_rng(2)
    tb = rng.uniform(0.01, 0.99, n)
    lv_logl, lv_tb, lv_logX, cnt = make_levels(logl, L_levels)
    level = np.array([rng.integers(0, 1 + np.searchsorted(lv_logl[1:], v)) for v in logl], np.int32)
    max_num_levels = L_levels if at_max else (0 if L_levels < 30 else 1000)
    ctr0 = rng.integers(0, 1000, T).astype(np.uint64)
    want = ref.mcmc(T, P, steps, S, nps, logl, tb, level, lv_logl, lv_tb, lv_logX, cnt, 10.0, 100.0,
                    max_num_levels, 43, ctr0)
    with LikelihoodEngine(data, flags, T) as eng:
        ds = DeviceSampler(eng, pri, SamplerOptions(lanes=T, num_particles=P, thread_steps=steps,
                                                    max_levels=max(64, 2 * L_levels)))
        ds.upload(S, nps, logl, tb, level)
        ds.set_counters(ctr0)
        ds.set_levels(lv_logl, lv_tb, lv_logX, cnt[:, 1], at_max, at_max)
        launches0 = eng.launch_count()
        ds.run(steps)
        gcnt, gabove = ds.collect()
        assert eng.launch_count() - launches0 >= 2 * steps + 1
        gS, gnp, glogl, gtb, glevel = ds.download()
        np.testing.assert_array_equal(ds.get_counters(), want["ctr"])
        np.testing.assert_array_equal(gnp, want["np"])
        np.testing.assert_array_equal(glevel, want["level"])
        np.testing.assert_array_equal(gcnt, want["counters"])
        assert gabove.shape == want["above"].shape
        np.testing.assert_allclose(gtb, want["tb"], rtol=1e-11, atol=1e-12)
        # hyperpriors draw log P from a Laplace: periods down to minutes, where M = 2 pi (t - Tp) / P
        # has lost 8 digits before any solver sees it (in the reference as much as here)
        ltol = 1e-6 if flags.hyperpriors else 1e-9
        np.testing.assert_allclose(glogl, want["logl"], rtol=ltol)
        np.testing.assert_allclose(gabove, want["above"], rtol=ltol)
        _zero_dead(gS, want["np"], L, flags.npmax, ds.row_u)
        np.testing.assert_allclose(gS, want["state"], rtol=1e-10, atol=1e-11)
        assert gcnt[:, 1].sum() == T * steps
        # particle bookkeeping entry points
        col, npc = ds.get_particle(3)
        np.testing.assert_array_equal(col, gS_raw(ds)[:, 3])
        ds.copy_particles([0, 5], [7, 9])
        S2, np2, l2, t2, lev2 = ds.download()
        for d, s_ in ((0, 7), (5, 9)):
            np.testing.assert_array_equal(S2[:, d], S2[:, s_])
            assert (np2[d], l2[d], t2[d], lev2[d]) == (np2[s_], l2[s_], t2[s_], lev2[s_])
        ds.set_particle(2, S2[:, 11], int(np2[11]), -12.5, 0.25, 1)
        S3, np3, l3, t3, lev3 = ds.download()
        np.testing.assert_array_equal(S3[:, 2], S2[:, 11])
        assert (np3[2], l3[2], t3[2], lev3[2]) == (np2[11], -12.5, 0.25, 1)
        ds.close()


def gS_raw(ds):
    return ds.download()[0]


def test_prior_is_the_stationary_distribution(gpu, ref):
    """With a single level at the bottom every scored proposal is accepted, so the chain must
    sample the prior: uniform number of planets (RJObject birth/death), uniform eccentricities and
    phases, Cprior-uniform systemic velocity (size-independent property of the moves)."""
    flags, data, fix = ModelFlags(npmax=3), case("white")[1], False
    ref.oracle.set_flags(flags, fix)
    ref.oracle.set_rvdata(data)
    ref.lib.rs_set_priors(None)
    pri = ref.default_priors()
    T = 8192
    S, nps, L = make_state(flags, data, pri, T, seed=3)
    with LikelihoodEngine(data, flags, T) as eng:
        ds = DeviceSampler(eng, pri, SamplerOptions(lanes=T, num_particles=1, thread_steps=200))
        ds.upload(S, nps, np.full(T, -1e200), np.full(T, 0.5), np.zeros(T, np.int32))
        ds.set_levels([-1e300], [0.0], [0.0], np.zeros(1, np.uint64), False, False)
        for _ in range(3):
            ds.run(200)
        cnt, above = ds.collect()
        S2, np2, logl2, tb2, lev2 = ds.download()
        ds.close()
    assert cnt[0, 1] == 3 * 200 * T and cnt[0, 0] > 0.95 * cnt[0, 1]
    hist = np.bincount(np2, minlength=4) / T
    assert np.all(np.abs(hist - 0.25) < 0.02), hist
    v = (S2[L["row_vsys"]] - data.get_RV_min()) / (data.get_RV_max() - data.get_RV_min())
    assert abs(v.mean() - 0.5) < 0.015 and abs(v.var() - 1 / 12) < 0.01
    alive = np2 >= 1
    e = S2[L["row_ecc"]][alive]
    assert abs(e.mean() - 0.5) < 0.02 and abs(e.var() - 1 / 12) < 0.01
    # the likelihoods stored with the particles are those of the particles
    want, _ = ref.oracle.eval_batch(S2[:L["n_rows"]], np2, nthreads=8)
    want = np.where(np.isfinite(want), want, -1e300)
    np.testing.assert_allclose(logl2, want, rtol=1e-9)


def test_many_barriers_on_device_match_reference(gpu, ref):
    """Blocks of steps with the barrier's hand-over in between (counters folded into the level
    table, `above` lists emptied, stream positions carried over): 12 barriers x 40 steps of 256
    lanes stay equal to the reference's mcmc_thread run barrier by barrier."""
    flags, data, pri = _setup(ref, "fixed_np")
    T, P, steps, L_levels = 256, 2, 40, 12
    n = T * P
    S, nps, L = make_state(flags, data, pri, n, seed=11)
    logl, _ = ref.oracle.eval_batch(S[:L["n_rows"]], nps, nthreads=8)
    logl = np.where(np.isfinite(logl), logl, -1e300)
    rng = np.random.default_rng(2)
    tb = rng.uniform(0.01, 0.99, n)
    lv_logl, lv_tb, lv_logX, cnt = make_levels(logl, L_levels)
    level = np.array([rng.integers(0, 1 + np.searchsorted(lv_logl[1:], v)) for v in logl], np.int32)
    ctr = np.zeros(T, np.uint64)
    with LikelihoodEngine(data, flags, T) as eng:
        ds = DeviceSampler(eng, pri, SamplerOptions(lanes=T, num_particles=P, thread_steps=steps, max_levels=64))
        ds.upload(S, nps, logl, tb, level)
        ds.set_counters(ctr)
        state = dict(state=S, np=nps, logl=logl, tb=tb, level=level, ctr=ctr)
        cnt = cnt.copy()
        for barrier in range(12):
            want = ref.mcmc(T, P, steps, state["state"], state["np"], state["logl"], state["tb"], state["level"],
                            lv_logl, lv_tb, lv_logX, cnt, 10.0, 100.0, L_levels, 43, state["ctr"])
            ds.set_levels(lv_logl, lv_tb, lv_logX, cnt[:, 1], True, True)
            ds.run(steps)
            gcnt, gabove = ds.collect()
            gS, gnp, glogl, gtb, glevel = ds.download()
            np.testing.assert_array_equal(ds.get_counters(), want["ctr"], err_msg="barrier %d" % barrier)
            np.testing.assert_array_equal(glevel, want["level"], err_msg="barrier %d" % barrier)
            np.testing.assert_array_equal(gcnt, want["counters"], err_msg="barrier %d" % barrier)
            # (the reference's incremental mu against from-scratch kernels: 1e-9 seen at logL ~ -1e3)
            np.testing.assert_allclose(glogl, want["logl"], rtol=1e-7)
            cnt = cnt + want["counters"]
            state = want
        ds.close()


def test_above_pool_in_hbm_orders_and_cuts_like_the_host(gpu, ref):
    """kc_sampler_collect_pool / kc_sampler_pool_select against numpy on the lists kc_sampler_collect
    returns: append in lane order over two barriers, lexicographic (value, tiebreaker) order with ties
    (a particle that stays above the top level is listed again every step), cut at `index`."""
    flags, data, pri = _setup(ref, "white")
    T, P, steps = 512, 1, 30
    S, nps, L = make_state(flags, data, pri, T, seed=5)
    logl, _ = ref.oracle.eval_batch(S[:L["n_rows"]], nps, nthreads=8)
    logl = np.where(np.isfinite(logl), logl, -1e300)
    tb = np.random.default_rng(3).uniform(0.01, 0.99, T)
    lv_logl, lv_tb, lv_logX, cnt = make_levels(logl, 3)
    with LikelihoodEngine(data, flags, T) as eng:
        ds = DeviceSampler(eng, pri, SamplerOptions(lanes=T, num_particles=P, thread_steps=steps, max_levels=16,
                                                    pool_reserve=1000))
        ds.upload(S, nps, logl, tb, np.zeros(T, np.int32))
        want = []
        for barrier in range(2):
            ds.set_levels(lv_logl, lv_tb, lv_logX, cnt[:, 1], False, False)
            ds.run(steps)
            c1, above = ds.collect()
            c2, n_new, n_pool = ds.collect_pool()
            np.testing.assert_array_equal(c1, c2)
            want.append(above)
            assert n_new == len(above) and n_pool == sum(len(a) for a in want)
        allp = np.concatenate(want)
        np.testing.assert_array_equal(ds.pool_fetch(len(allp) + 5), allp)
        assert len(np.unique(allp, axis=0)) < len(allp)  # ties are in
        order = np.lexsort((allp[:, 1], allp[:, 0]))
        srt = allp[order]
        index = int((1.0 - 1.0 / np.e) * len(allp))
        v, t, left = ds.pool_select(index)
        assert (v, t) == (srt[index, 0], srt[index, 1]) and left == len(allp) - index - 1
        np.testing.assert_array_equal(ds.pool_fetch(len(allp)), srt[index + 1:])
        v2, t2, left2 = ds.pool_select(0)
        assert (v2, t2) == (srt[index + 1, 0], srt[index + 1, 1]) and left2 == left - 1
        ds.pool_clear()
        assert len(ds.pool_fetch(10)) == 0
        ds.close()
