"""GPU: the factor cache of the shared-memory-resident GP path (include/kima_cuda.h: kc_gp_cache_enable).

The reference factors C on every log_likelihood() call (src/RVmodel.cpp:841-873) although most of
RVmodel::perturb's proposals (src/RVmodel.cpp:448-544) leave C alone; the cache keeps the last two factors
of every slot.  What has to hold: a hit returns the bits a miss returns (so chains do not depend on the
cache), entries are found by the parameters C depends on and by nothing else, and the older entry is the
one replaced.
"""
import numpy as np
import pytest

from conftest import load_golden
from kima_b200 import LikelihoodEngine, ModelFlags
from kima_b200.synthetic import make_data, make_params

pytestmark = pytest.mark.gpu


def _no_cache(data, flags, P, nps, mode=0):
    with LikelihoodEngine(data, flags, P.shape[1], kepler_mode=mode) as eng:
        return eng.evaluate(P, nps)


def _move_planets(P, L, rng):
    """A proposal that leaves C alone: planets, systemic velocity (and offsets / trend when present)."""
    Q = P.copy()
    n = max(L["npmax"], 1)
    Q[L["row_K"]:L["row_K"] + n] *= rng.uniform(0.5, 1.5, (n, P.shape[1]))
    Q[L["row_phi"]:L["row_phi"] + n] = rng.uniform(0, 2 * np.pi, (n, P.shape[1]))
    Q[L["row_vsys"]] += rng.normal(0, 1.0, P.shape[1])
    if L["n_off"]:
        Q[L["row_off"]:L["row_off"] + L["n_off"]] += rng.normal(0, 0.5, (L["n_off"], P.shape[1]))
    if L["row_trend"] >= 0:
        Q[L["row_trend"]] += rng.normal(0, 1e-3, P.shape[1])
    return Q


@pytest.mark.parametrize("mode", [0, 1])
@pytest.mark.parametrize("N,flags_kw", [
    (24, dict(npmax=1)),                                   # one partial block
    (32, dict(npmax=2)),                                   # exactly one block
    (64, dict(npmax=5)),                                   # 4-warp CTAs
    (90, dict(npmax=3, trend=True, degree=2)),             # 5-warp CTAs
    (100, dict(npmax=5)),                                  # 8-warp CTAs, rolled panels
    (128, dict(npmax=2)),                                  # full blocks only
    (177, dict(npmax=5, hyperpriors=True)),                # CoRoT7's shape: 16-warp CTAs, short last block
    (200, dict(npmax=1)),                                  # residual row alone in the last block's third strip
    (208, dict(npmax=2)),                                  # row N in a tile of its own
    (211, dict(npmax=1)),                                  # the largest N of the path
])
def test_hit_returns_the_bits_of_a_miss(gpu, N, flags_kw, mode):
    flags = ModelFlags(GP=True, **flags_kw)
    data = make_data(N)
    B = 96
    P, nps = make_params(flags, data, B, "mix", seed=N)
    rng = np.random.default_rng(N)
    ref0 = _no_cache(data, flags, P, nps, mode)
    with LikelihoodEngine(data, flags, B, kepler_mode=mode) as eng:
        L = eng.layout
        assert eng.enable_gp_cache(B) == B
        slots = np.arange(B, dtype=np.int32)
        eng.set_slots(slots)
        miss = eng.evaluate(P, nps)
        assert np.array_equal(miss, ref0)                   # a miss is the uncached path (and fills the cache)
        eng.set_slots(slots)
        hit = eng.evaluate(P, nps)
        assert np.array_equal(hit, ref0)                    # the same parameters again: every lane hits
        # proposals that leave C alone: hits, against a context that has never seen a cache
        Q = _move_planets(P, L, rng)
        eng.set_slots(slots)
        got = eng.evaluate(Q, nps)
        assert np.array_equal(got, _no_cache(data, flags, Q, nps, mode))
        # half the lanes change a hyperparameter or the jitter (misses), the others hit; uncached lanes mixed in
        R = _move_planets(Q, L, rng)
        R[L["row_eta"] + 1, 0::4] *= 1.01
        R[L["row_jit"], 1::4] *= 0.99
        s2 = slots.copy()
        s2[2::8] = -1
        eng.set_slots(s2)
        got = eng.evaluate(R, nps)
        assert np.array_equal(got, _no_cache(data, flags, R, nps, mode))
        # without slots the cache is not consulted
        assert np.array_equal(eng.evaluate(Q, nps), _no_cache(data, flags, Q, nps, mode))


def test_two_entries_per_slot_state_and_proposal(gpu):
    """The accept / reject pattern of a sampler: the state's factor survives a rejected C-changing proposal,
    an accepted one becomes the state, and the entry replaced is always the one used longer ago.  Timed:
    hits have to be visibly cheaper than misses, or the cache is not being used."""
    N, B = 177, 4096
    flags = ModelFlags(GP=True, npmax=5, hyperpriors=True)
    data = make_data(N)
    P, nps = make_params(flags, data, B, "mix", seed=5)
    rng = np.random.default_rng(6)
    with LikelihoodEngine(data, flags, B) as eng:
        L = eng.layout
        assert eng.enable_gp_cache(B) == B
        slots = np.arange(B, dtype=np.int32)

        def run(X):
            eng.set_slots(slots)
            out = eng.evaluate(X, nps)
            return out, eng.last_kernel_ms()

        state = P
        base, t_miss = run(state)                                   # fills entry 0
        prop = _move_planets(state, L, rng)
        prop[L["row_eta"] + 2] *= 1.05                              # changes C: fills entry 1
        lp, t_miss2 = run(prop)
        back, t_hit = run(_move_planets(state, L, rng))             # "rejected": the state's factor is still there
        assert t_hit < 0.6 * min(t_miss, t_miss2), (t_hit, t_miss, t_miss2)
        # a second C-changing proposal replaces the rejected one, not the state
        prop2 = _move_planets(state, L, rng)
        prop2[L["row_jit"]] *= 1.1
        _, t3 = run(prop2)
        _, t4 = run(_move_planets(state, L, rng))
        assert t4 < 0.6 * t3, (t4, t3)
        # "accepted": prop2 is the state now; its factor is there, and the next C-changing proposal
        # replaces the old state's entry
        _, t5 = run(_move_planets(prop2, L, rng))
        assert t5 < 0.6 * t3, (t5, t3)
        prop3 = _move_planets(prop2, L, rng)
        prop3[L["row_eta"]] *= 0.9
        _, t6 = run(prop3)
        _, t7 = run(_move_planets(prop2, L, rng))
        assert t7 < 0.6 * t6, (t7, t6)
    # and every one of those numbers is the uncached number
    with LikelihoodEngine(data, flags, B) as plain:
        assert np.array_equal(base, plain.evaluate(state, nps))
        assert np.array_equal(lp, plain.evaluate(prop, nps))


def test_cache_on_golden_multi_instrument_and_known_objects(gpu):
    """Golden GP cases of the small path (reference logL): multi-instrument jitters are part of the key,
    known objects and trends only move the residuals."""
    for name in ("synth_gp_N33_multi", "synth_gp_N130_trend_ko", "CoRoT7"):
        g = load_golden(name)
        B = g["nps"].size
        with LikelihoodEngine(g["data"], g["flags"], B) as eng:
            assert eng.enable_gp_cache(B) == B
            slots = np.arange(B, dtype=np.int32)
            for _ in range(2):                               # miss, then hit
                eng.set_slots(slots)
                logL = eng.evaluate(g["params"], g["nps"])
                fin = np.isfinite(g["logL"])
                assert np.array_equal(np.isfinite(logL), fin)
                assert np.max(np.abs(logL[fin] - g["logL"][fin]) / np.abs(g["logL"][fin])) <= 1e-10
            # one jitter of one instrument changed in half the lanes: those lanes miss, and say so in the result
            P2 = g["params"].copy()
            Lr = g["layout"]
            P2[Lr["row_jit"] + Lr["n_jit"] - 1, ::2] *= 1.5
            eng.set_slots(slots)
            got = eng.evaluate(P2, g["nps"])
        with LikelihoodEngine(g["data"], g["flags"], B) as plain:
            assert np.array_equal(got, plain.evaluate(P2, g["nps"]))


def test_cache_is_inactive_where_it_does_not_apply(gpu):
    data = make_data(300)   # blocked path
    with LikelihoodEngine(data, ModelFlags(GP=True, npmax=1), 8) as eng:
        assert eng.enable_gp_cache(8) == 0
        eng.set_slots(np.arange(8, dtype=np.int32))          # accepted and ignored
    with LikelihoodEngine(make_data(64), ModelFlags(npmax=1), 8) as eng:
        assert eng.enable_gp_cache(8) == 0
    with LikelihoodEngine(make_data(64), ModelFlags(GP=True, MA=True, npmax=1), 8) as eng:
        assert eng.enable_gp_cache(8) == 0
    with LikelihoodEngine(make_data(64), ModelFlags(GP=True, npmax=1), 8) as eng:
        assert eng.enable_gp_cache(8) == 8
        with pytest.raises(Exception):
            eng.set_slots(np.array([0, 1, 2, 3, 4, 5, 6, 8], np.int32))   # slot outside the cache
        assert eng.enable_gp_cache(0) == 0


def test_device_proposal_chains_do_not_depend_on_the_cache(gpu, tmp_path):
    """examples/CoRoT7 (GP, hyperpriors) with device-side proposals: kc_sampler_create attaches one slot per
    particle; with and without it the run writes the same files, byte for byte."""
    from test_device_proposals import run
    opts = "2\n500\n200\n50\n0\n10\n100\n400\n"   # 2 particles x 64 lanes, 50 thread steps, 400 saves
    off = run("gpu", "CoRoT7", tmp_path, "off", 64, 43, True, opts, {"KIMA_CUDA_GP_CACHE": "0"})
    on = run("gpu", "CoRoT7", tmp_path, "on", 64, 43, True, opts, {"KIMA_CUDA_GP_CACHE": "1"})
    for f in ("sample.txt", "sample_info.txt", "levels.txt", "sampler_state.txt"):
        assert (off / f).read_bytes() == (on / f).read_bytes(), f


def test_host_proposal_chains_do_not_depend_on_the_cache_or_on_the_host_threads(gpu, tmp_path):
    """examples/CoRoT7 with host proposals (the lock-step sampler, two lane groups, one factor-cache slot per
    particle of each group): the same files with and without the cache, and with 1 or 16 host threads.

    The thread count mattered once: with hyperpriors the reference re-parameterises SHARED prior objects before
    every use (src/RVConditionalPrior.cpp:111-112, 135-136, 150-151), which races between threads that perturb
    descendants of one particle; host/RVmodelDevice.h::own_lane_priors gives every lane objects of its own.
    512 lanes x 2 particles, enough steps for kill_lagging_particles to have spread copies over the lanes."""
    from test_device_proposals import run
    opts = "2\n20000\n51200\n100\n0\n10\n100\n40\n"   # 2 particles per lane; a save (and a bookkeeping round) per 100 steps of every lane, 40 of them
    base = run("gpu", "CoRoT7", tmp_path, "off", 512, 43, False, opts, {"KIMA_CUDA_GP_CACHE": "0", "KIMA_HOST_THREADS": "16"})
    for tag, env in (("on", {"KIMA_CUDA_GP_CACHE": "1", "KIMA_HOST_THREADS": "16"}),
                     ("on_again", {"KIMA_CUDA_GP_CACHE": "1", "KIMA_HOST_THREADS": "16"}),
                     ("serial", {"KIMA_CUDA_GP_CACHE": "1", "KIMA_HOST_THREADS": "1"})):
        d = run("gpu", "CoRoT7", tmp_path, tag, 512, 43, False, opts, env)
        for f in ("sample.txt", "sample_info.txt", "levels.txt", "sampler_state.txt"):
            assert (base / f).read_bytes() == (d / f).read_bytes(), (tag, f)


def test_cache_with_chunked_uploads(gpu):
    """kc_eval uploads large batches in chunks that overlap the kernels (option eval_chunk): every chunk is a
    launch sequence of its own over its part of the slots, sharing the classification scratch."""
    N, B = 100, 200
    flags = ModelFlags(GP=True, npmax=3)
    data = make_data(N)
    P, nps = make_params(flags, data, B, "mix", seed=9)
    rng = np.random.default_rng(10)
    with LikelihoodEngine(data, flags, B) as eng:
        eng.set_option("eval_chunk", 64)                     # 200 lanes: chunks of 64, 64, 64, 8
        L = eng.layout
        assert eng.enable_gp_cache(B) == B
        slots = rng.permutation(B).astype(np.int32)          # slots need not follow the lane order
        ref = _no_cache(data, flags, P, nps)
        for _ in range(2):                                   # misses, then hits
            eng.set_slots(slots)
            assert np.array_equal(eng.evaluate(P, nps), ref)
        Q = _move_planets(P, L, rng)
        Q[L["row_eta"] + 2, ::3] *= 1.02                     # a third of the lanes miss
        eng.set_slots(slots)
        assert np.array_equal(eng.evaluate(Q, nps), _no_cache(data, flags, Q, nps))
