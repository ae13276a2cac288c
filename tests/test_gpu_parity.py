"""GPU: the CUDA path, called through the C ABI, against the golden vectors (reference
outputs) and the CPU oracle on the same seeded inputs.

Tolerances (BASELINE.json north_star): logL <= 1e-10 relative, eccentric anomaly <= 1e-12 rad.
"""
import os

import numpy as np
import pytest

from conftest import GOLDEN, golden_cases, load_golden, relerr
from kima_b200 import KimaCudaError, LikelihoodEngine, ModelFlags
from kima_b200.synthetic import make_data, make_params

pytestmark = pytest.mark.gpu

LOGL_RTOL = 1e-10
E_ATOL = 1e-12


def test_kepler_E_and_true_anomaly(gpu):
    z = np.load(os.path.join(GOLDEN, "kepler.npz"))
    with LikelihoodEngine(make_data(8), ModelFlags(), 1) as eng:
        E = eng.kepler(z["M"], z["e"])
        assert np.array_equal(np.isnan(E), np.isnan(z["E"]))
        assert np.nanmax(np.abs(E - z["E"])) <= E_ATOL
        f = eng.true_anomaly(z["t"], z["P"], z["ecc"], z["Tp"])
        # acos is ill-conditioned at f = 0, pi: |df| <~ eps / |sin f|
        bound = 1e-12 + 4e-16 / np.maximum(np.abs(np.sin(z["f"])), 1e-9)
        assert np.all(np.abs(f - z["f"]) <= bound)


@pytest.mark.parametrize("mode", [0, 1])
@pytest.mark.parametrize("name", golden_cases())
def test_golden_logL_mu(gpu, name, mode):
    g = load_golden(name)
    B = g["nps"].size
    with LikelihoodEngine(g["data"], g["flags"], B, kepler_mode=mode) as eng:
        assert eng.layout == g["layout"]
        logL = eng.evaluate(g["params"], g["nps"])
        assert np.array_equal(np.isfinite(logL), np.isfinite(g["logL"]))
        assert np.all(np.isposinf(logL[~np.isfinite(logL)]))  # src/RVmodel.cpp:938-941
        assert relerr(logL, g["logL"]).max() <= LOGL_RTOL
        if np.all(np.isfinite(g["mu0"])):
            mu = eng.calculate_mu(0)
            scale = np.abs(g["mu0"]).max() + np.abs(g["params"][g["layout"]["row_K"]:g["layout"]["row_K"] + max(g["layout"]["npmax"], 1), 0]).sum() + 1.0
            assert np.max(np.abs(mu - g["mu0"])) <= 1e-10 * scale
        if g["C0"] is not None:
            # kc_gp_cov runs the product path's generator (phase table + table-driven exp, kc_gp_common.cuh):
            # the diagonal is the reference's bit for bit; off it, within 4e-15 eta1^2 of the reference's
            # entries (whose own distance from an 80-bit evaluation is ~1e-15 eta1^2, tools/gp_generator_error_budget.py)
            # and within |x| ulp(x) of the exponent, relatively, far down the exponential tail
            C = eng.calculate_C(0)
            C0 = g["C0"]
            eta1sq = g["params"][g["layout"]["row_eta"], 0] ** 2
            assert np.array_equal(np.diag(C), np.diag(C0))
            assert np.array_equal(C, C.T)
            err = np.abs(C - C0)
            assert err.max() <= 4e-15 * eta1sq
            assert np.all(err <= 2e-12 * np.abs(C0) + 1e-290 * eta1sq)


@pytest.mark.parametrize("mode", [0, 1])
def test_ill_conditioned_tail_within_the_reference_bound(gpu, mode):
    """The 48 worst-conditioned lanes of 65 536 prior draws (e -> 1 with an epoch at periastron:
    the reference's acos, src/RVmodel.cpp:1467, loses up to 8 digits of f there).  The GPU must
    stay within 1e-10 |logL| plus the bound derived from the reference's own f, mu and variances
    (tests/golden/make_golden_tail.py); three of the lanes have a bound above 1e-10 |logL|."""
    import json
    z = np.load(os.path.join(GOLDEN, "tail_kep_N200_np5_hyper.npz"))
    flags = ModelFlags(**{k: (bool(v) if isinstance(getattr(ModelFlags(), k), bool) else int(v))
                          for k, v in json.loads(str(z["flags"])).items()})
    d = make_data(200)
    with LikelihoodEngine(d, flags, z["nps"].size, kepler_mode=mode) as eng:
        got = eng.evaluate(z["params"], z["nps"])
    assert (z["bound"] > 1e-10 * np.abs(z["logL"])).sum() >= 3
    assert np.all(np.abs(got - z["logL"]) <= LOGL_RTOL * np.abs(z["logL"]) + z["bound"])
    # and the tail is a tail: the median lane of this worst-of-65536 set is still far inside 1e-10
    assert np.median(relerr(got, z["logL"])) <= 1e-12


FLAG_CASES = [
    ("plain_np0", ModelFlags(npmax=0), 57, 1, 0),
    ("np7", ModelFlags(npmax=7), 300, 1, 0),
    ("hyper", ModelFlags(npmax=4, hyperpriors=True), 190, 1, 0),
    ("studentt", ModelFlags(npmax=2, studentt=True), 68, 1, 0),
    ("trend1_ind", ModelFlags(npmax=1, trend=True, degree=1), 99, 1, 2),
    ("trend3_multi", ModelFlags(npmax=2, trend=True, degree=3, multi_instrument=True), 110, 4, 0),
    ("ko", ModelFlags(npmax=1, known_object=True, n_known_object=3), 64, 1, 0),
    ("npmax32_ko8", ModelFlags(npmax=32, known_object=True, n_known_object=8), 40, 1, 0),
    ("gp", ModelFlags(GP=True, npmax=2), 31, 1, 0),
    ("gp_33", ModelFlags(GP=True, npmax=0), 33, 1, 0),
    ("gp_multi_trend_ind", ModelFlags(GP=True, npmax=2, multi_instrument=True, trend=True, degree=2), 161, 3, 1),
    ("gp_ko_hyper", ModelFlags(GP=True, npmax=5, hyperpriors=True, known_object=True, n_known_object=1), 177, 1, 0),
    ("gp_257", ModelFlags(GP=True, npmax=1), 257, 1, 0),
    ("ma", ModelFlags(npmax=3, MA=True), 300, 1, 0),
    ("ma_1000", ModelFlags(npmax=2, MA=True), 1000, 1, 0),
    ("ma_studentt_multi_trend_ind", ModelFlags(npmax=2, MA=True, studentt=True, multi_instrument=True,
                                               trend=True, degree=3), 97, 3, 2),
    ("gp_ma_small", ModelFlags(GP=True, npmax=2, MA=True), 177, 1, 0),
    ("celerite_1000", ModelFlags(GP=True, kernel=1, npmax=3), 1000, 1, 0),
    ("celerite_ma_trend_ind", ModelFlags(GP=True, kernel=1, npmax=2, MA=True, trend=True, degree=1, hyperpriors=True), 177, 1, 1),
    ("celerite_tiny", ModelFlags(GP=True, kernel=1, npmax=0), 2, 1, 0),
    ("gp_ma_blocked", ModelFlags(GP=True, npmax=2, MA=True, multi_instrument=True), 300, 2, 0),
]


@pytest.mark.parametrize("case", FLAG_CASES, ids=lambda c: c[0])
def test_seeded_batches_vs_oracle(gpu, port, case):
    _, flags, N, n_inst, n_ind = case
    d = make_data(N, n_inst, n_ind, seed=1000 + N)
    B = 48 if flags.GP else 300   # 300 < 148*16: the CTA-per-particle tiling
    P, nps = make_params(flags, d, B, "mix", seed=N)
    port.set_rvdata(d)
    port.set_flags(flags)
    ref, _ = port.eval_batch(P, nps, nthreads=8)
    with LikelihoodEngine(d, flags, B) as eng:
        got = eng.evaluate(P, nps)
    assert relerr(got, ref).max() <= LOGL_RTOL


@pytest.mark.parametrize("N", [3, 63, 64, 65, 127, 128, 129, 191, 192, 193, 320])
def test_gp_block_column_edges(gpu, port, N):
    """The blocked factorisation at every alignment of N against its 64-wide block columns
    and 128-row tiles (partial last block, residual row alone in a tile, one block only)."""
    flags = ModelFlags(GP=True, npmax=2)
    d = make_data(N, seed=4000 + N)
    B = 40
    P, nps = make_params(flags, d, B, "mix", seed=N)
    port.set_rvdata(d)
    port.set_flags(flags)
    ref, _ = port.eval_batch(P, nps, nthreads=8)
    with LikelihoodEngine(d, flags, B, gp_slots=16) as eng:   # 40 lanes in waves of 16, 16, 8
        eng.set_option("gp_small", 0)                         # the blocked path at every N
        got = eng.evaluate(P, nps)
        again = eng.evaluate(P, nps)
    assert relerr(got, ref).max() <= LOGL_RTOL
    assert np.array_equal(got, again)  # deterministic: no atomics on the data path


@pytest.mark.parametrize("streams", [1, 2, 3, 4])
def test_gp_blocked_lane_groups_and_waves(gpu, port, streams):
    """The blocked path cuts a batch into concurrent lane groups (one stream each) and every
    group into waves of its share of the slots: 700 lanes, 4 groups x 3 waves of 60 slots at
    most, ragged last wave; results do not depend on the grouping."""
    flags = ModelFlags(GP=True, npmax=1)
    N = 70
    d = make_data(N, seed=4100)
    B = 700
    P, nps = make_params(flags, d, B, "mix", seed=N)
    port.set_rvdata(d)
    port.set_flags(flags)
    ref, _ = port.eval_batch(P, nps, nthreads=8)
    with LikelihoodEngine(d, flags, B, gp_slots=240, gp_streams=streams) as eng:
        eng.set_option("gp_small", 0)
        got = eng.evaluate(P, nps)
    assert relerr(got, ref).max() <= LOGL_RTOL
    with LikelihoodEngine(d, flags, B, gp_streams=1) as eng:
        eng.set_option("gp_small", 0)
        one = eng.evaluate(P, nps)
    assert np.array_equal(got, one)


@pytest.mark.parametrize("N", [2, 3, 31, 32, 33, 64, 65, 95, 96, 97, 160, 161, 177, 192, 208, 211, 212])
def test_gp_small_path_edges(gpu, port, N):
    """The shared-memory-resident kernel (one CTA per particle, N <= 211) at every alignment of
    N against its 32-wide block columns, 16-row update strips and 32-row substitution tiles;
    N = 212 is the first size that no longer fits and must take the blocked path.  Both paths
    against the oracle and against each other."""
    flags = ModelFlags(GP=True, npmax=2, multi_instrument=(N % 2 == 0 and N > 8))
    d = make_data(N, 3 if flags.multi_instrument else 1, seed=7000 + N)
    B = 200  # more CTAs than SMs
    P, nps = make_params(flags, d, B, "mix", seed=N)
    port.set_rvdata(d)
    port.set_flags(flags)
    ref, _ = port.eval_batch(P, nps, nthreads=8)
    with LikelihoodEngine(d, flags, B) as eng:
        n0 = eng.launch_count()
        got = eng.evaluate(P, nps)
        launches = eng.launch_count() - n0
        again = eng.evaluate(P, nps)
        eng.set_option("gp_small", 0)
        blocked = eng.evaluate(P, nps)
    assert (launches == 1) == (N <= 211)   # one launch for the whole likelihood
    assert relerr(got, ref).max() <= LOGL_RTOL
    assert np.array_equal(got, again)
    assert relerr(got, blocked).max() <= LOGL_RTOL


@pytest.mark.parametrize("N,mode", [(64, "mix"), (100, "mix"), (177, "mix"), (177, "max")])
def test_gp_small_path_lane_order_does_not_matter(gpu, port, N, mode):
    """Batches of >= 1024 lanes reach the SMs in order of planet count (gps_order_kernel, option gp_order:
    it keeps the SMs in step).  The order is scheduling only: the same bits with it and without it, from
    host buffers (kc_eval knows when all counts are equal and skips it) and from the device block, and
    the oracle's values."""
    flags = ModelFlags(GP=True, npmax=5, hyperpriors=True)
    d = make_data(N, seed=8000 + N)
    B = 3000
    P, nps = make_params(flags, d, B, mode, seed=N)
    port.set_rvdata(d)
    port.set_flags(flags)
    ref, _ = port.eval_batch(P[:, :256], nps[:256], nthreads=8)
    with LikelihoodEngine(d, flags, B) as eng:
        ordered = eng.evaluate(P, nps)
        n0 = eng.launch_count()
        eng.log_likelihood_device(B)
        eng.sync()
        assert eng.launch_count() - n0 == 2           # the ordering kernel + the likelihood kernel
        eng.set_option("gp_order", 0)
        plain = eng.evaluate(P, nps)
        n0 = eng.launch_count()
        eng.log_likelihood_device(B)
        eng.sync()
        assert eng.launch_count() - n0 == 1
    assert np.array_equal(ordered, plain)
    assert relerr(ordered[:256], ref).max() <= LOGL_RTOL


def test_ragged_batches_and_reuse(gpu, port):
    """B = 1, odd sizes, reuse of one context for shrinking and growing batches."""
    flags = ModelFlags(npmax=3)
    d = make_data(77)
    P, nps = make_params(flags, d, 4000, "mix", seed=5)
    port.set_rvdata(d)
    port.set_flags(flags)
    ref, _ = port.eval_batch(P, nps, nthreads=8)
    with LikelihoodEngine(d, flags, 4000) as eng:
        eng.set_params(P, nps)
        for B in (1, 2, 31, 33, 255, 2367, 2368, 2369, 4000, 7):
            got = eng.log_likelihood(B)
            assert relerr(got, ref[:B]).max() <= LOGL_RTOL
        assert eng.log_likelihood(0).size == 0


def test_full_size_keplerian_vs_oracle(gpu, port):
    """BASELINE workload K: N=1000, Np=10, 8192 lanes (warp-per-particle tiling)."""
    flags = ModelFlags(npmax=10)
    d = make_data(1000)
    B = 8192
    P, nps = make_params(flags, d, B, "max")
    port.set_rvdata(d)
    port.set_flags(flags)
    ref, _ = port.eval_batch(P, nps, nthreads=os.cpu_count() or 8)
    with LikelihoodEngine(d, flags, B) as eng:
        got = eng.evaluate(P, nps)
        rel = relerr(got, ref)
        assert rel.max() <= LOGL_RTOL
        assert np.median(rel) <= 1e-14
        # size-independent properties: lane permutation and batch splitting change nothing
        perm = np.random.default_rng(0).permutation(B)
        got_p = eng.evaluate(P[:, perm], nps[perm])
        assert np.array_equal(got_p, got[perm])
        half = eng.evaluate(P[:, :4096], nps[:4096])
        assert np.array_equal(half, got[:4096])
        # the two tilings (warp / CTA per particle) agree to rounding
        small = eng.evaluate(P[:, :256], nps[:256])
        assert relerr(small, got[:256]).max() <= 1e-13


def test_np0_equals_closed_form_and_K0_equals_np0(gpu):
    flags = ModelFlags(npmax=4)
    d = make_data(500)
    P, nps = make_params(flags, d, 64, "max")
    L = None
    with LikelihoodEngine(d, flags, 64) as eng:
        L = eng.layout
        with_planets = eng.evaluate(P, nps)
        P0 = P.copy()
        P0[L["row_K"]:L["row_K"] + 4] = 0.0       # zero amplitude == no planet
        k0 = eng.evaluate(P0, nps)
        none = eng.evaluate(P, np.zeros_like(nps))
    var = d.sig[:, None] ** 2 + P[L["row_jit"]][None, :] ** 2
    r = d.y[:, None] - P[L["row_vsys"]][None, :]
    closed = np.sum(-0.5 * np.log(2 * np.pi) - 0.5 * np.log(var) - 0.5 * r * r / var, axis=0)
    assert relerr(none, closed).max() <= 1e-13
    assert relerr(k0, none).max() <= 1e-15
    assert not np.allclose(with_planets, none)


def test_gp_with_vanishing_amplitude_equals_white_noise(gpu):
    """Cross-family property: eta1 -> 0 makes C diagonal, so the GP kernel family must
    reproduce the Gaussian white-noise family on the same parameters."""
    N = 300
    d = make_data(N)
    fg = ModelFlags(GP=True, npmax=3)
    fw = ModelFlags(npmax=3)
    Pg, nps = make_params(fg, d, 40, "mix", seed=3)
    with LikelihoodEngine(d, fg, 40) as eg:
        Lg = eg.layout
        Pg[Lg["row_eta"]] = 1e-160   # eta1^2 underflows to 0
        gp = eg.evaluate(Pg, nps)
    with LikelihoodEngine(d, fw, 40) as ew:
        Pw = Pg[:ew.layout["n_rows"]].copy()
        assert ew.layout["row_jit"] == Lg["row_jit"]
        wn = ew.evaluate(Pw, nps)
    assert relerr(gp, wn).max() <= 1e-12


def test_full_size_gp_vs_oracle(gpu, port):
    """BASELINE workload G: QP-GP, N=1000 (+ Np=10)."""
    flags = ModelFlags(GP=True, npmax=10)
    d = make_data(1000)
    B = 64
    P, nps = make_params(flags, d, B, "max")
    port.set_rvdata(d)
    port.set_flags(flags)
    ref, _ = port.eval_batch(P, nps, nthreads=os.cpu_count() or 8)
    with LikelihoodEngine(d, flags, 400, gp_slots=48) as eng:  # 400 lanes in waves of 48 (+16)
        got = eng.evaluate(P, nps)
        assert relerr(got, ref).max() <= LOGL_RTOL
        # more lanes than workspace slots: slots are recycled, results unchanged
        P2 = np.tile(P, (1, 6))[:, :400]
        n2 = np.tile(nps, 6)[:400]
        got2 = eng.evaluate(P2, n2)
        assert np.array_equal(got2[:64], got) and np.array_equal(got2[64:128], got)


def test_scale_point_gp_N4096_vs_reference(gpu):
    """BASELINE.json configs[4]: N=4096, QP-GP + Np<=10, against the reference's own
    calculate_mu / calculate_C / log_likelihood (tests/golden/synth_gp_N4096.npz, made by
    oracle/_ref/libkima_ref.so) at the north-star tolerance, over several waves of the factor
    storage (134 MB per lane in flight) and both lane-group settings."""
    g = load_golden("synth_gp_N4096")
    B = g["nps"].size
    with LikelihoodEngine(g["data"], g["flags"], 3 * B, gp_slots=3) as eng:   # waves of 3 slots
        got = eng.evaluate(g["params"], g["nps"])
        assert relerr(got, g["logL"]).max() <= LOGL_RTOL, relerr(got, g["logL"])
        # three times the lanes through the same three slots: 4 waves, results unchanged
        got3 = eng.evaluate(np.tile(g["params"], (1, 3)), np.tile(g["nps"], 3))
        assert np.array_equal(got3.reshape(3, B), np.tile(got, (3, 1)))
    with LikelihoodEngine(g["data"], g["flags"], B, gp_streams=1) as eng:     # all lanes in flight, one group
        assert np.array_equal(eng.evaluate(g["params"], g["nps"]), got)


def test_scale_point_gp_N4096_vs_lapack(gpu, port):
    """configs[4] shape: N=4096, QP-GP + Np<=10.  The C oracle supplies mu; LAPACK the factorisation."""
    from oracle.gp_numpy import gp_loglik, qp_cov
    N = 4096
    flags = ModelFlags(GP=True, npmax=10)
    d = make_data(N)
    P, nps = make_params(flags, d, 3, "mix", seed=11)
    port.set_rvdata(d)
    with LikelihoodEngine(d, flags, 3) as eng:
        L = eng.layout
        got = eng.evaluate(P, nps)
    port.set_flags(ModelFlags(npmax=10))  # same planet rows; mu does not depend on GP
    for b in range(3):
        Pw = P[:port.layout()["n_rows"]]
        mu = port.mu(Pw, b, int(nps[b]))
        C = qp_cov(d.t, d.sig, np.full(N, P[L["row_jit"], b] ** 2), *P[L["row_eta"]:L["row_eta"] + 4, b])
        ref = gp_loglik(d.y - mu, C)
        assert abs(got[b] - ref) <= 1e-9 * abs(ref)  # LAPACK's own rounding at N=4096, cond ~1e8


def test_gp_blocked_more_lanes_in_flight_than_a_grid_dimension(gpu):
    """The blocked path puts the lanes of a wave in gridDim.y (<= 65535): a batch whose lane
    groups exceed that must run as several waves, not fail with 'invalid configuration'."""
    flags = ModelFlags(GP=True, npmax=1)
    d = make_data(40)
    nd, B = 1000, 140000   # two lane groups of 70000 > 65535
    P, nps = make_params(flags, d, nd, "mix")
    with LikelihoodEngine(d, flags, B) as eng:
        eng.set_option("gp_small", 0)
        want = eng.evaluate(P, nps)
        got = eng.evaluate(np.tile(P, (1, B // nd)), np.tile(nps, B // nd))
        assert np.array_equal(got.reshape(B // nd, nd), np.tile(want, (B // nd, 1)))


def test_non_positive_definite_gives_plus_inf(gpu):
    flags = ModelFlags(GP=True, npmax=0)
    d = make_data(60)
    with LikelihoodEngine(d, flags, 2) as eng:
        L = eng.layout
        P = np.zeros((L["n_rows"], 2))
        P[L["row_eta"]:L["row_eta"] + 4, :] = [[5.0, 5.0], [50.0, 50.0], [20.0, 20.0], [1.0, 1.0]]
        P[L["row_jit"]] = [1.0, np.nan]
        out = eng.evaluate(P, np.zeros(2, np.int32))
        assert np.isfinite(out[0]) and np.isposinf(out[1])
        assert eng.likelihood_type_value(out)[1] == -1e300


def test_async_and_device_resident_paths(gpu):
    flags = ModelFlags(npmax=2)
    d = make_data(128)
    P, nps = make_params(flags, d, 512, "mix")
    with LikelihoodEngine(d, flags, 512) as eng:
        ref = eng.evaluate(P, nps)
        out = np.zeros(512)
        eng.log_likelihood_async(512, out)
        with pytest.raises(KimaCudaError):
            eng.log_likelihood_async(512, out)   # one evaluation in flight per context
        eng.wait()
        assert np.array_equal(out, ref)
        n0 = eng.launch_count()
        eng.log_likelihood_device(512)
        eng.sync()
        assert eng.launch_count() == n0 + 1 and eng.last_kernel_ms() > 0


def test_error_behaviour(gpu):
    d = make_data(32)
    with pytest.raises(KimaCudaError) as e:
        LikelihoodEngine(make_data(32, 2), ModelFlags(kernel=1, GP=True, multi_instrument=True), 4)
    assert e.value.code == 3  # KC_ERR_UNSUPPORTED: the reference reads an unset jitter there
    with pytest.raises(KimaCudaError):
        LikelihoodEngine(d, ModelFlags(kernel=2, GP=True), 4)
    with pytest.raises(KimaCudaError):
        LikelihoodEngine(d, ModelFlags(npmax=33), 4)
    with LikelihoodEngine(d, ModelFlags(), 4) as eng:
        with pytest.raises(KimaCudaError):
            eng.log_likelihood(5)
        with pytest.raises(ValueError):
            eng.set_params(np.zeros((99, 2)), np.zeros(2, np.int32))
        with pytest.raises(KimaCudaError):
            eng.calculate_C(0)  # needs GP=true
