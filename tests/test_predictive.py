"""Posterior-predictive path.

CPU part: the oracle's restatement (oracle/predictive.py) against vectors produced by
pykima's own keplerian() / GP.predict() (tests/golden/make_golden_predictive.py), and the
sample.txt column map against the reference's CI fixtures.
GPU part: kc_predict through the C ABI against the oracle and the golden vectors.

Tolerances: deterministic model 1e-10 of its scale against the oracle (same solver);
1e-7 against pykima (its Newton iteration stops at an L1 step of 1.234e-10 over the whole
array, keplerian.py:32).  GP mean: 1e-8 of max|mean| against the oracle and pykima (both
solve with LAPACK; the difference is cond(C) * eps).
"""
import json
import os

import numpy as np
import pytest

from conftest import GOLDEN
from kima_b200 import (KimaCudaError, LikelihoodEngine, ModelFlags, PosteriorPredictive, RVData,
                       params_from_samples, sample_columns)
from kima_b200.synthetic import make_data, make_params
from oracle import predictive as opred
from oracle.oracle import Oracle

CASES = ["predictive_gp_N60", "predictive_gp_N300_trend", "predictive_kep_N120_hyper"]


def load_case(name):
    z = np.load(os.path.join(GOLDEN, name + ".npz"))
    flags = ModelFlags(**{k: (bool(v) if isinstance(getattr(ModelFlags(), k), bool) else int(v))
                          for k, v in json.loads(str(z["flags"])).items()})
    data = RVData.from_arrays(z["t"], z["y"], z["sig"])
    return z, flags, data


@pytest.mark.parametrize("name", CASES)
def test_oracle_predictive_against_pykima(name):
    z, flags, data = load_case(name)
    orc = Oracle("port")
    orc.set_rvdata(data)
    orc.set_flags(flags)
    P, nps = z["params"], z["nps"]
    for b in range(nps.size):
        m = opred.model(orc, flags, data, P, b, nps[b], z["tgrid"])
        scale = np.abs(z["model_grid"][b]).max() + 1.0
        assert np.max(np.abs(m - z["model_grid"][b])) <= 1e-7 * scale
        md = opred.model(orc, flags, data, P, b, nps[b])
        assert np.max(np.abs(md - z["model_data"][b])) <= 1e-7 * scale
        if flags.GP:
            g = opred.gp_mean(orc, flags, data, P, b, nps[b], z["tgrid"])
            assert np.max(np.abs(g - z["gp_grid"][b])) <= 1e-8 * (np.abs(z["gp_grid"][b]).max() + 1.0)
            # return_std=True (GP.py:113-131): pykima forms C^-1 explicitly, which costs it digits
            sd = opred.gp_std(orc, flags, data, P, b, nps[b], z["tgrid"])
            eta1 = P[orc.layout()["row_eta"], b]
            assert np.max(np.abs(sd ** 2 - z["gp_std"][b] ** 2)) <= 1e-7 * eta1 ** 2


CELERITE_CASES = ["predictive_celerite_N200", "predictive_celerite_N90_trend"]


@pytest.mark.parametrize("name", CELERITE_CASES)
def test_oracle_celerite_predict_against_the_reference_solver(name):
    """oracle.predictive.gp_mean_celerite (numpy restatement of CholeskySolver::compute / solve / predict)
    against vectors made by the reference's own solver (make_golden_predictive.py: celerite_case)."""
    z, flags, data = load_case(name)
    orc = Oracle("port")
    orc.set_rvdata(data)
    orc.set_flags(flags)
    P, nps = z["params"], z["nps"]
    for b in range(nps.size):
        g = opred.gp_mean_celerite(orc, flags, data, P, b, nps[b], z["tgrid"])
        assert np.max(np.abs(g - z["gp_grid"][b])) <= 1e-11 * (np.abs(z["gp_grid"][b]).max() + 1.0)
        gd = opred.gp_mean_celerite(orc, flags, data, P, b, nps[b])
        assert np.max(np.abs(gd - z["gp_data"][b])) <= 1e-11 * (np.abs(z["gp_data"][b]).max() + 1.0)


@pytest.mark.gpu
@pytest.mark.parametrize("name", CELERITE_CASES)
def test_gpu_celerite_predict_against_the_reference_solver(gpu, name):
    """kc_predict with kernel=celerite (GP_celerite.predict, pykima/GP.py:349-355)."""
    z, flags, data = load_case(name)
    P, nps = z["params"], z["nps"]
    B = nps.size
    orc = Oracle("port")
    orc.set_rvdata(data)
    orc.set_flags(flags)
    with LikelihoodEngine(data, flags, B) as eng:
        m, g = eng.predict(P, nps, z["tgrid"])
        md, gd = eng.predict(P, nps, None)
        with pytest.raises(KimaCudaError):
            eng.predict(P, nps, z["tgrid"][::-1].copy())       # celerite walks a sorted grid
        with pytest.raises(KimaCudaError):
            eng.predict_std(P, nps, z["tgrid"])                 # no predictive variance for celerite
    for b in range(B):
        gs = np.abs(z["gp_grid"][b]).max() + 1.0
        assert np.max(np.abs(g[b] - z["gp_grid"][b])) <= 1e-8 * gs
        assert np.max(np.abs(gd[b] - z["gp_data"][b])) <= 1e-8 * gs
        assert np.max(np.abs(g[b] - opred.gp_mean_celerite(orc, flags, data, P, b, nps[b], z["tgrid"]))) <= 1e-8 * gs
        scale = np.abs(m[b]).max() + 1.0
        assert np.max(np.abs(m[b] - opred.model(orc, flags, data, P, b, nps[b], z["tgrid"]))) <= 1e-10 * scale


@pytest.mark.gpu
def test_gpu_celerite_predict_many_lanes_and_failures(gpu):
    """More lanes than a warp, MA residuals, a lane whose factorisation fails (NaN row)."""
    flags = ModelFlags(GP=True, kernel=1, npmax=2, MA=True)
    data = make_data(120)
    B = 200
    P, nps = make_params(flags, data, B, "mix", seed=21)
    orc = Oracle("port")
    orc.set_rvdata(data)
    orc.set_flags(flags)
    L = orc.layout()
    P[L["row_jit"], 5] = np.nan
    tg = np.linspace(data.t.min() - 2.0, data.t.max() + 2.0, 301)
    with LikelihoodEngine(data, flags, B) as eng:
        _, g = eng.predict(P, nps, tg, model=False)
    assert np.all(np.isnan(g[5]))
    for b in (0, 17, 64, B - 1):
        fac = opred.celerite_factor(opred.celerite_terms([P[L["row_eta"] + k, b] for k in range(3)]), data.t,
                                    data.sig ** 2 + P[L["row_jit"], b] ** 2)
        # MA residuals: y - mu with the MA(1) term, as the likelihood's celerite branch uses them
        alpha = opred.celerite_solve(fac, data.y - orc.mu(P, b, nps[b]))
        want = opred.celerite_predict(opred.celerite_terms([P[L["row_eta"] + k, b] for k in range(3)]), data.t, alpha, tg)
        assert np.max(np.abs(g[b] - want)) <= 1e-8 * (np.abs(want).max() + 1.0)


@pytest.mark.parametrize("example,flags,n_inst", [
    ("51Peg", ModelFlags(npmax=1), 1),
    ("CoRoT7", ModelFlags(GP=True, hyperpriors=True, npmax=5), 1),
    ("multi_instrument", ModelFlags(multi_instrument=True, npmax=1), 4),
    ("studentT", ModelFlags(studentt=True, npmax=1), 1)])
def test_sample_columns_match_reference_fixtures(example, flags, n_inst):
    """The column map reproduces the header RVmodel::description() wrote (src/RVmodel.cpp:1011-1091)."""
    path = os.path.join(GOLDEN, "e2e", example + "_sample_seed43.txt")
    with open(path) as fh:
        header = fh.readline().lstrip("#").split()
    rows = np.atleast_2d(np.loadtxt(path))
    idx = sample_columns(flags, n_inst, 0)
    assert idx["n_columns"] == rows.shape[1]
    # with hyperpriors description() runs "muK" and "Np" together without a separator
    # (src/RVmodel.cpp:1068-1071): one token short
    assert len(header) == rows.shape[1] - (1 if flags.hyperpriors else 0)
    if not flags.hyperpriors:
        assert header[idx["vsys"].start] == "vsys"
        assert header[idx["np"].start] == "Np"
        assert header[idx["P"].start] == "P0" and header[idx["w"].start] == "w0"
        assert header[idx["jitter"].start] in ("extra_sigma", "jitter1")
    assert np.all(rows[:, idx["ndim"]] == 5) and np.all(rows[:, idx["npmax"]] == flags.npmax)
    P, nps = params_from_samples(rows, flags, n_inst, 0)
    assert np.array_equal(nps, rows[:, idx["np"]].ravel().astype(int))
    assert np.all(nps >= 0) and np.all(nps <= flags.npmax)
    with pytest.raises(ValueError):
        params_from_samples(rows[:, :-1], flags, n_inst, 0)


# ---- GPU -------------------------------------------------------------------------------------
@pytest.mark.gpu
@pytest.mark.parametrize("name", CASES)
def test_gpu_predict_against_pykima_and_oracle(gpu, name):
    z, flags, data = load_case(name)
    P, nps = z["params"], z["nps"]
    B = nps.size
    orc = Oracle("port")
    orc.set_rvdata(data)
    orc.set_flags(flags)
    with LikelihoodEngine(data, flags, B) as eng:
        m, g = eng.predict(P, nps, z["tgrid"])
        md, gd = eng.predict(P, nps, None)
        assert eng.launch_count() >= 2
    for b in range(B):
        scale = np.abs(z["model_grid"][b]).max() + 1.0
        assert np.max(np.abs(m[b] - z["model_grid"][b])) <= 1e-7 * scale
        assert np.max(np.abs(m[b] - opred.model(orc, flags, data, P, b, nps[b], z["tgrid"]))) <= 1e-10 * scale
        assert np.max(np.abs(md[b] - opred.model(orc, flags, data, P, b, nps[b]))) <= 1e-10 * scale
        if flags.GP:
            gs = np.abs(z["gp_grid"][b]).max() + 1.0
            assert np.max(np.abs(g[b] - z["gp_grid"][b])) <= 1e-8 * gs
            assert np.max(np.abs(g[b] - opred.gp_mean(orc, flags, data, P, b, nps[b], z["tgrid"]))) <= 1e-8 * gs
            assert np.max(np.abs(gd[b] - opred.gp_mean(orc, flags, data, P, b, nps[b]))) <= 1e-8 * gs
    if not flags.GP:
        assert g is None


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["predictive_gp_N60", "predictive_gp_N300_trend"])
def test_gpu_predict_std_against_pykima_and_oracle(gpu, name):
    """kc_predict_std: GP.predict(return_std=True), pykima/GP.py:113-131."""
    z, flags, data = load_case(name)
    P, nps = z["params"], z["nps"]
    B = nps.size
    orc = Oracle("port")
    orc.set_rvdata(data)
    orc.set_flags(flags)
    eta1 = P[orc.layout()["row_eta"]]
    with LikelihoodEngine(data, flags, B) as eng:
        g, sd = eng.predict_std(P, nps, z["tgrid"])
        _, sdd = eng.predict_std(P, nps, None, mean=False)
        _, g0 = eng.predict(P, nps, z["tgrid"], model=False)
    assert np.array_equal(g, g0)   # the mean of kc_predict_std is kc_predict's
    for b in range(B):
        var = eta1[b] ** 2
        want = opred.gp_std(orc, flags, data, P, b, nps[b], z["tgrid"])
        assert np.max(np.abs(sd[b] ** 2 - want ** 2)) <= 1e-8 * var, b
        assert np.max(np.abs(sd[b] ** 2 - z["gp_std"][b] ** 2)) <= 1e-7 * var, b
        wd = opred.gp_std(orc, flags, data, P, b, nps[b])
        assert np.max(np.abs(sdd[b] ** 2 - wd ** 2)) <= 1e-8 * var, b


@pytest.mark.gpu
@pytest.mark.parametrize("N,B,nt,slots", [(1000, 6, 333, 4), (97, 40, 50, 7), (2100, 2, 21, 0)])
def test_gpu_predict_std_sizes_and_waves(gpu, N, B, nt, slots):
    """N=1000 (16 right-hand sides per tile), N=2100 (8 per tile), waves of the factor storage,
    grid sizes that are no multiple of the tile, a lane whose C is not positive definite."""
    flags = ModelFlags(GP=True, npmax=1)
    data = make_data(N)
    P, nps = make_params(flags, data, B, "mix", seed=5)
    L = None
    tg = np.linspace(data.t.min() - 3.0, data.t.max() + 3.0, nt)
    orc = Oracle("port")
    orc.set_rvdata(data)
    orc.set_flags(flags)
    L = orc.layout()
    P[L["row_jit"], 1] = np.nan
    with LikelihoodEngine(data, flags, B, gp_slots=slots) as eng:
        g, sd = eng.predict_std(P, nps, tg)
    assert np.all(np.isnan(sd[1])) and np.all(np.isnan(g[1]))
    for b in {0, B - 1} - {1}:
        want = opred.gp_std(orc, flags, data, P, b, nps[b], tg)
        assert np.max(np.abs(sd[b] ** 2 - want ** 2)) <= 1e-8 * P[L["row_eta"], b] ** 2


@pytest.mark.gpu
@pytest.mark.parametrize("flags,N,n_inst,n_ind", [
    (ModelFlags(GP=True, npmax=2, multi_instrument=True), 97, 3, 0),
    (ModelFlags(GP=True, npmax=1, known_object=True, n_known_object=2, trend=True, degree=3), 130, 1, 1),
    (ModelFlags(GP=True, npmax=3, hyperpriors=True), 1000, 1, 0)])
def test_gpu_predict_waves_and_flags(gpu, flags, N, n_inst, n_ind):
    """More lanes than factor slots (several waves, two lane groups), partial blocks,
    offsets / indicators at the data epochs, known objects on the grid."""
    data = make_data(N, n_inst=n_inst, n_ind=n_ind)
    B = 300 if N < 500 else 12
    P, nps = make_params(flags, data, B, "mix", seed=11)
    tg = np.linspace(data.t.min() - 5.0, data.t.max() + 5.0, 611)
    orc = Oracle("port")
    orc.set_rvdata(data)
    orc.set_flags(flags)
    with LikelihoodEngine(data, flags, B, gp_slots=7 if N < 500 else 5) as eng:
        m, g = eng.predict(P, nps, tg)
        md, gd = eng.predict(P, nps, None)
        logL = eng.evaluate(P, nps)
    for b in list(range(0, B, max(1, B // 9))) + [B - 1]:
        scale = np.abs(m[b]).max() + 1.0
        assert np.max(np.abs(m[b] - opred.model(orc, flags, data, P, b, nps[b], tg))) <= 1e-10 * scale
        assert np.max(np.abs(md[b] - opred.model(orc, flags, data, P, b, nps[b]))) <= 1e-10 * scale
        want = opred.gp_mean(orc, flags, data, P, b, nps[b], tg)
        if not np.isfinite(logL[b]):
            assert np.all(np.isnan(g[b]))
            continue
        gs = np.abs(want).max() + 1.0
        assert np.max(np.abs(g[b] - want)) <= 1e-8 * gs
        assert np.max(np.abs(gd[b] - opred.gp_mean(orc, flags, data, P, b, nps[b]))) <= 1e-8 * gs


@pytest.mark.gpu
def test_gpu_posterior_predictive_from_sample_rows(gpu):
    """Rows in sample.txt order -> model / model_stochastic / full_model / residuals."""
    flags = ModelFlags(GP=True, hyperpriors=True, npmax=5)
    data = RVData.load(os.path.join(GOLDEN, "e2e", "corot7.txt"), "ms", 2)
    rows = np.atleast_2d(np.loadtxt(os.path.join(GOLDEN, "e2e", "CoRoT7_sample_seed43.txt")))
    P, nps = params_from_samples(rows, flags, 1, 0)
    orc = Oracle("port")
    orc.set_rvdata(data)
    orc.set_flags(flags)
    tg = np.linspace(data.t.min(), data.t.max(), 400)
    with PosteriorPredictive(data, flags, max_batch=3) as pp:  # forces several calls
        m = pp.model(rows, tg)
        s = pp.model_stochastic(rows, tg)
        f = pp.full_model(rows, tg)
        res = pp.residuals(rows, full=True)
        s2, sd = pp.model_stochastic_std(rows, tg)
    assert np.array_equal(s2, s)
    for b in range(rows.shape[0]):
        want = opred.gp_std(orc, flags, data, P, b, nps[b], tg)
        assert np.max(np.abs(sd[b] ** 2 - want ** 2)) <= 1e-8 * P[orc.layout()["row_eta"], b] ** 2
    assert m.shape == (rows.shape[0], 400)
    np.testing.assert_allclose(f, m + s, rtol=0, atol=1e-9)
    for b in range(rows.shape[0]):
        scale = np.abs(m[b]).max() + 1.0
        assert np.max(np.abs(m[b] - opred.model(orc, flags, data, P, b, nps[b], tg))) <= 1e-10 * scale
        want = opred.gp_mean(orc, flags, data, P, b, nps[b], tg)
        assert np.max(np.abs(s[b] - want)) <= 1e-8 * (np.abs(want).max() + 1.0)
        full_d = opred.model(orc, flags, data, P, b, nps[b]) + opred.gp_mean(orc, flags, data, P, b, nps[b])
        assert np.max(np.abs(res[b] - (data.y - full_d))) <= 1e-8 * (np.abs(full_d).max() + 1.0)


@pytest.mark.gpu
def test_gpu_predict_errors(gpu):
    d = make_data(40)
    P, nps = make_params(ModelFlags(npmax=1), d, 4, "mix")
    with LikelihoodEngine(d, ModelFlags(npmax=1), 4) as eng:
        with pytest.raises(KimaCudaError):
            eng.predict(P, nps, np.arange(5.0), gp=True)          # no GP in this model
        with pytest.raises(KimaCudaError):
            eng.predict(P, nps, np.arange(5.0), model=False, gp=False)
        m, g = eng.predict(P, nps, np.arange(5.0) + 50000.0)
        assert m.shape == (4, 5) and g is None
    fc = ModelFlags(GP=True, kernel=1, npmax=1)
    Pc, npc = make_params(fc, d, 2, "mix")
    with LikelihoodEngine(d, fc, 2) as eng:
        with pytest.raises(KimaCudaError):
            eng.predict_std(Pc, npc, np.arange(5.0))              # celerite: the mean only
