"""CPU: the plain-C oracle against the committed golden vectors (generated from the
unmodified reference by tests/golden/make_golden.py) and the reference's own KATs."""
import math
import os

import numpy as np
import pytest

from conftest import GOLDEN, golden_cases, load_golden, relerr


def test_kepler_golden_exact(port):
    z = np.load(os.path.join(GOLDEN, "kepler.npz"))
    E = port.kepler(z["M"], z["e"])
    # same libm, same operation order: the restatement is bit-identical here
    assert np.array_equal(E, z["E"], equal_nan=True)
    f = port.true_anomaly(z["t"], z["P"], z["ecc"], z["Tp"])
    assert np.array_equal(f, z["f"], equal_nan=True)


def test_kepler_reference_known_answers(port):
    # /root/reference/tests/test_kima.py:15-29
    E = port.kepler(np.array([0.0, np.pi, 2 * np.pi, 0.0]), np.array([0.0, 0.0, 0.0, 0.2]))
    np.testing.assert_allclose(E[:2], [0.0, np.pi], atol=1e-15)
    assert E[3] == 0.0
    # npy_mod wraps 2*pi to 0 for the C++ solver (the Python one returns 2*pi)
    assert abs(math.sin(E[2])) < 1e-15
    E80 = port.kepler(np.array([math.radians(80.0)]), np.array([0.2]))[0]
    assert math.degrees(E80) == pytest.approx(91.45545886486815, rel=1e-8)


def test_keplerian_circular_equals_K_at_t0(port):
    # /root/reference/tests/test_kima.py:50-51: circular orbit, K=1, at t=t0 -> 1
    from kima_b200.data import RVData
    from kima_b200.flags import ModelFlags
    t = np.array([10.0, 11.0, 12.5])
    d = RVData.from_arrays(t, np.zeros(3), np.ones(3))
    port.set_rvdata(d)
    port.set_flags(ModelFlags(npmax=1))
    L = port.layout()
    P = np.zeros((L["n_rows"], 1))
    P[L["row_P"]], P[L["row_K"]], P[L["row_phi"]], P[L["row_ecc"]], P[L["row_w"]] = 7.0, 1.0, 0.0, 0.0, 0.0
    mu = port.mu(P, 0, 1)
    assert mu[0] == pytest.approx(1.0, abs=1e-15)  # t[0] is the epoch: M=0 -> v=K


@pytest.mark.parametrize("name", golden_cases())
def test_logL_mu_cov_golden(port, name):
    g = load_golden(name)
    port.set_rvdata(g["data"])
    port.set_flags(g["flags"])
    assert port.layout() == g["layout"]
    logL, _ = port.eval_batch(g["params"], g["nps"])
    tol = 2e-12 if g["flags"].GP else 0.0  # GP: summation order of the LLT differs from Eigen's
    assert np.array_equal(np.isfinite(logL), np.isfinite(g["logL"]))
    assert relerr(logL, g["logL"]).max() <= tol
    mu, lo = port.mu(g["params"], 0, int(g["nps"][0]), with_lo=True)
    assert np.array_equal(mu, g["mu0"], equal_nan=True)
    assert np.array_equal(lo, g["mu0_lo"], equal_nan=True)
    if g["C0"] is not None:
        assert np.array_equal(port.cov(g["params"], 0), g["C0"])


def test_scale_point_N4096_port_matches_reference(port):
    """BASELINE.json configs[4] (N=4096, QP-GP): the plain-C port against the reference's values
    on two lanes (the port's unblocked LLT takes ~6 s per lane here)."""
    g = load_golden("synth_gp_N4096")
    port.set_rvdata(g["data"])
    port.set_flags(g["flags"])
    logL, _ = port.eval_batch(g["params"][:, :2].copy(), g["nps"][:2], 2)
    assert relerr(logL, g["logL"][:2]).max() <= 2e-12


def test_tail_case_port_is_the_reference_bit_for_bit(port):
    """The ill-conditioned lanes (tests/golden/tail_kep_N200_np5_hyper.npz): the plain-C port follows the
    reference's acos chain operation by operation, so even here it reproduces the reference exactly."""
    import json
    import os
    from kima_b200.flags import ModelFlags
    from kima_b200.synthetic import make_data
    z = np.load(os.path.join(GOLDEN, "tail_kep_N200_np5_hyper.npz"))
    flags = ModelFlags(**{k: (bool(v) if isinstance(getattr(ModelFlags(), k), bool) else int(v))
                          for k, v in json.loads(str(z["flags"])).items()})
    port.set_rvdata(make_data(200))
    port.set_flags(flags)
    logL, _ = port.eval_batch(z["params"], z["nps"])
    assert np.array_equal(logL, z["logL"])


def test_nonfinite_maps_to_minus_1e300(port):
    # src/RVmodel.cpp:938-941 then DNest4/code/LikelihoodType.cpp:17-24
    g = load_golden("synth_edge")
    assert np.isposinf(g["logL"]).sum() == 2
    assert port.likelihood_type_value(float("inf")) == -1e300
    assert port.likelihood_type_value(float("nan")) == -1e300
    assert port.likelihood_type_value(-12.5) == -12.5


def test_cholesky_matches_numpy(port):
    rng = np.random.default_rng(0)
    for n in (1, 7, 31, 32, 33, 100, 257):
        A = rng.normal(size=(n, n))
        A = A @ A.T + n * np.eye(n)
        L, info = port.cholesky(A)
        assert info == -1
        np.testing.assert_allclose(L, np.linalg.cholesky(A), rtol=1e-11, atol=1e-12)
    A = np.eye(5)
    A[3, 3] = -1.0
    assert port.cholesky(A)[1] == 3  # Eigen's info: index of the failing pivot
