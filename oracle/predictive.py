"""TEST INFRASTRUCTURE — CPU restatement of the posterior-predictive curves.

Restates, for one parameter vector at a time,
  KimaResults.model()          /root/reference/pykima/results.py:723-797
  GP.predict(y - model, t)     /root/reference/pykima/GP.py:96-131 (mean, and the standard
                               deviation of return_std=True), with the kernel of pykima/GP.py:36-55
on top of the reference's C++ arithmetic as the oracle exposes it: the Keplerians use
kepler::true_anomaly (src/RVmodel.cpp:1457-1475) through ``Oracle.true_anomaly``, the
covariance is calculate_C's (``Oracle.cov``: sig^2 + jitter^2 on the diagonal,
src/RVmodel.cpp:186-216) and the residual is y - calculate_mu (``Oracle.mu``), i.e. what
log_likelihood factorises (:841-873).  pykima's own ``GP`` object adds ``yerr`` and the jitter
unsquared to the diagonal (results.py:307-308, GP.py:75-78); that is a plotting
approximation of its own and is not what is restated here — tests/golden/
make_golden_predictive.py calls pykima's classes with the C++ variances instead.

Parity pinned: against pykima's keplerian() and GP.predict() run in the build container
(tests/golden/predictive_*.npz).  Only tests/ and bench.py's cpu_baseline leg import this.
"""
from __future__ import annotations

import numpy as np
from scipy.linalg import cho_factor, cho_solve, cholesky, solve_triangular

from kima_b200.flags import ModelFlags
from .oracle import Oracle


def _planets(orc: Oracle, flags: ModelFlags, P, b: int, np_b: int):
    L = orc.layout()
    out = []
    for j in range(L["n_ko"]):
        out.append(tuple(P[L[r] + j, b] for r in ("row_KO_P", "row_KO_K", "row_KO_phi", "row_KO_e", "row_KO_w")))
    for j in range(int(np_b)):
        per = P[L["row_P"] + j, b]
        if flags.hyperpriors:
            per = np.exp(per)  # src/RVmodel.cpp:357-358
        out.append((per, P[L["row_K"] + j, b], P[L["row_phi"] + j, b], P[L["row_ecc"] + j, b],
                    P[L["row_w"] + j, b]))
    return out


def model(orc: Oracle, flags: ModelFlags, data, P, b: int, np_b: int, t=None) -> np.ndarray:
    """results.py:723-797.  t=None: the data epochs, offsets and indicator terms included
    (that is calculate_mu); a grid: Keplerians + vsys + trend only."""
    if t is None:
        return orc.mu(P, b, np_b)
    t = np.asarray(t, np.float64)
    L = orc.layout()
    v = np.zeros_like(t)
    for per, K, phi, ecc, w in _planets(orc, flags, P, b, np_b):
        Tp = data.M0_epoch - (per * phi) / (2.0 * np.pi)  # src/RVmodel.cpp:369
        f = orc.true_anomaly(t, np.full_like(t, per), np.full_like(t, ecc), np.full_like(t, Tp))
        v += K * (np.cos(f + w) + ecc * np.cos(w))      # :371
    v += P[L["row_vsys"], b]
    if flags.trend:
        tm = t - data.get_t_middle()
        s = [P[L["row_trend"] + k, b] for k in range(3)]
        v += s[0] * tm + s[1] * tm ** 2 + s[2] * tm ** 3
    return v


def qp_kernel(eta, x1, x2) -> np.ndarray:
    """pykima/GP.py:46-55 (cdist branch)."""
    d = np.abs(np.asarray(x1)[:, None] - np.asarray(x2)[None, :])
    return eta[0] ** 2 * np.exp(-0.5 * (d / eta[1]) ** 2 - 2.0 * (np.sin(np.pi * d / eta[2]) / eta[3]) ** 2)


def gp_mean(orc: Oracle, flags: ModelFlags, data, P, b: int, np_b: int, t=None) -> np.ndarray:
    """GP.py:96-112: K(t*, t) C^-1 (y - mu); NaN where C is not positive definite."""
    L = orc.layout()
    tt = data.t if t is None else np.asarray(t, np.float64)
    C = orc.cov(P, b)
    r = data.y - orc.mu(P, b, np_b)
    try:
        alpha = cho_solve(cho_factor(C, lower=True), r)
    except np.linalg.LinAlgError:
        return np.full(tt.size, np.nan)
    eta = [P[L["row_eta"] + k, b] for k in range(4)]
    return qp_kernel(eta, tt, data.t) @ alpha


def gp_std(orc: Oracle, flags: ModelFlags, data, P, b: int, np_b: int, t=None) -> np.ndarray:
    """GP.py:113-131 (return_std=True): sqrt(k(t*,t*) - k(t*,t) C^-1 k(t,t*)), negative variances
    set to zero; NaN where C is not positive definite.  pykima forms C^-1 explicitly; here
    |L^-1 k(t,t*)|^2 by forward substitution (the same number, better conditioned)."""
    L = orc.layout()
    tt = data.t if t is None else np.asarray(t, np.float64)
    C = orc.cov(P, b)
    try:
        Lc = cholesky(C, lower=True)
    except np.linalg.LinAlgError:
        return np.full(tt.size, np.nan)
    eta = [P[L["row_eta"] + k, b] for k in range(4)]
    v = solve_triangular(Lc, qp_kernel(eta, data.t, tt), lower=True)
    var = eta[0] ** 2 - np.einsum("ij,ij->j", v, v)
    return np.sqrt(np.maximum(var, 0.0))
