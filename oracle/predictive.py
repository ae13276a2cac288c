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


# ---- kernel = celerite ----------------------------------------------------------------------------
def celerite_terms(eta):
    """The three complex terms of kima's quasi-periodic approximation (src/RVmodel.cpp:231-250):
    (a_comp, b_comp, c_comp, d_comp); the kernel has no eta4."""
    wbeat, wrot, amp = 1.0 / eta[1], 2.0 * np.pi / eta[2], eta[0] * eta[0]
    c, d = wbeat, wrot
    x = c / d
    a, b, e, f = amp / 2, amp * x / 2, amp / 8, amp * x / 4
    g = amp * (3.0 / 8.0 + 0.001)
    return (np.array([a, e, g]), np.array([b, f, 0.0]), np.array([c, c, c]), np.array([d, 2 * d, 0.0]))


def celerite_factor(terms, t, diag):
    """CholeskySolver::compute (celerite/cpp/include/celerite/solver/cholesky.h:95-190) for J = 6
    complex-term columns: D[N], W[6][N], u[6][N-1], phi[6][N-1]; None if a pivot goes negative."""
    a_c, b_c, c_c, d_c = terms
    N, J = t.size, 6
    D = diag + a_c.sum()
    W = np.zeros((J, N))
    u = np.zeros((J, N - 1))
    phi = np.zeros((J, N - 1))
    Dn = D[0]
    for j in range(3):
        W[2 * j, 0] = np.cos(d_c[j] * t[0]) / Dn
        W[2 * j + 1, 0] = np.sin(d_c[j] * t[0]) / Dn
    S = np.zeros((J, J))
    D = D.copy()
    for n in range(1, N):
        dx = t[n] - t[n - 1]
        for j in range(3):
            cd, sd = np.cos(d_c[j] * t[n]), np.sin(d_c[j] * t[n])
            phi[2 * j, n - 1] = phi[2 * j + 1, n - 1] = np.exp(-c_c[j] * dx)
            u[2 * j, n - 1] = a_c[j] * cd + b_c[j] * sd
            u[2 * j + 1, n - 1] = a_c[j] * sd - b_c[j] * cd
            W[2 * j, n], W[2 * j + 1, n] = cd, sd
        for j in range(J):
            xj = Dn * W[j, n - 1]
            for k in range(j + 1):
                S[k, j] = phi[j, n - 1] * (phi[k, n - 1] * (S[k, j] + xj * W[k, n - 1]))
        Dn = D[n]
        for j in range(J):
            uj, xj = u[j, n - 1], W[j, n]
            for k in range(j):
                tmp = u[k, n - 1] * S[k, j]
                Dn -= 2.0 * (uj * tmp)
                xj -= tmp
                W[k, n] -= uj * S[k, j]
            tmp = uj * S[j, j]
            Dn -= uj * tmp
            W[j, n] = xj - tmp
        if Dn < 0:
            return None
        D[n] = Dn
        W[:, n] /= Dn
    return D, W, u, phi


def celerite_solve(fac, b):
    """CholeskySolver::solve (cholesky.h:217-262): forward and backward sweep."""
    D, W, u, phi = fac
    N, J = b.size, 6
    x = b.astype(np.float64).copy()
    f = np.zeros(J)
    for n in range(1, N):
        f = phi[:, n - 1] * (f + W[:, n - 1] * x[n - 1])
        x[n] = b[n] - u[:, n - 1] @ f
    x /= D
    f = np.zeros(J)
    for n in range(N - 2, -1, -1):
        f = phi[:, n] * (f + u[:, n] * x[n + 1])
        x[n] -= W[:, n] @ f
    return x


def celerite_predict(terms, t, alpha, xs):
    """CholeskySolver::predict (cholesky.h:598-690): O(N + M) forward and backward pass over the
    sorted prediction times xs."""
    a_c, b_c, c_c, d_c = terms
    N, M = t.size, xs.size
    pred = np.zeros(M)
    Q = np.zeros(6)
    m = 0
    while m < M and xs[m] <= t[0]:
        m += 1
    for n in range(N):
        tref = t[n + 1] if n < N - 1 else t[N - 1]
        tn = t[n]
        dt = tref - tn
        for j in range(3):
            tmp = np.exp(-c_c[j] * dt)
            Q[2 * j] = (Q[2 * j] + alpha[n] * np.cos(d_c[j] * tn)) * tmp
            Q[2 * j + 1] = (Q[2 * j + 1] + alpha[n] * np.sin(d_c[j] * tn)) * tmp
        while m < M and (n == N - 1 or xs[m] <= tref):
            xm = xs[m]
            dt = xm - tref
            pm = 0.0
            for j in range(3):
                cd, sd = np.cos(d_c[j] * xm), np.sin(d_c[j] * xm)
                tmp = np.exp(-c_c[j] * dt)
                pm += (a_c[j] * cd + b_c[j] * sd) * tmp * Q[2 * j]
                pm += (a_c[j] * sd - b_c[j] * cd) * tmp * Q[2 * j + 1]
            pred[m] = pm
            m += 1
    m = M - 1
    while m >= 0 and xs[m] > t[N - 1]:
        m -= 1
    Q = np.zeros(6)
    for n in range(N - 1, -1, -1):
        tref = t[n - 1] if n > 0 else t[0]
        tn = t[n]
        dt = tn - tref
        for j in range(3):
            cd, sd = np.cos(d_c[j] * tn), np.sin(d_c[j] * tn)
            tmp = np.exp(-c_c[j] * dt)
            Q[2 * j] = (Q[2 * j] + alpha[n] * (a_c[j] * cd + b_c[j] * sd)) * tmp
            Q[2 * j + 1] = (Q[2 * j + 1] + alpha[n] * (a_c[j] * sd - b_c[j] * cd)) * tmp
        while m >= 0 and (n == 0 or xs[m] > tref):
            xm = xs[m]
            dt = tref - xm
            pm = 0.0
            for j in range(3):
                tmp = np.exp(-c_c[j] * dt)
                pm += np.cos(d_c[j] * xm) * tmp * Q[2 * j]
                pm += np.sin(d_c[j] * xm) * tmp * Q[2 * j + 1]
            pred[m] += pm
            m -= 1
    return pred


def gp_mean_celerite(orc: Oracle, flags: ModelFlags, data, P, b: int, np_b: int, t=None) -> np.ndarray:
    """GP_celerite.predict (pykima/GP.py:349-355) with the C++ likelihood's set-up of the solver
    (src/RVmodel.cpp:218-266: jitter^2 and sig^2 on the diagonal) and residual y - calculate_mu:
    K_* (K + sigma)^-1 r through celerite's O(N) recursions; NaN where a pivot goes negative."""
    L = orc.layout()
    tt = data.t if t is None else np.asarray(t, np.float64)
    eta = [P[L["row_eta"] + k, b] for k in range(3)]
    terms = celerite_terms(eta)
    fac = celerite_factor(terms, data.t, data.sig ** 2 + P[L["row_jit"], b] ** 2)
    if fac is None or not np.all(np.isfinite(fac[0])):
        return np.full(tt.size, np.nan)
    alpha = celerite_solve(fac, data.y - orc.mu(P, b, np_b))
    return celerite_predict(terms, data.t, alpha, tt)
