"""Posterior samples → batched predictive curves: the host-side mirror of the part of
pykima's ``KimaResults`` that re-evaluates the model per posterior sample
(/root/reference/pykima/results.py:723-855: ``model``, ``model_stochastic``, ``full_model``,
``residuals``), scoring all samples in one ``kc_predict`` call instead of a Python loop.

A row of ``sample.txt`` / ``posterior_sample.txt`` is what ``RVmodel::print`` wrote
(/root/reference/src/RVmodel.cpp:946-1009, ``RJObject::print``
DNest4/code/RJObject/RJObjectImpl.h:241-257):

    jitters[n_inst] | extra_sigma, slope [quadr [cubic]], offsets[n_inst-1], betas[n_ind],
    eta1..eta4 (3 for celerite), sigmaMA tauMA, KO_P KO_K KO_phi KO_e KO_w (n_ko each),
    ndim npmax [center width muK] np, P[npmax] K[npmax] phi[npmax] e[npmax] w[npmax],
    staleness, [nu], vsys
"""
from __future__ import annotations

from typing import Optional

import numpy as np

from .data import RVData
from .engine import LikelihoodEngine
from .flags import ModelFlags, layout_from_flags


def sample_columns(flags: ModelFlags, n_inst: int, n_ind: int) -> dict:
    """Column index of every field of a sample row (the ``indices`` pykima builds at
    results.py:232-330), from the flags alone."""
    c = 0
    idx = {}

    def take(name, n):
        nonlocal c
        idx[name] = slice(c, c + n)
        c += n

    take("jitter", n_inst if flags.multi_instrument else 1)
    take("trend", flags.degree if flags.trend else 0)
    take("inst_offsets", n_inst - 1 if flags.multi_instrument else 0)
    take("betas", n_ind)
    take("GPpars", (4 if flags.kernel == 0 else 3) if flags.GP else 0)
    take("MA", 2 if flags.MA else 0)
    n_ko = flags.n_known_object if flags.known_object else 0
    for name in ("KO_P", "KO_K", "KO_phi", "KO_e", "KO_w"):
        take(name, n_ko)
    take("ndim", 1)
    take("npmax", 1)
    take("hyper", 3 if flags.hyperpriors else 0)
    take("np", 1)
    for name in ("P", "K", "phi", "ecc", "w"):
        take(name, flags.npmax)
    take("staleness", 1)
    take("nu", 1 if flags.studentt else 0)
    take("vsys", 1)
    idx["n_columns"] = c
    return idx


def params_from_samples(samples: np.ndarray, flags: ModelFlags, n_inst: int, n_ind: int):
    """Rows of sample.txt → the SoA block ``params[n_rows][B]``, ``np[B]`` the library takes."""
    S = np.atleast_2d(np.asarray(samples, np.float64))
    idx = sample_columns(flags, n_inst, n_ind)
    if S.shape[1] != idx["n_columns"]:
        # the message of results.py:741-745
        raise ValueError("`sample` has wrong dimensions, should be %d got %d" % (idx["n_columns"], S.shape[1]))
    L = layout_from_flags(flags, n_inst, n_ind)
    B = S.shape[0]
    P = np.zeros((L["n_rows"], B))

    def put(row, name, n):
        if n:
            P[row:row + n, :] = S[:, idx[name]].T

    put(L["row_P"], "P", flags.npmax)  # log P under hyperpriors, as the sampler holds it
    put(L["row_K"], "K", flags.npmax)
    put(L["row_phi"], "phi", flags.npmax)
    put(L["row_ecc"], "ecc", flags.npmax)
    put(L["row_w"], "w", flags.npmax)
    put(L["row_vsys"], "vsys", 1)
    put(L["row_jit"], "jitter", L["n_jit"])
    put(L["row_off"], "inst_offsets", L["n_off"])
    put(L["row_trend"], "trend", flags.degree if flags.trend else 0)
    put(L["row_beta"], "betas", L["n_beta"])
    if flags.GP:
        put(L["row_eta"], "GPpars", 4 if flags.kernel == 0 else 3)
    if flags.MA:
        put(L["row_MA"], "MA", 2)
    if flags.studentt:
        put(L["row_nu"], "nu", 1)
    if L["n_ko"]:
        for row, name in (("row_KO_P", "KO_P"), ("row_KO_K", "KO_K"), ("row_KO_phi", "KO_phi"),
                          ("row_KO_e", "KO_e"), ("row_KO_w", "KO_w")):
            put(L[row], name, L["n_ko"])
    nps = S[:, idx["np"]].ravel().astype(np.int32)
    return P, nps


class PosteriorPredictive:
    """``model`` / ``model_stochastic`` / ``full_model`` / ``residuals`` of pykima's
    KimaResults for a whole array of posterior samples at once."""

    def __init__(self, data: RVData, flags: ModelFlags, max_batch: int = 4096, device: int = 0):
        self.data = data
        self.flags = flags
        self.max_batch = int(max_batch)
        self.engine = LikelihoodEngine(data, flags, self.max_batch, device=device)

    def close(self) -> None:
        self.engine.close()

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def _run(self, samples, t, model, gp):
        P, nps = params_from_samples(samples, self.flags, self.data.number_instruments,
                                     self.data.number_indicators)
        B = P.shape[1]
        nt = self.data.N() if t is None else np.asarray(t).size
        m = np.empty((B, nt)) if model else None
        g = np.empty((B, nt)) if gp else None
        for lo in range(0, B, self.max_batch):
            hi = min(B, lo + self.max_batch)
            mm, gg = self.engine.predict(P[:, lo:hi], nps[lo:hi], t, model=model, gp=gp)
            if model:
                m[lo:hi] = mm
            if gp:
                g[lo:hi] = gg
        return m, g

    def model(self, samples, t: Optional[np.ndarray] = None) -> np.ndarray:
        """Deterministic part (results.py:723-797), one row per sample.

        Deviation from pykima at the data epochs (``t=None``): this is the C++ likelihood's
        ``calculate_mu`` without the MA term, so it also holds the indicator (beta) terms and the
        known-object Keplerians, and the instrument offsets whenever ``multi_instrument`` is set;
        pykima's ``model()`` adds neither betas nor known objects and adds offsets only for
        multi-file data.  On a grid ``t`` the two agree except for known objects (included here).
        ``residuals`` inherits the difference: it is y minus what the likelihood subtracts."""
        return self._run(samples, t, True, False)[0]

    def model_stochastic(self, samples, t: Optional[np.ndarray] = None) -> np.ndarray:
        """GP predictive mean (results.py:799-830); zeros without a GP."""
        if not self.flags.GP:
            S = np.atleast_2d(samples)
            return np.zeros((S.shape[0], self.data.N() if t is None else np.asarray(t).size))
        return self._run(samples, t, False, True)[1]

    def model_stochastic_std(self, samples, t: Optional[np.ndarray] = None):
        """GP predictive mean and standard deviation, ``GP.predict_with_hyperpars(..., return_std=True)``
        (pykima/GP.py:134-175, :113-131), one row per sample: ``(mean, std)``."""
        if not self.flags.GP:
            raise ValueError("model_stochastic_std needs a GP model")
        P, nps = params_from_samples(samples, self.flags, self.data.number_instruments,
                                     self.data.number_indicators)
        B = P.shape[1]
        nt = self.data.N() if t is None else np.asarray(t).size
        g, sd = np.empty((B, nt)), np.empty((B, nt))
        for lo in range(0, B, self.max_batch):
            hi = min(B, lo + self.max_batch)
            g[lo:hi], sd[lo:hi] = self.engine.predict_std(P[:, lo:hi], nps[lo:hi], t)
        return g, sd

    def full_model(self, samples, t: Optional[np.ndarray] = None) -> np.ndarray:
        """Deterministic part + GP predictive mean (results.py:832-855)."""
        m, g = self._run(samples, t, True, bool(self.flags.GP))
        return m if g is None else m + g

    def residuals(self, samples, full: bool = False) -> np.ndarray:
        """y - model (results.py:857-861)."""
        v = self.full_model(samples) if full else self.model(samples)
        return self.data.y[None, :] - v
