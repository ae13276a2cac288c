"""Deterministic synthetic RV data sets and parameter batches.

Data: ``default_rng(20261104)``; parameters: seed 43, drawn from the
reference's default priors (src/RVConditionalPrior.cpp:47-59,
src/RVmodel.cpp:34-38,59-68,78-79).  The same generator feeds the GPU path,
the CPU oracle and the reference harness, so every arm scores identical
parameter vectors.
"""
from __future__ import annotations

import numpy as np

from .data import RVData
from .flags import ModelFlags, layout_from_flags

DATA_SEED = 20261104
PARAM_SEED = 43


def make_data(N: int, n_inst: int = 1, n_ind: int = 0, seed: int = DATA_SEED) -> RVData:
    rng = np.random.default_rng(seed)
    t = np.sort(rng.uniform(0.0, 2000.0, N)) + 50000.0
    sig = rng.uniform(0.5, 2.0, N)
    y = rng.normal(0.0, 5.0, N)
    obsi = None
    if n_inst > 1:
        obsi = 1 + rng.integers(0, n_inst, N)
        # every instrument must appear at least once
        obsi[:n_inst] = np.arange(1, n_inst + 1)
    act = None
    if n_ind:
        act = rng.normal(0.0, 1.0, (n_ind, N))
        act -= act.mean(axis=1, keepdims=True)
    return RVData.from_arrays(t, y, sig, obsi, act, n_inst=n_inst)


def _mlu(u, knee, upper):
    """ModifiedLogUniform cdf inverse (DNest4 Distributions/LogUniform.cpp)."""
    return knee * ((1.0 + upper / knee) ** u - 1.0)


def make_params(flags: ModelFlags, data: RVData, B: int, np_mode: str = "max",
                seed: int = PARAM_SEED):
    """params[n_rows][B] (float64, C-contiguous) and np[B] (int32).

    np_mode: "max" -> every lane has npmax planets (worst case);
             "mix" -> Np ~ U{0..npmax} (a DNest4-realistic mixture).
    """
    rng = np.random.default_rng(seed)
    L = layout_from_flags(flags, data.number_instruments, data.number_indicators)
    P = np.zeros((L["n_rows"], B))
    npm = flags.npmax
    if np_mode == "max":
        nps = np.full(B, npm, np.int32)
    elif np_mode == "mix":
        nps = rng.integers(0, npm + 1, B).astype(np.int32)
    else:
        raise ValueError(np_mode)

    def planets(rowP, rowK, rowphi, rowe, roww, n, hyper):
        if n == 0:
            return
        logP = rng.uniform(np.log(1.0), np.log(1e5), (n, B))
        P[rowP:rowP + n] = logP if hyper else np.exp(logP)
        P[rowK:rowK + n] = _mlu(rng.uniform(size=(n, B)), 1.0, 1e3)
        P[rowphi:rowphi + n] = rng.uniform(0, 2 * np.pi, (n, B))
        P[rowe:rowe + n] = rng.uniform(0, 1, (n, B))
        P[roww:roww + n] = rng.uniform(0, 2 * np.pi, (n, B))

    planets(L["row_P"], L["row_K"], L["row_phi"], L["row_ecc"], L["row_w"], npm,
            flags.hyperpriors)
    span = data.get_max_RV_span()
    P[L["row_vsys"]] = rng.uniform(data.get_RV_min(), data.get_RV_max(), B)
    P[L["row_jit"]:L["row_jit"] + L["n_jit"]] = _mlu(
        rng.uniform(size=(L["n_jit"], B)), min(1.0, 0.1 * span), span)
    if L["n_off"]:
        s = data.get_RV_span()
        P[L["row_off"]:L["row_off"] + L["n_off"]] = rng.uniform(-s, s, (L["n_off"], B))
    if flags.trend:
        for k in range(3):
            if flags.degree >= k + 1:
                P[L["row_trend"] + k] = rng.normal(
                    0.0, 10.0 ** data.get_trend_magnitude(k + 1), B)
    if L["n_beta"]:
        P[L["row_beta"]:L["row_beta"] + L["n_beta"]] = rng.normal(0, 1, (L["n_beta"], B))
    if flags.GP:
        P[L["row_eta"] + 0] = np.exp(rng.uniform(-5, 5, B))
        P[L["row_eta"] + 1] = np.exp(rng.uniform(np.log(1.0), np.log(100.0), B))
        P[L["row_eta"] + 2] = rng.uniform(10, 40, B)
        P[L["row_eta"] + 3] = np.exp(rng.uniform(-1, 1, B))
    if flags.studentt:
        P[L["row_nu"]] = np.exp(rng.uniform(np.log(2.0), np.log(1000.0), B))
    if L["n_ko"]:
        planets(L["row_KO_P"], L["row_KO_K"], L["row_KO_phi"], L["row_KO_e"],
                L["row_KO_w"], L["n_ko"], False)
    if flags.MA:
        # sigmaMA ~ U(-1, 1), tauMA ~ LogUniform(1, 100)  (src/RVmodel.cpp:31-32); drawn last so
        # that the other rows of a seeded batch do not depend on the MA flag
        P[L["row_MA"]] = rng.uniform(-1, 1, B)
        P[L["row_MA"] + 1] = np.exp(rng.uniform(np.log(1.0), np.log(100.0), B))
    return np.ascontiguousarray(P), nps
