"""Device-side proposals and accept logic: host mirror of include/kima_sampler.h
(SURVEY.md section 8, row f1).

``DeviceSampler`` keeps the particles of all lanes on the GPU and runs blocks of lock-step
MCMC steps there (the reference's ``Sampler::mcmc_thread``, DNest4/code/SamplerImpl.h:134-165,
with ``RVmodel::perturb``, src/RVmodel.cpp:435-820, inside); the caller does DNest4's
bookkeeping between blocks.  Names follow the reference.
"""
from __future__ import annotations

import ctypes
from dataclasses import dataclass
from typing import Optional

import numpy as np

from . import _lib
from .engine import LikelihoodEngine

from .sampler_types import (CPrior, CPriorTable, CSamplerConfig, CSamplerLayout, KC_MAX_KO, KC_PRIOR_NONE,  # noqa: F401
                            PRIOR_TYPES, _d, _i, _u)


@dataclass
class SamplerOptions:
    """The fields of DNest4's Options (DNest4/code/Options.h) the MCMC steps read."""
    lanes: int
    num_particles: int = 1
    thread_steps: int = 100
    max_levels: int = 256
    lambda_: float = 10.0
    beta: float = 100.0
    seed: int = 43
    lane0: int = 0
    pool_reserve: int = 0

    def to_c(self) -> CSamplerConfig:
        return CSamplerConfig(self.lanes, self.num_particles, self.thread_steps, self.max_levels,
                              self.lambda_, self.beta, self.seed, self.lane0, self.pool_reserve)


class DeviceSampler:
    """Particles, level counters and `above` lists of all lanes in HBM; steps run on the GPU."""

    def __init__(self, engine: LikelihoodEngine, priors: CPriorTable, opt: SamplerOptions):
        self.engine = engine
        self.lib = engine.lib
        self.ctx = engine._ctx
        self.opt = opt
        cfg = opt.to_c()
        _lib.check(self.lib.kc_sampler_create(self.ctx, ctypes.byref(priors), ctypes.byref(cfg)))
        SL = CSamplerLayout()
        _lib.check(self.lib.kc_sampler_get_layout(self.ctx, ctypes.byref(SL)))
        self.n_state, self.row_u, self.row_hyper, self.n_particles = SL.n_state, SL.row_u, SL.row_hyper, SL.n_particles
        self.n_levels = 0

    def close(self) -> None:
        if self.ctx is not None and self.ctx.value:
            self.lib.kc_sampler_destroy(self.ctx)
        self.ctx = None

    def upload(self, state=None, np_=None, logl=None, tiebreaker=None, level=None) -> None:
        a = [None if x is None else np.ascontiguousarray(x, t) for x, t in
             ((state, np.float64), (np_, np.int32), (logl, np.float64), (tiebreaker, np.float64), (level, np.int32))]
        if a[0] is not None:
            assert a[0].shape == (self.n_state, self.n_particles), a[0].shape
        _lib.check(self.lib.kc_sampler_upload(self.ctx, _d(a[0]), _i(a[1]), _d(a[2]), _d(a[3]), _i(a[4])))

    def download(self):
        n = self.n_particles
        state = np.empty((self.n_state, n))
        np_, level = np.empty(n, np.int32), np.empty(n, np.int32)
        logl, tb = np.empty(n), np.empty(n)
        _lib.check(self.lib.kc_sampler_download(self.ctx, _d(state), _i(np_), _d(logl), _d(tb), _i(level)))
        return state, np_, logl, tb, level

    def get_particle(self, which: int):
        col = np.empty(self.n_state)
        n = ctypes.c_int32()
        _lib.check(self.lib.kc_sampler_get_particle(self.ctx, which, _d(col), ctypes.byref(n)))
        return col, n.value

    def set_particle(self, which: int, col, np_: int, logl: float, tiebreaker: float, level: int) -> None:
        col = np.ascontiguousarray(col, np.float64)
        assert col.size == self.n_state
        _lib.check(self.lib.kc_sampler_set_particle(self.ctx, which, _d(col), np_, logl, tiebreaker, level))

    def copy_particles(self, dst, src) -> None:
        dst = np.ascontiguousarray(dst, np.int32)
        src = np.ascontiguousarray(src, np.int32)
        _lib.check(self.lib.kc_sampler_copy_particles(self.ctx, dst.size, _i(dst), _i(src)))

    def set_levels(self, logl, tiebreaker, log_X, tries, enough: bool, at_max: bool) -> None:
        logl = np.ascontiguousarray(logl, np.float64)
        tb = np.ascontiguousarray(tiebreaker, np.float64)
        lx = np.ascontiguousarray(log_X, np.float64)
        tr = np.ascontiguousarray(tries, np.uint64)
        self.n_levels = logl.size
        _lib.check(self.lib.kc_sampler_set_levels(self.ctx, logl.size, _d(logl), _d(tb), _d(lx), _u(tr),
                                                  int(enough), int(at_max)))

    def run(self, steps: int) -> None:
        _lib.check(self.lib.kc_sampler_run(self.ctx, steps))

    def collect(self, max_pairs: Optional[int] = None):
        """(counters[L][4] = accepts, tries, visits, exceeds; above[n][2])"""
        if max_pairs is None:
            max_pairs = self.opt.lanes * self.opt.thread_steps
        cnt = np.zeros((max(self.n_levels, 1), 4), np.uint64)
        above = np.empty((max(max_pairs, 1), 2))
        n = ctypes.c_int32()
        _lib.check(self.lib.kc_sampler_collect(self.ctx, _u(cnt), _d(above), max_pairs, ctypes.byref(n)))
        return cnt[:self.n_levels], above[:min(n.value, max_pairs)].copy()

    # ---- `all_above` kept in HBM ----------------------------------------------------------
    def collect_pool(self):
        """(counters[L][4], pairs appended, pool size)"""
        cnt = np.zeros((max(self.n_levels, 1), 4), np.uint64)
        a, b = ctypes.c_int64(), ctypes.c_int64()
        _lib.check(self.lib.kc_sampler_collect_pool(self.ctx, _u(cnt), ctypes.byref(a), ctypes.byref(b)))
        return cnt[:self.n_levels], a.value, b.value

    def pool_select(self, index: int):
        """(value, tiebreaker) at sorted position `index`; the pool keeps what lies above it."""
        v, t, n = ctypes.c_double(), ctypes.c_double(), ctypes.c_int64()
        _lib.check(self.lib.kc_sampler_pool_select(self.ctx, index, ctypes.byref(v), ctypes.byref(t), ctypes.byref(n)))
        return v.value, t.value, n.value

    def pool_clear(self) -> None:
        _lib.check(self.lib.kc_sampler_pool_clear(self.ctx))

    def pool_fetch(self, max_pairs: int) -> np.ndarray:
        out = np.empty((max(max_pairs, 1), 2))
        n = ctypes.c_int64()
        _lib.check(self.lib.kc_sampler_pool_fetch(self.ctx, _d(out), max_pairs, ctypes.byref(n)))
        return out[:n.value].copy()

    def get_counters(self) -> np.ndarray:
        c = np.empty(self.opt.lanes, np.uint64)
        _lib.check(self.lib.kc_sampler_get_counters(self.ctx, _u(c)))
        return c

    def set_counters(self, c) -> None:
        c = np.ascontiguousarray(c, np.uint64)
        assert c.size == self.opt.lanes
        _lib.check(self.lib.kc_sampler_set_counters(self.ctx, _u(c)))

    # ---- test hooks ------------------------------------------------------------------------
    def test_perturb(self):
        T = self.opt.lanes
        state = np.empty((self.n_state, T))
        np_ = np.empty(T, np.int32)
        logH = np.empty(T)
        _lib.check(self.lib.kc_sampler_test_perturb(self.ctx, _d(state), _i(np_), _d(logH)))
        return state, np_, logH

    def prior_eval(self, prior: CPrior, what: int, x) -> np.ndarray:
        x = np.ascontiguousarray(x, np.float64)
        out = np.empty_like(x)
        _lib.check(self.lib.kc_prior_eval(self.ctx, ctypes.byref(prior), what, x.size, _d(x), _d(out)))
        return out

    def rng_draws(self, key: int, ctr0: int, what: int, n: int):
        out = np.empty(n)
        end = ctypes.c_uint64()
        _lib.check(self.lib.kc_rng_draws(self.ctx, key, ctr0, what, n, _d(out), ctypes.byref(end)))
        return out, end.value

    def last_run_ms(self) -> float:
        ms = ctypes.c_float()
        _lib.check(self.lib.kc_sampler_last_run_ms(self.ctx, ctypes.byref(ms)))
        return ms.value


def make_prior(name: str, *pars) -> CPrior:
    """``make_prior<T>(args...)`` (src/RVmodel.h:165-166) as data: kc_prior_make."""
    q = CPrior()
    p = np.asarray(pars, np.float64)
    _lib.check(_lib.load_library().kc_prior_make(PRIOR_TYPES[name], _d(p), p.size, ctypes.byref(q)))
    return q


def default_prior_table(flags, data, fix: bool = False) -> CPriorTable:
    """The priors an RVmodel has when the user's constructor sets none: RVmodel::setPriors
    (src/RVmodel.cpp:23-85) and the RVConditionalPrior constructor
    (src/RVConditionalPrior.cpp:10-54).  Known-object priors have no defaults there (setPriors
    throws): the caller fills KO_*prior in."""
    T = CPriorTable()
    for name, _ in CPriorTable._fields_:
        f = getattr(T, name)
        if isinstance(f, CPrior):
            f.type = KC_PRIOR_NONE
        elif hasattr(f, "__len__"):
            for q in f:
                q.type = KC_PRIOR_NONE
    T.betaprior = make_prior("Gaussian", 0.0, 1.0)
    T.sigmaMA_prior = make_prior("Uniform", -1.0, 1.0)
    T.tauMA_prior = make_prior("LogUniform", 1.0, 100.0)
    T.Cprior = make_prior("Uniform", data.get_RV_min(), data.get_RV_max())
    span = data.get_max_RV_span()
    T.Jprior = make_prior("ModifiedLogUniform", min(1.0, 0.1 * span), span)
    if flags.trend:
        if flags.degree >= 1:
            T.slope_prior = make_prior("Gaussian", 0.0, 10.0 ** data.get_trend_magnitude(1))
        if flags.degree >= 2:
            T.quadr_prior = make_prior("Gaussian", 0.0, 10.0 ** data.get_trend_magnitude(2))
        if flags.degree == 3:
            T.cubic_prior = make_prior("Gaussian", 0.0, 10.0 ** data.get_trend_magnitude(3))
    T.offsets_prior = make_prior("Uniform", -data.get_RV_span(), data.get_RV_span())
    if flags.GP:
        T.log_eta1_prior = make_prior("Uniform", -5.0, 5.0)
        T.eta2_prior = make_prior("LogUniform", 1.0, 100.0)
        T.eta3_prior = make_prior("Uniform", 10.0, 40.0)
        if flags.kernel == 0:
            T.log_eta4_prior = make_prior("Uniform", -1.0, 1.0)
    if flags.studentt:
        T.nu_prior = make_prior("LogUniform", 2.0, 1000.0)
    if flags.hyperpriors:
        T.log_muP_prior = make_prior("TruncatedCauchy", np.log(365), 1.0, np.log(365) - 21, np.log(365) + 21)
        T.wP_prior = make_prior("Uniform", 0.1, 3.0)
        T.log_muK_prior = make_prior("TruncatedCauchy", 0.0, 1.0, 0.0 - 21, 0.0 + 21)
        T.Pprior.type = PRIOR_TYPES["Laplace"]
        T.Kprior.type = PRIOR_TYPES["Exponential"]
    else:
        T.Pprior = make_prior("LogUniform", 1.0, 1e5)
        T.Kprior = make_prior("ModifiedLogUniform", 1.0, 1e3)
    T.eprior = make_prior("Uniform", 0.0, 1.0)
    T.phiprior = make_prior("Uniform", 0.0, 2 * np.pi)
    T.wprior = make_prior("Uniform", 0.0, 2 * np.pi)
    T.fix = int(fix)
    return T


def draw_from_prior(sampler: DeviceSampler, flags, priors: CPriorTable, n: int, seed: int = 43):
    """n particle states drawn from the prior, as RVmodel::from_prior leaves them
    (src/RVmodel.cpp:88-160, RJObject::from_prior RJObjectImpl.h:19-42): every value is
    ``prior.cdf_inverse(u)`` of a uniform u (ContinuousDistribution::generate), evaluated by the
    library; the uniforms here are numpy's (the reference draws them from thread 0's RNG).
    Returns (state [n_state][n], np [n])."""
    L = sampler.engine.layout
    rng = np.random.default_rng(seed)
    S = np.zeros((sampler.n_state, n))
    npm = flags.npmax

    def gen(prior, rows=1):
        u = rng.uniform(0.0, 1.0, (rows, n))
        return u, sampler.prior_eval(prior, 1, u.ravel()).reshape(rows, n)

    rh = sampler.row_hyper
    if flags.hyperpriors:
        S[rh + 0] = gen(priors.log_muP_prior)[1][0]
        S[rh + 1] = gen(priors.wP_prior)[1][0]
        S[rh + 2] = np.exp(gen(priors.log_muK_prior)[1][0])
    nps = np.full(n, npm, np.int32) if priors.fix else rng.integers(0, npm + 1, n).astype(np.int32)
    cond = (priors.Pprior, priors.Kprior, priors.phiprior, priors.eprior, priors.wprior)
    for j in range(npm):
        alive = nps > j
        for d in range(5):
            u = rng.uniform(0.0, 1.0, n)
            if flags.hyperpriors and d == 0:      # Laplace(center, width)
                c, w = S[rh + 0], S[rh + 1]
                v = c - w * np.sign(u - 0.5) * np.log(1.0 - 2.0 * np.abs(u - 0.5))
            elif flags.hyperpriors and d == 1:    # Exponential(muK)
                v = -S[rh + 2] * np.log(1.0 - u)
            else:
                v = sampler.prior_eval(cond[d], 1, u)
            S[sampler.row_u + d * npm + j] = np.where(alive, u, 0.0)
            S[L["row_P"] + d * npm + j] = np.where(alive, v, 0.0)
    S[L["row_vsys"]] = gen(priors.Cprior)[1][0]
    S[L["row_jit"]:L["row_jit"] + L["n_jit"]] = gen(priors.Jprior, L["n_jit"])[1]
    if L["n_off"]:
        S[L["row_off"]:L["row_off"] + L["n_off"]] = gen(priors.offsets_prior, L["n_off"])[1]
    if flags.trend:
        for k, pr in enumerate((priors.slope_prior, priors.quadr_prior, priors.cubic_prior)):
            if flags.degree >= k + 1:
                S[L["row_trend"] + k] = gen(pr)[1][0]
    if L["n_beta"]:
        S[L["row_beta"]:L["row_beta"] + L["n_beta"]] = gen(priors.betaprior, L["n_beta"])[1]
    if flags.GP:
        S[L["row_eta"] + 0] = np.exp(gen(priors.log_eta1_prior)[1][0])
        S[L["row_eta"] + 1] = gen(priors.eta2_prior)[1][0]
        S[L["row_eta"] + 2] = gen(priors.eta3_prior)[1][0]
        if flags.kernel == 0:
            S[L["row_eta"] + 3] = np.exp(gen(priors.log_eta4_prior)[1][0])
    if flags.MA:
        S[L["row_MA"] + 0] = gen(priors.sigmaMA_prior)[1][0]
        S[L["row_MA"] + 1] = gen(priors.tauMA_prior)[1][0]
    if flags.studentt:
        S[L["row_nu"]] = gen(priors.nu_prior)[1][0]
    for i in range(L["n_ko"]):
        S[L["row_KO_P"] + i] = gen(priors.KO_Pprior[i])[1][0]
        S[L["row_KO_K"] + i] = gen(priors.KO_Kprior[i])[1][0]
        S[L["row_KO_e"] + i] = gen(priors.KO_eprior[i])[1][0]
        S[L["row_KO_phi"] + i] = gen(priors.KO_phiprior[i])[1][0]
        S[L["row_KO_w"] + i] = gen(priors.KO_wprior[i])[1][0]
    return np.ascontiguousarray(S), nps
