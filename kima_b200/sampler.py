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

    def to_c(self) -> CSamplerConfig:
        return CSamplerConfig(self.lanes, self.num_particles, self.thread_steps, self.max_levels,
                              self.lambda_, self.beta, self.seed, self.lane0, 0)


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
