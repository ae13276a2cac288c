"""TEST INFRASTRUCTURE — ctypes front-end of oracle/_ref/libkima_ref_philox.so: the unmodified
reference sampler code (RVmodel::perturb, Sampler::mcmc_thread, the distributions, ...) driven
by the counter-based random stream of include/kc_philox.h (oracle/philox_rng.cpp,
oracle/ref_sampler_harness.cpp).  The checker of the device-side sampler.

Only tests/ may import this module.  Nothing in kima_b200/ does.
"""
from __future__ import annotations

import ctypes
import os

import numpy as np

from kima_b200.sampler_types import CPrior, CPriorTable, CSamplerLayout, PRIOR_TYPES

from .oracle import PHILOX_SO, Oracle

_dp = ctypes.POINTER(ctypes.c_double)
_ip = ctypes.POINTER(ctypes.c_int32)
_up = ctypes.POINTER(ctypes.c_uint64)


def have_philox_reference() -> bool:
    return os.path.exists(PHILOX_SO)


class RefSampler:
    """The reference's perturb / mcmc_thread on chosen states and stream positions."""

    def __init__(self):
        self.oracle = Oracle("philox")  # data, flags, likelihoods of the same shared object
        self.lib = self.oracle.lib
        self.lib.rs_rng_draws.restype = ctypes.c_uint64
        self.lib.rs_rng_draws.argtypes = [ctypes.c_uint64, ctypes.c_uint64, ctypes.c_int, ctypes.c_int, _dp]
        self.lib.rs_perturb.argtypes = [ctypes.c_int, ctypes.c_int, _dp, _ip, ctypes.c_uint64, _up, _dp]
        self.lib.rs_mcmc.argtypes = [ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_int, _dp, _ip, _dp, _dp,
                                     _ip, ctypes.c_int, _dp, _dp, _dp, _up, ctypes.c_double, ctypes.c_double,
                                     ctypes.c_int, ctypes.c_uint64, _up, _dp, ctypes.c_int, _ip]

    # ---- priors ----------------------------------------------------------------------------
    def default_priors(self) -> CPriorTable:
        """setPriors() + the RVConditionalPrior constructor for the current flags and data."""
        T = CPriorTable()
        rc = self.lib.rs_default_priors(ctypes.byref(T))
        assert rc == 0, "a default prior has no table form"
        return T

    def set_priors(self, table: CPriorTable) -> None:
        self.lib.rs_set_priors(ctypes.byref(table))

    def make_prior(self, name: str, *pars) -> CPrior:
        q = CPrior()
        p = np.asarray(pars, np.float64)
        rc = self.lib.rs_prior_make(PRIOR_TYPES[name], p.ctypes.data_as(_dp), p.size, ctypes.byref(q))
        if rc:
            raise ValueError("%s%r: the reference's constructor would throw" % (name, pars))
        return q

    def prior_eval(self, prior: CPrior, what: int, x) -> np.ndarray:
        x = np.ascontiguousarray(x, np.float64)
        out = np.empty_like(x)
        self.lib.rs_prior_eval(ctypes.byref(prior), what, x.size, x.ctypes.data_as(_dp), out.ctypes.data_as(_dp))
        return out

    def rng_draws(self, key: int, ctr0: int, what: int, n: int):
        out = np.empty(n)
        end = self.lib.rs_rng_draws(key, ctr0, what, n, out.ctypes.data_as(_dp))
        return out, int(end)

    def sampler_layout(self, n_particles: int) -> CSamplerLayout:
        SL = CSamplerLayout()
        self.lib.rs_sampler_layout(n_particles, ctypes.byref(SL))
        return SL

    # ---- the reference's code on the shared stream ----------------------------------------
    def perturb(self, state, np_, key0: int, ctr):
        """RVmodel::perturb of column b with stream (key0 + b, ctr[b]); returns
        (state', np', logH, ctr')."""
        state = np.array(state, np.float64, order="C")
        np_ = np.array(np_, np.int32)
        ctr = np.array(ctr, np.uint64)
        B = state.shape[1]
        logH = np.empty(B)
        self.lib.rs_perturb(B, B, state.ctypes.data_as(_dp), np_.ctypes.data_as(_ip), key0,
                            ctr.ctypes.data_as(_up), logH.ctypes.data_as(_dp))
        return state, np_, logH, ctr

    def mcmc(self, T, P, steps, state, np_, logl, tb, level, lv_logl, lv_tb, lv_logX, lv_counters,
             lambda_, beta, max_num_levels, key0, ctr):
        """`steps` iterations of Sampler::mcmc_thread for every thread.  Returns a dict with the
        new state arrays, the summed counter increments [L][4], the `above` pairs and the
        stream positions."""
        state = np.array(state, np.float64, order="C")
        np_ = np.array(np_, np.int32)
        logl = np.array(logl, np.float64)
        tb = np.array(tb, np.float64)
        level = np.array(level, np.int32)
        cnt = np.array(lv_counters, np.uint64).reshape(-1, 4).copy()
        L = cnt.shape[0]
        a = [np.ascontiguousarray(x, np.float64) for x in (lv_logl, lv_tb, lv_logX)]
        ctr = np.array(ctr, np.uint64)
        max_pairs = T * steps
        above = np.zeros((max(max_pairs, 1), 2))
        n_lane = np.zeros(T, np.int32)
        n = self.lib.rs_mcmc(T, P, steps, T * P, state.ctypes.data_as(_dp), np_.ctypes.data_as(_ip),
                             logl.ctypes.data_as(_dp), tb.ctypes.data_as(_dp), level.ctypes.data_as(_ip), L,
                             a[0].ctypes.data_as(_dp), a[1].ctypes.data_as(_dp), a[2].ctypes.data_as(_dp),
                             cnt.ctypes.data_as(_up), lambda_, beta, max_num_levels, key0,
                             ctr.ctypes.data_as(_up), above.ctypes.data_as(_dp), max_pairs,
                             n_lane.ctypes.data_as(_ip))
        return dict(state=state, np=np_, logl=logl, tb=tb, level=level, counters=cnt, above=above[:n].copy(),
                    n_above_lane=n_lane, ctr=ctr)
