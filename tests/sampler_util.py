"""Shared helpers of the device-side sampler tests: consistent random
sampler states, a level table, and the host instantiation of kc_sampler_logic.cuh
(libkc_hostcheck.so) behind the same method names as kima_b200.sampler.DeviceSampler."""
import ctypes
import os

import numpy as np

from kima_b200.flags import ModelFlags, layout_from_flags
from kima_b200.sampler_types import CPrior, CPriorTable, CSamplerConfig, CSamplerLayout, PRIOR_TYPES
from kima_b200.synthetic import make_data, make_params

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
_dp = ctypes.POINTER(ctypes.c_double)
_ip = ctypes.POINTER(ctypes.c_int32)
_up = ctypes.POINTER(ctypes.c_uint64)

# flag combinations x data shapes covering every branch of RVmodel::perturb (src/RVmodel.cpp:435-820)
CASES = {
    "white": (ModelFlags(npmax=3), dict(N=40)),
    "fixed_np": (ModelFlags(npmax=2), dict(N=30), dict(fix=True)),
    "studentt_multi": (ModelFlags(npmax=2, studentt=True, multi_instrument=True), dict(N=45, n_inst=3)),
    "trend3_ind": (ModelFlags(npmax=1, trend=True, degree=3), dict(N=35, n_ind=2)),
    "hyper": (ModelFlags(npmax=4, hyperpriors=True), dict(N=40)),
    "known_object": (ModelFlags(npmax=1, known_object=True, n_known_object=2), dict(N=30)),
    "gp": (ModelFlags(GP=True, npmax=2), dict(N=25)),
    "gp_multi_trend": (ModelFlags(GP=True, npmax=1, multi_instrument=True, trend=True, degree=2), dict(N=30, n_inst=2)),
    "gp_celerite": (ModelFlags(GP=True, kernel=1, npmax=1), dict(N=30)),
    "ma": (ModelFlags(MA=True, npmax=2, trend=True, degree=1), dict(N=30)),
    "np0": (ModelFlags(npmax=0), dict(N=20)),
}


def case(name):
    c = CASES[name]
    flags, dkw = c[0], c[1]
    extra = c[2] if len(c) > 2 else {}
    return flags, make_data(**dkw), bool(extra.get("fix", False))


class HostTwin:
    """kc_sampler_logic.cuh compiled for the host; the caller supplies the likelihoods."""

    def __init__(self, flags: ModelFlags, data, priors: CPriorTable, lanes, num_particles=1, max_steps=64,
                 max_levels=64, lambda_=10.0, beta=100.0, seed=43, lane0=0):
        self.lib = ctypes.CDLL(os.path.join(ROOT, "kima_b200", "libkc_hostcheck.so"))
        self.lib.hc_smp_create.restype = ctypes.c_void_p
        self.lib.hc_rng_draws.restype = ctypes.c_uint64
        self.lib.hc_rng_draws.argtypes = [ctypes.c_uint64, ctypes.c_uint64, ctypes.c_int, ctypes.c_int, _dp]
        cf = flags.to_c()
        cfg = CSamplerConfig(lanes, num_particles, max_steps, max_levels, lambda_, beta, seed, lane0, 0)
        self.h = ctypes.c_void_p(self.lib.hc_smp_create(ctypes.byref(cf), data.number_instruments,
                                                        data.number_indicators, ctypes.byref(priors),
                                                        ctypes.byref(cfg)))
        SL = CSamplerLayout()
        self.lib.hc_smp_layout(self.h, ctypes.byref(SL))
        self.n_state, self.row_u, self.row_hyper, self.n_particles = SL.n_state, SL.row_u, SL.row_hyper, SL.n_particles
        self.lanes, self.max_steps = lanes, max_steps
        self.n_levels = 0

    def __del__(self):
        try:
            self.lib.hc_smp_destroy(self.h)
        except Exception:
            pass

    def upload(self, state, np_, logl, tb, level):
        a = [np.ascontiguousarray(state, np.float64), np.ascontiguousarray(np_, np.int32),
             np.ascontiguousarray(logl, np.float64), np.ascontiguousarray(tb, np.float64),
             np.ascontiguousarray(level, np.int32)]
        self.lib.hc_smp_upload(self.h, a[0].ctypes.data_as(_dp), a[1].ctypes.data_as(_ip), a[2].ctypes.data_as(_dp),
                               a[3].ctypes.data_as(_dp), a[4].ctypes.data_as(_ip))

    def download(self):
        n = self.n_particles
        state = np.empty((self.n_state, n))
        np_, level = np.empty(n, np.int32), np.empty(n, np.int32)
        logl, tb = np.empty(n), np.empty(n)
        self.lib.hc_smp_download(self.h, state.ctypes.data_as(_dp), np_.ctypes.data_as(_ip), logl.ctypes.data_as(_dp),
                                 tb.ctypes.data_as(_dp), level.ctypes.data_as(_ip))
        return state, np_, logl, tb, level

    def set_levels(self, logl, tb, logX, tries, enough, at_max):
        a = [np.ascontiguousarray(x, np.float64) for x in (logl, tb, logX)]
        tr = np.ascontiguousarray(tries, np.uint64)
        self.n_levels = a[0].size
        self.lib.hc_smp_set_levels(self.h, self.n_levels, a[0].ctypes.data_as(_dp), a[1].ctypes.data_as(_dp),
                                   a[2].ctypes.data_as(_dp), tr.ctypes.data_as(_up), int(enough), int(at_max))

    def propose(self):
        self.lib.hc_smp_propose(self.h)
        Q = np.empty((self.n_state, self.lanes))
        npq, flags = np.empty(self.lanes, np.int32), np.empty(self.lanes, np.int32)
        self.lib.hc_smp_get_proposals(self.h, Q.ctypes.data_as(_dp), npq.ctypes.data_as(_ip), flags.ctypes.data_as(_ip))
        return Q, npq, flags

    def accept(self, logl_q):
        a = np.ascontiguousarray(logl_q, np.float64)
        self.lib.hc_smp_accept(self.h, a.ctypes.data_as(_dp))

    def collect(self):
        cnt = np.zeros((max(self.n_levels, 1), 4), np.uint64)
        above = np.zeros((self.lanes * self.max_steps, 2))
        n_lane = np.zeros(self.lanes, np.int32)
        n = self.lib.hc_smp_collect(self.h, cnt.ctypes.data_as(_up), above.ctypes.data_as(_dp), above.shape[0],
                                    n_lane.ctypes.data_as(_ip))
        return cnt[:self.n_levels], above[:n].copy(), n_lane

    def get_counters(self):
        c = np.empty(self.lanes, np.uint64)
        self.lib.hc_smp_counters(self.h, c.ctypes.data_as(_up), 0)
        return c

    def set_counters(self, c):
        c = np.ascontiguousarray(c, np.uint64)
        self.lib.hc_smp_counters(self.h, c.ctypes.data_as(_up), 1)

    def test_perturb(self):
        state = np.empty((self.n_state, self.lanes))
        np_ = np.empty(self.lanes, np.int32)
        logH = np.empty(self.lanes)
        self.lib.hc_smp_perturb(self.h, state.ctypes.data_as(_dp), np_.ctypes.data_as(_ip), logH.ctypes.data_as(_dp))
        return state, np_, logH

    def prior_eval(self, prior: CPrior, what, x):
        x = np.ascontiguousarray(x, np.float64)
        out = np.empty_like(x)
        self.lib.hc_prior_eval(ctypes.byref(prior), what, x.size, x.ctypes.data_as(_dp), out.ctypes.data_as(_dp))
        return out

    def rng_draws(self, key, ctr0, what, n):
        out = np.empty(n)
        end = self.lib.hc_rng_draws(key, ctr0, what, n, out.ctypes.data_as(_dp))
        return out, int(end)


def host_prior_make(name, *pars) -> CPrior:
    lib = ctypes.CDLL(os.path.join(ROOT, "kima_b200", "libkc_hostcheck.so"))
    q = CPrior()
    p = np.asarray(pars, np.float64)
    rc = lib.hc_prior_make(PRIOR_TYPES[name], p.ctypes.data_as(_dp), p.size, ctypes.byref(q))
    if rc:
        raise ValueError((name, pars))
    return q


def host_prior_eval(prior: CPrior, what, x):
    lib = ctypes.CDLL(os.path.join(ROOT, "kima_b200", "libkc_hostcheck.so"))
    x = np.ascontiguousarray(x, np.float64)
    out = np.empty_like(x)
    lib.hc_prior_eval(ctypes.byref(prior), what, x.size, x.ctypes.data_as(_dp), out.ctypes.data_as(_dp))
    return out


def make_state(flags: ModelFlags, data, priors: CPriorTable, n: int, seed: int = 7):
    """A consistent random sampler state for n particles: state [n_state][n], np [n].
    Components are from_uniform(u) of uniform u's (as RJObject::add_component leaves them,
    RJObjectImpl.h:85-108), everything else a draw from the default priors."""
    L = layout_from_flags(flags, data.number_instruments, data.number_indicators)
    P, nps = make_params(flags, data, n, "mix", seed=seed)
    if priors.fix:
        nps[:] = flags.npmax
    npm = flags.npmax
    n_state = L["n_rows"] + 5 * npm + 3
    row_u, row_h = L["n_rows"], L["n_rows"] + 5 * npm
    S = np.zeros((n_state, n))
    S[:L["n_rows"]] = P
    rng = np.random.default_rng(seed + 1000)
    # hyperparameters: center ~ log_muP-ish, width ~ U(0.1, 3), muK ~ exp(U(-2, 3))
    S[row_h + 0] = rng.normal(np.log(365.0), 1.0, n)
    S[row_h + 1] = rng.uniform(0.1, 3.0, n)
    S[row_h + 2] = np.exp(rng.uniform(-2.0, 3.0, n))
    cond = [priors.Pprior, priors.Kprior, priors.phiprior, priors.eprior, priors.wprior]
    for j in range(npm):
        for d in range(5):
            u = rng.uniform(0.0, 1.0, n)
            S[row_u + d * npm + j] = u
            if flags.hyperpriors and d < 2:
                v = np.empty(n)
                for i in range(n):  # Laplace(center, width) / Exponential(muK) of the particle
                    q = (host_prior_make("Laplace", S[row_h, i], S[row_h + 1, i]) if d == 0
                         else host_prior_make("Exponential", S[row_h + 2, i]))
                    v[i] = host_prior_eval(q, 1, u[i:i + 1])[0]
            else:
                v = host_prior_eval(cond[d], 1, u)
            S[L["row_P"] + d * npm + j] = v
    # rows of absent components are zero, as model_to_column writes them
    for j in range(npm):
        dead = nps <= j
        for d in range(5):
            S[L["row_P"] + d * npm + j, dead] = 0.0
            S[row_u + d * npm + j, dead] = 0.0
    return np.ascontiguousarray(S), nps.astype(np.int32), L


def make_levels(logl, L_levels: int, seed: int = 3):
    """A plausible level table under the particles' likelihoods: level 0 at -1e300, the others
    at increasing quantiles; counters with some history."""
    rng = np.random.default_rng(seed)
    fin = np.sort(logl[np.isfinite(logl) & (logl > -1e299)])
    qs = np.linspace(0.05, 0.8, max(L_levels - 1, 1))
    lv_logl = np.concatenate([[-1e300], np.quantile(fin, qs)[:L_levels - 1]]) if L_levels > 1 else np.array([-1e300])
    lv_tb = rng.uniform(0.05, 0.95, L_levels)
    lv_tb[0] = 0.0  # Level() default: LikelihoodType(lowest, 0)
    lv_logX = -np.arange(L_levels, dtype=np.float64) * 0.9
    cnt = np.zeros((L_levels, 4), np.uint64)
    cnt[:, 1] = rng.integers(50, 500, L_levels)        # tries
    cnt[:, 0] = (cnt[:, 1] * rng.uniform(0.1, 0.6, L_levels)).astype(np.uint64)  # accepts
    cnt[:, 2] = rng.integers(50, 500, L_levels)        # visits
    cnt[:, 3] = (cnt[:, 2] * rng.uniform(0.2, 0.5, L_levels)).astype(np.uint64)  # exceeds
    return lv_logl, lv_tb, lv_logX, cnt
