"""ctypes mirrors of the structs of include/kima_sampler.h and the signatures of its entry points."""
from __future__ import annotations

import ctypes

KC_MAX_KO = 8
PRIOR_TYPES = {
    "Fixed": 0, "Uniform": 1, "LogUniform": 2, "ModifiedLogUniform": 3, "Gaussian": 4,
    "TruncatedGaussian": 5, "Cauchy": 6, "TruncatedCauchy": 7, "Exponential": 8,
    "TruncatedExponential": 9, "Laplace": 10, "Kumaraswamy": 11, "Rayleigh": 12,
    "TruncatedRayleigh": 13, "Triangular": 14, "Pareto": 15, "mixGaussianLogUniform": 16,
}
KC_PRIOR_NONE = -1

_dp = ctypes.POINTER(ctypes.c_double)
_ip = ctypes.POINTER(ctypes.c_int32)
_up = ctypes.POINTER(ctypes.c_uint64)


def _d(a):
    return None if a is None else a.ctypes.data_as(_dp)


def _i(a):
    return None if a is None else a.ctypes.data_as(_ip)


def _u(a):
    return None if a is None else a.ctypes.data_as(_up)


class CPrior(ctypes.Structure):
    """``kc_prior``"""
    _fields_ = [("type", ctypes.c_int32), ("reserved", ctypes.c_int32),
                ("p", ctypes.c_double * 4), ("k", ctypes.c_double * 3)]


_SINGLE = ("Cprior", "Jprior", "slope_prior", "quadr_prior", "cubic_prior", "offsets_prior", "betaprior",
           "sigmaMA_prior", "tauMA_prior", "log_eta1_prior", "eta2_prior", "eta3_prior", "log_eta4_prior",
           "nu_prior")
_KO = ("KO_Pprior", "KO_Kprior", "KO_eprior", "KO_phiprior", "KO_wprior")
_COND = ("Pprior", "Kprior", "eprior", "phiprior", "wprior", "log_muP_prior", "wP_prior", "log_muK_prior")


class CPriorTable(ctypes.Structure):
    """``kc_prior_table``"""
    _fields_ = ([(n, CPrior) for n in _SINGLE] + [(n, CPrior * KC_MAX_KO) for n in _KO] +
                [(n, CPrior) for n in _COND] + [("fix", ctypes.c_int32), ("reserved", ctypes.c_int32)])


class CSamplerConfig(ctypes.Structure):
    """``kc_sampler_config``"""
    _fields_ = [("lanes", ctypes.c_int32), ("num_particles", ctypes.c_int32), ("max_steps", ctypes.c_int32),
                ("max_levels", ctypes.c_int32), ("lambda_", ctypes.c_double), ("beta", ctypes.c_double),
                ("seed", ctypes.c_uint64), ("lane0", ctypes.c_int32), ("pool_reserve", ctypes.c_int32)]


class CSamplerLayout(ctypes.Structure):
    """``kc_sampler_layout``"""
    _fields_ = [(n, ctypes.c_int32) for n in ("n_state", "row_u", "row_hyper", "n_particles")]


SAMPLER_SIGNATURES = {
    "kc_prior_make": (ctypes.c_int, [ctypes.c_int, _dp, ctypes.c_int, ctypes.POINTER(CPrior)]),
    "kc_sampler_create": (ctypes.c_int, [ctypes.c_void_p, ctypes.POINTER(CPriorTable),
                                         ctypes.POINTER(CSamplerConfig)]),
    "kc_sampler_destroy": (ctypes.c_int, [ctypes.c_void_p]),
    "kc_sampler_get_layout": (ctypes.c_int, [ctypes.c_void_p, ctypes.POINTER(CSamplerLayout)]),
    "kc_sampler_upload": (ctypes.c_int, [ctypes.c_void_p, _dp, _ip, _dp, _dp, _ip]),
    "kc_sampler_download": (ctypes.c_int, [ctypes.c_void_p, _dp, _ip, _dp, _dp, _ip]),
    "kc_sampler_get_particle": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_int, _dp, _ip]),
    "kc_sampler_set_particle": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_int, _dp, ctypes.c_int32, ctypes.c_double,
                                              ctypes.c_double, ctypes.c_int32]),
    "kc_sampler_copy_particles": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_int, _ip, _ip]),
    "kc_sampler_set_levels": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_int, _dp, _dp, _dp, _up,
                                             ctypes.c_int, ctypes.c_int]),
    "kc_sampler_run": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_int]),
    "kc_sampler_collect": (ctypes.c_int, [ctypes.c_void_p, _up, _dp, ctypes.c_int, _ip]),
    "kc_sampler_collect_pool": (ctypes.c_int, [ctypes.c_void_p, _up, ctypes.POINTER(ctypes.c_int64),
                                              ctypes.POINTER(ctypes.c_int64)]),
    "kc_sampler_pool_select": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_int64, _dp, _dp, ctypes.POINTER(ctypes.c_int64)]),
    "kc_sampler_pool_clear": (ctypes.c_int, [ctypes.c_void_p]),
    "kc_sampler_pool_fetch": (ctypes.c_int, [ctypes.c_void_p, _dp, ctypes.c_int64, ctypes.POINTER(ctypes.c_int64)]),
    "kc_sampler_last_run_ms": (ctypes.c_int, [ctypes.c_void_p, ctypes.POINTER(ctypes.c_float)]),
    "kc_sampler_get_counters": (ctypes.c_int, [ctypes.c_void_p, _up]),
    "kc_sampler_set_counters": (ctypes.c_int, [ctypes.c_void_p, _up]),
    "kc_prior_eval": (ctypes.c_int, [ctypes.c_void_p, ctypes.POINTER(CPrior), ctypes.c_int, ctypes.c_int, _dp, _dp]),
    "kc_rng_draws": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_uint64, ctypes.c_uint64, ctypes.c_int,
                                    ctypes.c_int, _dp, _up]),
    "kc_sampler_test_perturb": (ctypes.c_int, [ctypes.c_void_p, _dp, _ip, _dp]),
}
