"""Batched likelihood engine: the host-side mirror of the reference's hot path.

Where the reference calls, per particle and per MCMC step,

    RVmodel::calculate_mu()   (src/RVmodel.cpp:286-395)
    RVmodel::calculate_C()    (src/RVmodel.cpp:170-216)
    RVmodel::log_likelihood() (src/RVmodel.cpp:827-943)

this class scores a whole batch of particles ("lanes") per call through the
C ABI of libkima_cuda (include/kima_cuda.h).  Names follow the reference.
"""
from __future__ import annotations

import ctypes
from typing import Optional

import numpy as np

from . import _lib
from .data import RVData
from .flags import CLayout, ModelFlags

_dp = ctypes.POINTER(ctypes.c_double)
_ip = ctypes.POINTER(ctypes.c_int32)


def _d(a: np.ndarray):
    return a.ctypes.data_as(_dp)


class LikelihoodEngine:
    def __init__(self, data: RVData, flags: ModelFlags, max_batch: int, device: int = 0,
                 kepler_mode: int = 0, gp_slots: int = 0, gp_streams: int = 0):
        self.lib = _lib.load_library()
        self.data = data
        self.flags = flags
        self.max_batch = int(max_batch)
        self._ctx = ctypes.c_void_p()
        _lib.check(self.lib.kc_create(device, ctypes.byref(self._ctx)))
        try:
            self._upload_data(data)
            _lib.check(self.lib.kc_set_option(self._ctx, b"kepler_mode", kepler_mode))
            if gp_slots:
                _lib.check(self.lib.kc_set_option(self._ctx, b"gp_slots", gp_slots))
            if gp_streams:
                _lib.check(self.lib.kc_set_option(self._ctx, b"gp_streams", gp_streams))
            cf = flags.to_c()
            _lib.check(self.lib.kc_configure(self._ctx, ctypes.byref(cf), self.max_batch))
            L = CLayout()
            _lib.check(self.lib.kc_get_layout(self._ctx, ctypes.byref(L)))
            self.layout = L.as_dict()
            pp, npp, ld = _dp(), _ip(), ctypes.c_int()
            _lib.check(self.lib.kc_host_params(self._ctx, ctypes.byref(pp), ctypes.byref(npp),
                                               ctypes.byref(ld)))
            self.ld = ld.value
            n_rows = self.layout["n_rows"]
            # numpy views of the pinned staging block the sampler shim fills
            self.params = np.ctypeslib.as_array(pp, shape=(n_rows, self.ld))
            self.np = np.ctypeslib.as_array(npp, shape=(self.ld,))
        except Exception:
            self.close()
            raise

    # ---- lifetime ----------------------------------------------------------------
    def close(self) -> None:
        if getattr(self, "_ctx", None) and self._ctx.value:
            self.params = None
            self.np = None
            self.lib.kc_destroy(self._ctx)
            self._ctx = ctypes.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def _upload_data(self, d: RVData) -> None:
        N = d.N()
        if N == 0:
            raise ValueError("empty data set")
        multi = d.number_instruments > 1
        obsi = np.ascontiguousarray(d.obsi, np.int32) if multi else None
        act = np.ascontiguousarray(d.actind, np.float64) if d.number_indicators else None
        _lib.check(self.lib.kc_upload_data(
            self._ctx, N, _d(np.ascontiguousarray(d.t)), _d(np.ascontiguousarray(d.y)),
            _d(np.ascontiguousarray(d.sig)),
            None if obsi is None else obsi.ctypes.data_as(_ip), d.number_instruments,
            d.number_indicators, None if act is None else _d(act), d.M0_epoch, d.get_t_middle()))

    # ---- parameters ----------------------------------------------------------------
    def set_params(self, params: np.ndarray, nps: np.ndarray) -> int:
        """Copy params[n_rows][B], nps[B] into the pinned staging block."""
        params = np.asarray(params, np.float64)
        nps = np.asarray(nps, np.int32)
        n_rows, B = params.shape
        if n_rows != self.layout["n_rows"]:
            raise ValueError("params has %d rows, layout needs %d" % (n_rows, self.layout["n_rows"]))
        if B > self.max_batch or nps.shape != (B,):
            raise ValueError("batch of %d does not fit max_batch=%d" % (B, self.max_batch))
        self.params[:, :B] = params
        self.np[:B] = nps
        return B

    # ---- the hot path ----------------------------------------------------------------
    def log_likelihood(self, B: int, out: Optional[np.ndarray] = None) -> np.ndarray:
        """logL of lanes [0,B) of the staged block (H2D, kernels, D2H)."""
        if out is None:
            out = np.empty(B, np.float64)
        _lib.check(self.lib.kc_eval(self._ctx, B, _d(out)))
        return out

    def set_option(self, key: str, value: int) -> None:
        _lib.check(self.lib.kc_set_option(self._ctx, key.encode(), int(value)))

    # ---- factor cache of the shared-memory-resident GP path (kima_cuda.h: kc_gp_cache_enable) ----
    def enable_gp_cache(self, n_slots: int) -> int:
        """Keep the last two factors of every slot; returns the active slot count (0: the cache does not
        apply to this context).  No counterpart in the reference, which factors C on every call."""
        _lib.check(self.lib.kc_gp_cache_enable(self._ctx, int(n_slots)))
        n = ctypes.c_int(0)
        _lib.check(self.lib.kc_gp_cache_slots(self._ctx, ctypes.byref(n)))
        return n.value

    def set_slots(self, slots) -> None:
        """Slots of the lanes of the next evaluation (consumed by it; < 0 = uncached lane)."""
        slots = np.ascontiguousarray(slots, np.int32)
        _lib.check(self.lib.kc_set_slots(self._ctx, slots.size, slots.ctypes.data_as(_ip)))

    def log_likelihood_async(self, B: int, out: np.ndarray) -> None:
        _lib.check(self.lib.kc_eval_async(self._ctx, B, _d(out)))

    def wait(self) -> None:
        _lib.check(self.lib.kc_wait(self._ctx))

    def log_likelihood_device(self, B: int) -> None:
        """Score the device-resident block; result stays on the device (async)."""
        _lib.check(self.lib.kc_eval_device(self._ctx, B, None))

    def sync(self) -> None:
        _lib.check(self.lib.kc_sync(self._ctx))

    def evaluate(self, params: np.ndarray, nps: np.ndarray) -> np.ndarray:
        B = self.set_params(params, nps)
        return self.log_likelihood(B)

    # ---- posterior-predictive curves (pykima/results.py:723-855) ------------------------
    def predict(self, params: np.ndarray, nps: np.ndarray, t: Optional[np.ndarray] = None,
                model: bool = True, gp: Optional[bool] = None):
        """Deterministic model and GP predictive mean of every parameter vector on the grid
        ``t`` (None = the data epochs).  Returns ``(model[B, nt] | None, gp[B, nt] | None)``."""
        B = self.set_params(params, nps)
        if gp is None:
            gp = bool(self.flags.GP)  # kernel=celerite: the solver's own predict (sorted grid)
        tt = None if t is None else np.ascontiguousarray(t, np.float64)
        nt = self.data.N() if tt is None else tt.size
        what = (_lib.KC_PRED_MODEL if model else 0) | (_lib.KC_PRED_GP if gp else 0)
        m = np.empty((B, nt)) if model else None
        g = np.empty((B, nt)) if gp else None
        _lib.check(self.lib.kc_predict(self._ctx, B, nt, None if tt is None else _d(tt), what,
                                       None if m is None else _d(m), None if g is None else _d(g)))
        return m, g

    def predict_std(self, params: np.ndarray, nps: np.ndarray, t: Optional[np.ndarray] = None,
                    mean: bool = True):
        """GP predictive mean and standard deviation, pykima's ``GP.predict(y, t, return_std=True)``
        (pykima/GP.py:113-131) for every parameter vector.  Returns ``(mean[B, nt] | None, std[B, nt])``."""
        B = self.set_params(params, nps)
        tt = None if t is None else np.ascontiguousarray(t, np.float64)
        nt = self.data.N() if tt is None else tt.size
        g = np.empty((B, nt)) if mean else None
        sd = np.empty((B, nt))
        _lib.check(self.lib.kc_predict_std(self._ctx, B, nt, None if tt is None else _d(tt),
                                           None if g is None else _d(g), _d(sd)))
        return g, sd

    @staticmethod
    def likelihood_type_value(logL: np.ndarray) -> np.ndarray:
        """What DNest4's LikelihoodType stores (LikelihoodType.cpp:17-24)."""
        logL = np.asarray(logL, np.float64)
        return np.where(np.isfinite(logL), logL, -1e300)

    # ---- instrumentation ------------------------------------------------------------
    def last_kernel_ms(self) -> float:
        a, b = ctypes.c_float(), ctypes.c_float()
        _lib.check(self.lib.kc_last_kernel_ms(self._ctx, ctypes.byref(a), ctypes.byref(b)))
        return a.value + b.value

    def flush_l2(self) -> None:
        _lib.check(self.lib.kc_flush_l2(self._ctx))

    def launch_count(self) -> int:
        return int(self.lib.kc_launch_count(self._ctx))

    def measure_fp64_peak(self, reps: int = 5):
        a, b = ctypes.c_double(), ctypes.c_double()
        _lib.check(self.lib.kc_measure_fp64_peak(self._ctx, reps, ctypes.byref(a), ctypes.byref(b)))
        return a.value, b.value

    # ---- test hooks --------------------------------------------------------------------
    def kepler(self, M, ecc) -> np.ndarray:
        M = np.ascontiguousarray(M, np.float64)
        ecc = np.ascontiguousarray(ecc, np.float64)
        E = np.empty(M.size)
        _lib.check(self.lib.kc_kepler_E(self._ctx, M.size, _d(M), _d(ecc), _d(E)))
        return E

    def true_anomaly(self, t, P, ecc, Tp) -> np.ndarray:
        a = [np.ascontiguousarray(x, np.float64) for x in (t, P, ecc, Tp)]
        f = np.empty(a[0].size)
        _lib.check(self.lib.kc_true_anomaly(self._ctx, a[0].size, _d(a[0]), _d(a[1]), _d(a[2]),
                                            _d(a[3]), _d(f)))
        return f

    def calculate_mu(self, lane: int) -> np.ndarray:
        mu = np.empty(self.data.N())
        _lib.check(self.lib.kc_model_mu(self._ctx, lane, _d(mu)))
        return mu

    def calculate_C(self, lane: int) -> np.ndarray:
        N = self.data.N()
        C = np.empty((N, N))
        _lib.check(self.lib.kc_gp_cov(self._ctx, lane, _d(C)))
        return C.T.copy()
