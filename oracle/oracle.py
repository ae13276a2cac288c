"""TEST INFRASTRUCTURE — ctypes front-end of the two CPU checkers.

* ``Oracle("port")``      -> oracle/libkima_oracle.so  (plain-C restatement)
* ``Oracle("reference")`` -> oracle/_ref/libkima_ref.so (the unmodified reference
  sources compiled by oracle/Makefile; exists only where it was built)

Only tests/, ``__graft_entry__.smoke()`` and bench.py's cpu_baseline /
``--impl reference`` legs may import this module.  Nothing in kima_b200/ does.
"""
from __future__ import annotations

import ctypes
import os
import subprocess
from typing import Optional, Sequence

import numpy as np

from kima_b200.flags import CLayout, CModelFlags, ModelFlags

HERE = os.path.dirname(os.path.abspath(__file__))
PORT_SO = os.path.join(HERE, "libkima_oracle.so")
REF_SO = os.path.join(HERE, "_ref", "libkima_ref.so")
# the reference sources minus DNest4/code/RNG.cpp, on the counter-based stream (oracle/philox_rng.cpp)
PHILOX_SO = os.path.join(HERE, "_ref", "libkima_ref_philox.so")
REFERENCE_ROOT = "/root/reference"

_dp = ctypes.POINTER(ctypes.c_double)
_ip = ctypes.POINTER(ctypes.c_int)


def _d(a):
    return a.ctypes.data_as(_dp)


def _i(a):
    return a.ctypes.data_as(_ip)


def build(ref: Optional[bool] = None, quiet: bool = True) -> None:
    """Compile the checkers (building a checker is not using it)."""
    targets = ["oracle"]
    if ref is None:
        ref = os.path.isdir(os.path.join(REFERENCE_ROOT, "src"))
    if ref:
        targets.append("ref")
    subprocess.run(["make", "-C", HERE, "-j8"] + targets, check=True,
                   stdout=subprocess.DEVNULL if quiet else None)


def have_reference() -> bool:
    return os.path.exists(REF_SO)


class Oracle:
    def __init__(self, kind: str = "port"):
        if kind not in ("port", "reference", "philox"):
            raise ValueError(kind)
        path = PORT_SO if kind == "port" else (REF_SO if kind == "reference" else PHILOX_SO)
        # "philox" is the reference harness as well, in a shared object of its own (own Data
        # singleton, own flags) that also holds the sampler harness (oracle/ref_sampler.py)
        kind = "port" if kind == "port" else "reference"
        self.kind = kind
        if not os.path.exists(path):
            raise FileNotFoundError(
                "%s is not built: run `make -C oracle %s`" % (path, "oracle" if kind == "port" else "ref"))
        self.lib = ctypes.CDLL(path)
        self.p = "ko_" if kind == "port" else "ref_"
        L = self.lib
        f = getattr(L, self.p + "eval_batch")
        f.restype = ctypes.c_double
        f.argtypes = [ctypes.c_int, ctypes.c_int, _dp, _ip, _dp, ctypes.c_int]
        self._eval = f
        self._flags = ModelFlags()
        self._fix = False
        self.N = 0
        self.n_inst = 1
        self.n_ind = 0

    # ---- data ----------------------------------------------------------------
    def set_data(self, t, y, sig, obsi=None, n_inst: int = 1, actind=None) -> None:
        t = np.ascontiguousarray(t, np.float64)
        y = np.ascontiguousarray(y, np.float64)
        sig = np.ascontiguousarray(sig, np.float64)
        N = t.size
        ob = None if obsi is None else np.ascontiguousarray(obsi, np.int32)
        act = None
        n_ind = 0
        if actind is not None and np.size(actind):
            act = np.ascontiguousarray(actind, np.float64).reshape(-1, N)
            n_ind = act.shape[0]
        fn = getattr(self.lib, self.p + "set_data")
        fn(ctypes.c_int(N), _d(t), _d(y), _d(sig), None if ob is None else _i(ob),
           ctypes.c_int(n_inst), ctypes.c_int(n_ind), None if act is None else _d(act))
        self.N, self.n_inst, self.n_ind = N, n_inst, n_ind
        self.set_flags(self._flags, self._fix)

    def set_rvdata(self, d) -> None:
        """From a kima_b200.data.RVData."""
        multi = d.number_instruments > 1
        self.set_data(d.t, d.y, d.sig, d.obsi if multi else None, d.number_instruments,
                      d.actind if d.number_indicators else None)

    # reference-only: the reference's own file parsers (src/Data.cpp)
    def ref_load(self, filename: str, units: str, skip: int, indicators: Sequence[str] = ()):
        assert self.kind == "reference"
        if indicators:
            arr = (ctypes.c_char_p * len(indicators))(*[s.encode() for s in indicators])
            self.lib.ref_load_indicators(filename.encode(), units.encode(), skip, len(indicators), arr)
        else:
            self.lib.ref_load(filename.encode(), units.encode(), ctypes.c_int(skip))
        return self._after_load()

    def ref_load_multi(self, filename_or_files, units: str, skip: int):
        assert self.kind == "reference"
        if isinstance(filename_or_files, str):
            self.lib.ref_load_multi(filename_or_files.encode(), units.encode(), ctypes.c_int(skip))
        else:
            fs = list(filename_or_files)
            arr = (ctypes.c_char_p * len(fs))(*[s.encode() for s in fs])
            self.lib.ref_load_multi_files(len(fs), arr, units.encode(), ctypes.c_int(skip))
        return self._after_load()

    def _after_load(self):
        N, ni, nd = ctypes.c_int(), ctypes.c_int(), ctypes.c_int()
        self.lib.ref_data_shape(ctypes.byref(N), ctypes.byref(ni), ctypes.byref(nd))
        self.N, self.n_inst, self.n_ind = N.value, ni.value, nd.value
        self.set_flags(self._flags, self._fix)
        return self.get_data()

    def get_data(self) -> dict:
        assert self.kind == "reference"
        N = self.N
        t, y, sig = np.zeros(N), np.zeros(N), np.zeros(N)
        obsi = np.zeros(N, np.int32)
        act = np.zeros((self.n_ind, N))
        m0, tm = ctypes.c_double(), ctypes.c_double()
        self.lib.ref_get_data(_d(t), _d(y), _d(sig), _i(obsi), _d(act), ctypes.byref(m0), ctypes.byref(tm))
        return dict(t=t, y=y, sig=sig, obsi=obsi, actind=act, M0_epoch=m0.value,
                    t_middle=tm.value, n_inst=self.n_inst, n_ind=self.n_ind)

    # ---- flags ---------------------------------------------------------------
    def set_flags(self, flags: ModelFlags, fix: bool = False) -> None:
        self._flags, self._fix = flags, fix
        c = flags.to_c()
        if self.kind == "reference":
            self.lib.ref_set_flags(ctypes.byref(c), ctypes.c_int(int(fix)))
        else:
            self.lib.ko_set_flags(ctypes.byref(c))

    def layout(self) -> dict:
        L = CLayout()
        getattr(self.lib, self.p + "get_layout")(ctypes.byref(L))
        return L.as_dict()

    # ---- the hot path --------------------------------------------------------
    def eval_batch(self, params, np_, nthreads: int = 1):
        """logL[B] and the seconds spent, for params[n_rows][B], np_[B]."""
        params = np.ascontiguousarray(params, np.float64)
        np_ = np.ascontiguousarray(np_, np.int32)
        n_rows, B = params.shape
        assert n_rows == self.layout()["n_rows"], (n_rows, self.layout()["n_rows"])
        assert np_.size == B
        out = np.zeros(B)
        secs = self._eval(B, B, _d(params), _i(np_), _d(out), nthreads)
        return out, secs

    def mu(self, params, b: int, np_b: int, with_lo: bool = False):
        params = np.ascontiguousarray(params, np.float64)
        mu = np.zeros(self.N)
        lo = np.zeros(self.N)
        getattr(self.lib, self.p + "mu")(params.shape[1], _d(params), b, int(np_b), _d(mu), _d(lo))
        return (mu, lo) if with_lo else mu

    def cov(self, params, b: int):
        params = np.ascontiguousarray(params, np.float64)
        C = np.zeros((self.N, self.N))
        getattr(self.lib, self.p + "cov")(params.shape[1], _d(params), b, 0, _d(C))
        return C.T.copy()  # column-major -> C[i, j]

    def celerite_predict(self, params, b: int, np_b: int, x):
        """kernel=celerite: the reference solver's own conditional mean at the sorted times x
        (CholeskySolver::predict, celerite/.../solver/cholesky.h:590-690).  Reference kind only."""
        if self.kind != "reference":
            raise RuntimeError("celerite_predict calls the reference's solver: Oracle('reference') only")
        params = np.ascontiguousarray(params, np.float64)
        x = np.ascontiguousarray(x, np.float64)
        out = np.zeros(x.size)
        fn = self.lib.ref_celerite_predict
        fn.restype = ctypes.c_int
        rc = fn(ctypes.c_int(params.shape[1]), _d(params), ctypes.c_int(b), ctypes.c_int(int(np_b)),
                ctypes.c_int(x.size), _d(x), _d(out))
        return out if rc == 0 else np.full(x.size, np.nan)

    def kepler(self, M, ecc):
        M = np.ascontiguousarray(M, np.float64)
        ecc = np.ascontiguousarray(ecc, np.float64)
        E = np.zeros(M.size)
        name = "ko_kepler_n" if self.kind == "port" else "ref_kepler"
        getattr(self.lib, name)(M.size, _d(M), _d(ecc), _d(E))
        return E

    def true_anomaly(self, t, P, ecc, Tp):
        a = [np.ascontiguousarray(x, np.float64) for x in (t, P, ecc, Tp)]
        f = np.zeros(a[0].size)
        name = "ko_true_anomaly_n" if self.kind == "port" else "ref_true_anomaly"
        getattr(self.lib, name)(a[0].size, _d(a[0]), _d(a[1]), _d(a[2]), _d(a[3]), _d(f))
        return f

    def likelihood_type_value(self, v: float) -> float:
        fn = getattr(self.lib, self.p + "likelihood_type_value")
        fn.restype = ctypes.c_double
        fn.argtypes = [ctypes.c_double]
        return fn(v)

    def cholesky(self, A):
        """port only: Eigen-style blocked LLT of a column-major copy; (L, info)."""
        assert self.kind == "port"
        A = np.array(A, dtype=np.float64, order="F")
        n = A.shape[0]
        info = self.lib.ko_cholesky(A.ctypes.data_as(_dp), n)
        return np.tril(A), info
