"""ctypes binding of libkima_cuda.so (include/kima_cuda.h).  Fails loudly when the
library is missing or no GPU is present: there is no CPU fallback by design."""
from __future__ import annotations

import ctypes
import os

from .flags import CLayout, CModelFlags
from .sampler_types import SAMPLER_SIGNATURES

HERE = os.path.dirname(os.path.abspath(__file__))
# KIMA_CUDA_LIB: a developer build of the same library (e.g. with the in-kernel phase profile)
LIB_PATH = os.environ.get("KIMA_CUDA_LIB") or os.path.join(HERE, "libkima_cuda.so")

KC_OK = 0
KC_ERR_ARG, KC_ERR_CUDA, KC_ERR_UNSUPPORTED, KC_ERR_COMM = 1, 2, 3, 4
KC_PRED_MODEL, KC_PRED_GP, KC_PRED_GP_STD = 1, 2, 4


class KimaCudaError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__("libkima_cuda error %d: %s" % (code, msg))
        self.code = code


_dp = ctypes.POINTER(ctypes.c_double)
_ip = ctypes.POINTER(ctypes.c_int32)
_vp = ctypes.c_void_p

# name -> (restype, argtypes); every symbol include/kima_cuda.h declares
SIGNATURES = {
    "kc_create": (ctypes.c_int, [ctypes.c_int, ctypes.POINTER(_vp)]),
    "kc_destroy": (ctypes.c_int, [_vp]),
    "kc_last_error": (ctypes.c_char_p, []),
    "kc_version": (ctypes.c_char_p, []),
    "kc_upload_data": (ctypes.c_int, [_vp, ctypes.c_int, _dp, _dp, _dp, _ip, ctypes.c_int,
                                      ctypes.c_int, _dp, ctypes.c_double, ctypes.c_double]),
    "kc_configure": (ctypes.c_int, [_vp, ctypes.POINTER(CModelFlags), ctypes.c_int]),
    "kc_get_layout": (ctypes.c_int, [_vp, ctypes.POINTER(CLayout)]),
    "kc_host_params": (ctypes.c_int, [_vp, ctypes.POINTER(_dp), ctypes.POINTER(_ip),
                                      ctypes.POINTER(ctypes.c_int)]),
    "kc_device_params": (ctypes.c_int, [_vp, ctypes.POINTER(_vp), ctypes.POINTER(_vp),
                                        ctypes.POINTER(ctypes.c_int)]),
    "kc_eval": (ctypes.c_int, [_vp, ctypes.c_int, _dp]),
    "kc_eval_async": (ctypes.c_int, [_vp, ctypes.c_int, _dp]),
    "kc_wait": (ctypes.c_int, [_vp]),
    "kc_eval_device": (ctypes.c_int, [_vp, ctypes.c_int, _vp]),
    "kc_eval_celerite_noalloc": (ctypes.c_int, [_vp, ctypes.c_int, _vp, _vp, ctypes.c_int, _vp]),
    "kc_sync": (ctypes.c_int, [_vp]),
    "kc_stream": (ctypes.c_int, [_vp, ctypes.POINTER(_vp)]),
    "kc_set_option": (ctypes.c_int, [_vp, ctypes.c_char_p, ctypes.c_int]),
    "kc_gp_cache_enable": (ctypes.c_int, [_vp, ctypes.c_int]),
    "kc_gp_cache_slots": (ctypes.c_int, [_vp, ctypes.POINTER(ctypes.c_int)]),
    "kc_set_slots": (ctypes.c_int, [_vp, ctypes.c_int, _ip]),
    "kc_last_kernel_ms": (ctypes.c_int, [_vp, ctypes.POINTER(ctypes.c_float),
                                         ctypes.POINTER(ctypes.c_float)]),
    "kc_flush_l2": (ctypes.c_int, [_vp]),
    "kc_launch_count": (ctypes.c_int64, [_vp]),
    "kc_measure_fp64_peak": (ctypes.c_int, [_vp, ctypes.c_int, _dp, _dp]),
    "kc_kepler_E": (ctypes.c_int, [_vp, ctypes.c_int, _dp, _dp, _dp]),
    "kc_true_anomaly": (ctypes.c_int, [_vp, ctypes.c_int, _dp, _dp, _dp, _dp, _dp]),
    "kc_model_mu": (ctypes.c_int, [_vp, ctypes.c_int, _dp]),
    "kc_gp_cov": (ctypes.c_int, [_vp, ctypes.c_int, _dp]),
    "kc_predict": (ctypes.c_int, [_vp, ctypes.c_int, ctypes.c_int, _dp, ctypes.c_int, _dp, _dp]),
    "kc_predict_std": (ctypes.c_int, [_vp, ctypes.c_int, ctypes.c_int, _dp, _dp, _dp]),
    "kc_comm_unique_id": (ctypes.c_int, [_vp]),
    "kc_comm_init": (ctypes.c_int, [_vp, ctypes.c_int, ctypes.c_int, _vp]),
    "kc_allgather_above": (ctypes.c_int, [_vp, _dp, ctypes.c_int, _dp, ctypes.c_int,
                                          ctypes.POINTER(ctypes.c_int)]),
    "kc_allgather_bytes": (ctypes.c_int, [_vp, _vp, ctypes.c_size_t, _vp]),
    "kc_allreduce_levels": (ctypes.c_int, [_vp, ctypes.POINTER(ctypes.c_uint64), ctypes.c_int]),
    "kc_comm_destroy": (ctypes.c_int, [_vp]),
}

# ... and every symbol include/kima_sampler.h declares
SIGNATURES.update(SAMPLER_SIGNATURES)

_lib = None


def load_library() -> ctypes.CDLL:
    """dlopen the in-tree library and bind every declared entry point."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise FileNotFoundError(
            "%s is missing: build it with `python -m kima_b200.build` "
            "(nvcc, sm_100a).  There is no CPU fallback." % LIB_PATH)
    lib = ctypes.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the symbol is not exported
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(rc: int) -> None:
    if rc != KC_OK:
        msg = load_library().kc_last_error()
        raise KimaCudaError(rc, msg.decode() if msg else "")
