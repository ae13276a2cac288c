"""kc_eval_celerite_noalloc (include/kima_cuda.h): the allocation-free variant of the
kernel=celerite likelihood, whose residual scratch is a file-scope `static __device__` array.

The entry point has no counterpart in the reference.
It is checked against the existing entry point (kc_eval) on the same parameters,
through the C ABI, on two calls in a row (different parameters and batch sizes, so that what the
first call leaves in the shared array would show).  This file sorts last on
purpose.
"""
import ctypes

import numpy as np
import pytest

from kima_b200 import LikelihoodEngine, ModelFlags, _lib
from kima_b200.synthetic import make_data, make_params
from conftest import relerr

pytestmark = pytest.mark.gpu

TOL = 1e-10  # the relative tolerance of the GPU parity tests (north_star)
MAX_DOUBLES = 1 << 19  # KC_NOALLOC_MAX_DOUBLES


def _device_block(eng):
    dpar, dnp, ld = ctypes.c_void_p(), ctypes.c_void_p(), ctypes.c_int()
    _lib.check(eng.lib.kc_device_params(eng._ctx, ctypes.byref(dpar), ctypes.byref(dnp), ctypes.byref(ld)))
    return dpar, dnp, ld.value


@pytest.mark.parametrize("flags,N", [
    (ModelFlags(GP=True, kernel=1, npmax=3), 300),
    (ModelFlags(GP=True, kernel=1, npmax=2, MA=True, trend=True, degree=2), 177),
])
def test_noalloc_matches_kc_eval_twice_in_a_row(gpu, flags, N):
    import torch
    d = make_data(N)
    Bmax = 1024
    with LikelihoodEngine(d, flags, Bmax, device=0) as eng:
        dpar, dnp, ld = _device_block(eng)
        out = torch.full((Bmax,), float("nan"), dtype=torch.float64, device="cuda:0")
        torch.cuda.synchronize()
        # set-up only (B = 0): resolves the array's address on this device
        _lib.check(eng.lib.kc_eval_celerite_noalloc(eng._ctx, 0, None, None, 0, None))
        for call, (B, seed) in enumerate(((1024, 43), (200, 44))):
            P, nps = make_params(flags, d, B, "mix", seed=seed)
            want = eng.evaluate(P, nps)  # existing entry point; leaves the block staged in HBM
            out.fill_(float("nan"))
            torch.cuda.synchronize()
            _lib.check(eng.lib.kc_eval_celerite_noalloc(eng._ctx, B, dpar, dnp, ld,
                                                        ctypes.c_void_p(out.data_ptr())))
            eng.sync()
            got = out[:B].cpu().numpy()
            assert np.all(np.isfinite(want) | np.isposinf(want))
            assert relerr(got, want).max() <= TOL, "call %d: %g" % (call, relerr(got, want).max())
            assert torch.isnan(out[B:]).all()  # nothing written past lane B


def test_noalloc_rejects_sizes_beyond_the_static_scratch(gpu):
    import torch
    flags = ModelFlags(GP=True, kernel=1, npmax=1)
    N = 300
    d = make_data(N)
    B = 2048  # 300 * 2048 doubles > KC_NOALLOC_MAX_DOUBLES
    assert N * B > MAX_DOUBLES
    with LikelihoodEngine(d, flags, B, device=0) as eng:
        dpar, dnp, ld = _device_block(eng)
        out = torch.zeros(B, dtype=torch.float64, device="cuda:0")
        rc = eng.lib.kc_eval_celerite_noalloc(eng._ctx, B, dpar, dnp, ld, ctypes.c_void_p(out.data_ptr()))
        assert rc == _lib.KC_ERR_ARG
        msg = eng.lib.kc_last_error().decode()
        assert "KC_NOALLOC_MAX_DOUBLES" in msg and str(MAX_DOUBLES) in msg
        # the largest batch that fits still runs
        Bfit = (MAX_DOUBLES // N) // 32 * 32
        P, nps = make_params(flags, d, Bfit, "max")
        want = eng.evaluate(P, nps)
        _lib.check(eng.lib.kc_eval_celerite_noalloc(eng._ctx, Bfit, dpar, dnp, ld,
                                                    ctypes.c_void_p(out.data_ptr())))
        eng.sync()
        assert relerr(out[:Bfit].cpu().numpy(), want).max() <= TOL


def test_noalloc_needs_the_celerite_kernel(gpu):
    flags = ModelFlags(GP=True, npmax=1)  # standard kernel
    d = make_data(50)
    with LikelihoodEngine(d, flags, 8, device=0) as eng:
        dpar, dnp, ld = _device_block(eng)
        rc = eng.lib.kc_eval_celerite_noalloc(eng._ctx, 8, dpar, dnp, ld, dpar)
        assert rc == _lib.KC_ERR_ARG
