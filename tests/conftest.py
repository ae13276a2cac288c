import glob
import json
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")
    config.addinivalue_line("markers", "reference: needs oracle/_ref built from /root/reference")


def _build_once():
    # building a checker / the library is not using it; both are cheap no-ops when fresh
    from kima_b200 import build as kbuild
    from oracle import oracle as orc
    try:
        kbuild.build()
    except Exception as e:  # no nvcc on this machine: CPU tests that need the .so will say so
        print("libkima_cuda build skipped:", e)
    if not os.path.exists(orc.PORT_SO) or (os.path.isdir("/root/reference/src") and not orc.have_reference()):
        orc.build()


_build_once()


def golden_cases():
    return sorted(os.path.basename(p)[:-4] for p in glob.glob(os.path.join(GOLDEN, "*.npz"))
                  if not p.endswith("kepler.npz") and not os.path.basename(p).startswith("predictive_")
                  and "N4096" not in p and not os.path.basename(p).startswith("tail_"))  # the scale point has its own tests (12 s per lane on the CPU)


def load_golden(name):
    from kima_b200.data import RVData
    from kima_b200.flags import ModelFlags
    z = np.load(os.path.join(GOLDEN, name + ".npz"))
    f = json.loads(str(z["flags"]))
    flags = ModelFlags(**{k: (bool(v) if k not in ("degree", "n_known_object", "npmax", "kernel") else int(v))
                          for k, v in f.items()})
    n_inst = int(z["n_inst"])
    data = RVData.from_arrays(z["t"], z["y"], z["sig"], z["obsi"] if n_inst > 1 else None,
                              z["actind"] if z["actind"].size else None, n_inst=n_inst)
    return dict(name=name, flags=flags, data=data, params=z["params"], nps=z["nps"],
                logL=z["logL"], mu0=z["mu0"], mu0_lo=z["mu0_lo"],
                C0=z["C0"] if "C0" in z.files else None, layout=json.loads(str(z["layout"])),
                load_info=json.loads(str(z["load_info"])))


@pytest.fixture(scope="session")
def port():
    from oracle.oracle import Oracle
    return Oracle("port")


@pytest.fixture(scope="session")
def reference():
    from oracle.oracle import Oracle, have_reference
    if not have_reference():
        pytest.skip("oracle/_ref/libkima_ref.so not built (no /root/reference here)")
    return Oracle("reference")


def have_gpu():
    try:
        import ctypes
        from kima_b200 import _lib
        lib = _lib.load_library()
        ctx = ctypes.c_void_p()
        if lib.kc_create(0, ctypes.byref(ctx)) != 0:
            return False
        lib.kc_destroy(ctx)
        return True
    except Exception:
        return False


@pytest.fixture(scope="session")
def gpu():
    if not have_gpu():
        pytest.fail("no CUDA device or libkima_cuda.so missing: GPU tests cannot run "
                    "(there is no CPU fallback)")
    return True


def relerr(a, b):
    a = np.asarray(a, np.float64)
    b = np.asarray(b, np.float64)
    both_inf = np.isinf(a) & np.isinf(b) & (np.sign(a) == np.sign(b))
    d = np.abs(a - b) / np.maximum(np.abs(b), 1e-300)
    return np.where(both_inf, 0.0, d)
