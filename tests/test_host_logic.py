"""CPU: host-side logic, the C-ABI surface, and the device math headers compiled for the host."""
import ctypes
import os
import re

import numpy as np
import pytest

from conftest import GOLDEN, ROOT, load_golden
from kima_b200 import _lib
from kima_b200.data import RVData
from kima_b200.flags import CLayout, ModelFlags, layout_from_flags
from kima_b200.dist import merge_above_and_create_level, shard_lanes


def test_library_exports_every_declared_symbol():
    hdr = open(os.path.join(ROOT, "include", "kima_cuda.h")).read() + \
        open(os.path.join(ROOT, "include", "kima_sampler.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    declared = set(re.findall(r"\b(kc_[A-Za-z0-9_]+)\s*\(", hdr)) - {"kc_layout_from_flags"}
    assert declared == set(_lib.SIGNATURES), declared ^ set(_lib.SIGNATURES)
    lib = _lib.load_library()  # raises if the .so is missing or lacks a symbol
    for name in declared:
        assert hasattr(lib, name)
    assert b"sm_100a" in lib.kc_version()


def test_no_gpu_means_loud_failure_not_fallback():
    lib = _lib.load_library()
    ctx = ctypes.c_void_p()
    rc = lib.kc_create(0, ctypes.byref(ctx))
    if rc == 0:  # a GPU is present: nothing to check here
        lib.kc_destroy(ctx)
        pytest.skip("GPU present")
    assert rc == _lib.KC_ERR_CUDA and b"no CPU fallback" in lib.kc_last_error()
    from kima_b200 import LikelihoodEngine, KimaCudaError
    from kima_b200.synthetic import make_data
    with pytest.raises(KimaCudaError):
        LikelihoodEngine(make_data(16), ModelFlags(), 4)


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "kima_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                src = open(os.path.join(dirpath, f)).read()
                for pat in (r"(from|import)\s+oracle", r"libkima_oracle", r"libkima_ref", r"oracle/", r"ko_eval", r"ref_eval"):
                    assert not re.search(pat, src), (f, pat)


def test_product_host_build_uses_the_gpu_evaluator_only():
    """host/overlay/kima.h lets a TEST name another Evaluator (tests/cpp/RefEvaluator.h: the reference's CPU
    likelihood, to test the sampler without a GPU); the product build never does."""
    mk = open(os.path.join(ROOT, "host", "Makefile")).read()
    assert "KIMA_B200_EVALUATOR" not in mk and "tests/cpp" not in mk and "RefEvaluator" not in mk
    for f in os.listdir(os.path.join(ROOT, "host")):
        if f.endswith((".h", ".hpp", ".cpp")):
            src = open(os.path.join(ROOT, "host", f)).read()
            assert not re.search(r'#\s*include\s*"RefEvaluator.h"', src), f
    overlay = open(os.path.join(ROOT, "host", "overlay", "kima.h")).read()
    assert '#define KIMA_B200_EVALUATOR_HEADER "RVmodelDevice.h"' in overlay   # the default


@pytest.mark.parametrize("flags,n_inst,n_ind", [
    (ModelFlags(), 1, 0), (ModelFlags(npmax=10, hyperpriors=True), 1, 0),
    (ModelFlags(GP=True, npmax=5), 1, 0), (ModelFlags(npmax=2, multi_instrument=True), 4, 0),
    (ModelFlags(npmax=3, studentt=True, multi_instrument=True, trend=True, degree=3,
                known_object=True, n_known_object=2), 3, 2)])
def test_layout_python_equals_c(port, flags, n_inst, n_ind):
    N = 8
    obsi = (np.arange(N) % n_inst + 1) if n_inst > 1 else None
    act = np.zeros((n_ind, N)) if n_ind else None
    port.set_data(np.arange(N) * 1.0, np.zeros(N), np.ones(N), obsi, n_inst, act)
    port.set_flags(flags)
    L = layout_from_flags(flags, n_inst, n_ind)
    assert L == port.layout()


def test_model_flags_errors_follow_setPriors():
    with pytest.raises(ValueError, match="degree=0"):
        ModelFlags(trend=True, degree=0)
    with pytest.raises(ValueError, match="3rd degree"):
        ModelFlags(trend=True, degree=4)


def _write(tmp_path, name, lines):
    p = tmp_path / name
    p.write_text("\n".join(lines) + "\n")
    return str(p)


def test_rvdata_load_formats(tmp_path):
    f = _write(tmp_path, "a.rv", ["jdb vrad svrad", "--- ---- -----", "1.5 0.010 0.002", "0.5\t0.020\t0.001", "2.5 0.0 0.003"])
    d = RVData.load(f, "kms")  # default skip=2
    assert np.array_equal(d.t, [1.5, 0.5, 2.5]) and np.allclose(d.y, [10.0, 20.0, 0.0])
    assert np.allclose(d.sig, [2.0, 1.0, 3.0]) and d.M0_epoch == 1.5 and d.get_t_middle() == 1.5
    assert d.number_instruments == 1 and d.number_indicators == 0 and not d.datamulti
    # indicators, with a skipped column; means are removed (src/Data.cpp:166-173)
    f2 = _write(tmp_path, "b.txt", ["1 1 1 10 99 1", "2 2 1 20 99 3", "3 3 1 30 99 5"])
    d2 = RVData.load(f2, "ms", 0, ["fwhm", "", "bis"])
    assert d2.indicator_names == ["fwhm", "bis"] and np.allclose(d2.actind, [[-10, 0, 10], [-2, 0, 2]])
    # joined multi-instrument file: consecutive ids renumbered from 1 (src/Data.cpp:247-269)
    f3 = _write(tmp_path, "c.txt", ["# h", "1 1 1 0", "2 2 1 0", "3 3 1 7", "4 4 1 2"])
    d3 = RVData.load_multi(f3, "ms", 1)
    assert d3.obsi.tolist() == [1, 1, 2, 3] and d3.number_instruments == 3 and d3.datamulti
    # several files: merged and sorted by time (src/Data.cpp:423-458)
    fa = _write(tmp_path, "i1.txt", ["h", "h", "10 1 1", "30 3 1"])
    fb = _write(tmp_path, "i2.txt", ["h", "h", "20 2 1", "5 0.5 1"])
    d4 = RVData.load_multi([fa, fb], "ms", 2)
    assert d4.t.tolist() == [5, 10, 20, 30] and d4.obsi.tolist() == [2, 1, 2, 1]
    assert d4.get_max_RV_span() == 2.0 and d4.get_RV_span() == 2.5
    with pytest.raises(FileNotFoundError):
        RVData.load(str(tmp_path / "missing.rv"))


def test_rvdata_matches_golden_example_arrays():
    # the arrays the reference's parsers left behind are in the fixtures; when the example
    # files themselves are present (build container) the Python loader must reproduce them
    for name in ("51Peg", "many_planets", "CoRoT7", "multi_instrument", "multi_instrument_joined",
                 "activity_correlations"):
        g = load_golden(name)
        li = g["load_info"]
        paths = li.get("files") or [li["file"]]
        paths = [os.path.join("/root/reference", p) for p in paths]
        if not all(os.path.exists(p) for p in paths):
            pytest.skip("example files not on this machine")
        if "files" in li:
            d = RVData.load_multi(paths, li["units"], li["skip"])
        elif name == "multi_instrument_joined":
            d = RVData.load_multi(paths[0], li["units"], li["skip"])
        else:
            d = RVData.load(paths[0], li["units"], li["skip"], li.get("indicators", ()))
        assert np.array_equal(d.t, g["data"].t) and np.array_equal(d.y, g["data"].y)
        assert np.array_equal(d.sig, g["data"].sig)
        if d.number_instruments > 1:
            assert np.array_equal(d.obsi, g["data"].obsi)


def test_shard_lanes_partition():
    for total, world in ((8192, 8), (10, 4), (3, 8), (0, 2)):
        spans = [shard_lanes(total, r, world) for r in range(world)]
        assert spans[0][0] == 0 and spans[-1][1] == total
        assert all(spans[i][1] == spans[i + 1][0] for i in range(world - 1))
        sizes = [b - a for a, b in spans]
        assert max(sizes) - min(sizes) <= 1


def test_level_creation_rule():
    rng = np.random.default_rng(5)
    above = np.stack([rng.normal(size=200), rng.uniform(size=200)], axis=1)
    assert merge_above_and_create_level(above[:10], 50)[0] is None
    lvl, rest = merge_above_and_create_level(above, 50)
    srt = above[np.lexsort((above[:, 1], above[:, 0]))]
    idx = int((1 - 1 / np.e) * 200)
    assert np.array_equal(lvl, srt[idx]) and np.array_equal(rest, srt[idx + 1:])


# ---- device math headers, compiled for the host (MUFU seeds emulated) ---------------------
@pytest.fixture(scope="module")
def hc():
    path = os.path.join(ROOT, "kima_b200", "libkc_hostcheck.so")
    if not os.path.exists(path):
        pytest.skip("host check library not built")
    return ctypes.CDLL(path)


_dp = ctypes.POINTER(ctypes.c_double)


def _d(a):
    return a.ctypes.data_as(_dp)


def test_device_kepler_E_within_1e12_of_golden(hc):
    z = np.load(os.path.join(GOLDEN, "kepler.npz"))
    M, e = np.ascontiguousarray(z["M"]), np.ascontiguousarray(z["e"])
    E = np.zeros(M.size)
    hc.hc_kepler_E(M.size, _d(M), _d(e), _d(E))
    assert np.nanmax(np.abs(E - z["E"])) <= 1e-12  # north_star: eccentric anomaly <= 1e-12 rad
    assert np.array_equal(np.isnan(E), np.isnan(z["E"]))


def test_device_mod_two_pi_is_exact_fmod(hc):
    import math
    rng = np.random.default_rng(1)
    M = np.concatenate([rng.uniform(-1e6, 1e6, 20000), rng.uniform(0, 1e12, 2000),
                        np.array([0.0, 2 * math.pi, -2 * math.pi, 4 * math.pi, math.pi, 1e-300])])
    r = np.zeros(M.size)
    hc.hc_mod_two_pi(M.size, _d(M), _d(r))
    ref = np.array([math.fmod(x, 2 * math.pi) for x in M])
    ref = np.where(ref < 0, ref + 2 * math.pi, ref)
    assert np.array_equal(r, ref)


def test_device_exp_dd_is_correctly_rounded(hc):
    from decimal import Decimal, getcontext
    getcontext().prec = 60
    x = np.random.default_rng(2).uniform(-15.0, 27.0, 3000)
    y = np.zeros(x.size)
    hc.hc_exp_dd(x.size, _d(x), _d(y))
    cr = np.array([float(Decimal(float(v)).exp()) for v in x])
    assert np.array_equal(y, cr)


def test_device_sincos_accuracy(hc):
    import math
    rng = np.random.default_rng(3)
    for which, lo, hi in ((0, -1e4, 1e4), (1, 0.0, 7.0), (2, -0.03, 0.03)):
        x = np.ascontiguousarray(rng.uniform(lo, hi, 20000))
        s, c = np.zeros(x.size), np.zeros(x.size)
        hc.hc_sincos(which, x.size, _d(x), _d(s), _d(c))
        rs = np.array([math.sin(v) for v in x])
        rc = np.array([math.cos(v) for v in x])
        assert np.max(np.abs(s - rs)) <= 3e-16 and np.max(np.abs(c - rc)) <= 3e-16


def test_dnest_post_is_pykimas_postprocess():
    """tests/dnest_post.py against pykima.classic.postprocess (/root/reference/pykima/classic.py:20-135)
    on a committed run of the reference's CPU sampler (tests/golden/make_golden_postprocess.py)."""
    import dnest_post
    run = os.path.join(GOLDEN, "postprocess_run")
    z = np.load(os.path.join(run, "pykima_postprocess.npz"))
    logz, P, H = dnest_post.log_evidence(run)
    assert abs(logz - float(z["logz"])) <= 1e-10 and abs(H - float(z["H"])) <= 1e-10
    lv = np.loadtxt(os.path.join(run, "levels.txt"), ndmin=2)
    si = np.loadtxt(os.path.join(run, "sample_info.txt"), ndmin=2)
    assert np.max(np.abs(dnest_post.postprocess(lv, si)[2] - z["logx"])) <= 1e-12
    assert abs(P.sum() - 1.0) <= 1e-12
