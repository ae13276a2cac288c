"""The bookkeeping draws of kill_lagging_particles shard over ranks without changing them
(host/BatchedSampler.h::kill_lagging_device; DNest4/code/SamplerImpl.h:543-558): tests/cpp/sharded_draws.cpp
restates the first loop with the reference's engine and distribution types and compares the sharded flags and
engine state with the serial loop's.  The GPU-side check of the real code path is the two-rank byte-equality
test (tests/test_device_proposals.py) run with KIMA_CHECK_SHARDED_KILL=1."""
import os
import shutil
import subprocess

import pytest

HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.fixture(scope="module")
def exe(tmp_path_factory):
    cxx = shutil.which("g++")
    if not cxx:
        pytest.skip("no g++")
    out = tmp_path_factory.mktemp("sharded") / "sharded_draws"
    subprocess.run([cxx, "-O2", "-std=c++11", os.path.join(HERE, "cpp", "sharded_draws.cpp"), "-o", str(out)], check=True)
    return str(out)


@pytest.mark.parametrize("n,ranks,seed", [(1, 2, 1), (7, 8, 2), (1000, 2, 43), (100003, 8, 44), (262144, 4, 45)])
def test_sharded_flags_and_engine_state_equal_the_serial_loop(exe, n, ranks, seed):
    r = subprocess.run([exe, str(n), str(ranks), str(seed)], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, r.stdout + r.stderr
    tag, per_draw, n_bad = r.stdout.split()
    assert tag == "ok"
    assert int(per_draw) == 1          # mt19937_64: one 64-bit output per double
    if n >= 1000:
        assert int(n_bad) > 0          # some particles were flagged: the comparison is not vacuous
