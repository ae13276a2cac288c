"""CPU: the lock-step sampler itself (host/BatchedSampler.h, host/LanePool.h, host/LanePriors.h), without a GPU.

tests/cpp/RefEvaluator.h is a test-only Evaluator that scores the staged particles with the reference's own
RVmodel::log_likelihood(); built with it, the UNMODIFIED examples/*/kima_setup.cpp run through the batched
sampler on the host cores and must write, byte for byte, the files of the reference's sampler
(DNest4/code/SamplerImpl.h) started with the same -t and -s.  What this covers that the GPU tests cannot reach
on this machine: update_particle split around the likelihood call, the two lane groups, lanes on pool threads,
the level bookkeeping, and the per-lane hyperprior objects.  Needs /root/reference (build container only)."""
import os
import shutil
import subprocess
from concurrent.futures import ThreadPoolExecutor

import pytest

from conftest import ROOT

REF = "/root/reference"
OBJ = os.path.join(ROOT, "oracle", "_ref", "obj")
HOST_BUILD = os.path.join(ROOT, "host", "_build")
OUT = os.path.join(ROOT, "tests", "cpp", "_build")
E2E = os.path.join(ROOT, "tests", "golden", "e2e")
DATA = {"BL2009": ["BL2009_dataset2.kms.rv"], "CoRoT7": ["corot7.txt"], "many_planets": ["HD10180.kms.rv"],
        "multi_instrument": ["HD106252_ELODIE.txt", "HD106252_HET.txt", "HD106252_HJS.txt", "HD106252_Lick.txt"]}
FILES = ("sample.txt", "sample_info.txt", "levels.txt", "sampler_state.txt")
CI_OPTIONS = "2\n50\n10\n5\n0\n10\n100\n100\n"   # the reference's CI options (tests/golden/e2e/OPTIONS)
INC = ["-I" + os.path.join(REF, d) for d in ("src", "DNest4/code", "eigen", "celerite/cpp/include")]
SKIP_OBJS = ("ref_harness.o", "ref_flags.o", "philox_rng.o", "ref_sampler_harness.o")


def _ref_objs():
    return sorted(os.path.join(OBJ, f) for f in os.listdir(OBJ) if f.endswith(".o") and f not in SKIP_OBJS)


def _compile(args):
    r = subprocess.run(args, capture_output=True, text=True)
    assert r.returncode == 0, " ".join(args[:6]) + " ...\n" + r.stderr[-3000:]


@pytest.fixture(scope="module")
def built():
    cxx = shutil.which("g++")
    if not cxx or not os.path.isdir(REF) or not os.path.isdir(OBJ):
        pytest.skip("needs g++, /root/reference and the reference objects (oracle/Makefile)")
    for n in DATA:
        if not os.path.exists(os.path.join(HOST_BUILD, "kima_cpu_%s" % n)):
            pytest.skip("host/_build/kima_cpu_%s is missing (make -C host)" % n)
    os.makedirs(OUT, exist_ok=True)
    flags = [cxx, "-pthread", "-std=c++11", "-w", "-DEIGEN_MPL2_ONLY"]
    jobs = [flags + ["-O1", "-I" + os.path.join(ROOT, "host")] + INC +
            ["-o", os.path.join(OUT, "lane_priors_check"), os.path.join(ROOT, "tests", "cpp", "lane_priors_check.cpp")] +
            _ref_objs()]
    for n in DATA:
        # host/overlay FIRST: its kima.h stands in for the reference's src/kima.h (as in host/Makefile)
        jobs.append(flags + ["-O2", "-DNDEBUG", '-DKIMA_B200_EVALUATOR_HEADER="RefEvaluator.h"',
                             "-DKIMA_B200_EVALUATOR=RefEvaluator", "-I" + os.path.join(ROOT, "host", "overlay"),
                             "-I" + os.path.join(ROOT, "tests", "cpp"), "-I" + os.path.join(ROOT, "host"),
                             "-I" + os.path.join(ROOT, "include")] + INC +
                            ["-o", os.path.join(OUT, "kima_batched_cpu_%s" % n),
                             os.path.join(REF, "examples", n, "kima_setup.cpp")] + _ref_objs())
    with ThreadPoolExecutor(min(5, os.cpu_count() or 2)) as ex:
        list(ex.map(_compile, jobs))
    return OUT


def _run(exe, name, d, lanes, options, env=None, seed=43, one_core=False):
    """one_core: pin the process to one core.  The REFERENCE needs that to be reproducible with hyperpriors: its
    threads share the prior objects the hyperprior code re-parameterises before every use (host/LanePriors.h), and
    on several cores examples/CoRoT7 -t 8 writes a different chain every time (profiles/host_repro_r02.txt).  On
    one core its threads never run at the same time and the chain is the serial one."""
    d.mkdir()
    for f in DATA[name]:
        shutil.copy(os.path.join(E2E, f), d)
    (d / "OPTIONS").write_text(options)
    cmd = [exe, "-t", str(lanes), "-s", str(seed)]
    if one_core and shutil.which("taskset"):
        cmd = ["taskset", "-c", str(sorted(os.sched_getaffinity(0))[0])] + cmd
    r = subprocess.run(cmd, cwd=d, capture_output=True, text=True, timeout=900,
                       env=dict(os.environ, **(env or {})))
    assert r.returncode == 0, r.stdout[-1500:] + r.stderr[-1500:]
    return d


def test_the_binary_under_test_is_the_batched_sampler(built, tmp_path):
    d = tmp_path / "probe"
    d.mkdir()
    shutil.copy(os.path.join(E2E, DATA["BL2009"][0]), d)
    (d / "OPTIONS").write_text(CI_OPTIONS)
    r = subprocess.run([os.path.join(built, "kima_batched_cpu_BL2009"), "-t", "4", "-s", "43"], cwd=d, capture_output=True,
                       text=True, timeout=300, env=dict(os.environ, KIMA_PROFILE="1", KIMA_HOST_THREADS="2"))
    assert r.returncode == 0 and "# host threads 2" in r.stderr, r.stderr[-1500:]   # BatchedSampler::run's profile line


def _same(a, b, tag):
    for f in FILES:
        assert (a / f).read_bytes() == (b / f).read_bytes(), (tag, f)


def test_lane_priors_against_the_reference_classes(built):
    r = subprocess.run([os.path.join(built, "lane_priors_check"), os.path.join(E2E, "BL2009_dataset2.kms.rv")],
                       capture_output=True, text=True, timeout=60)
    assert r.returncode == 0 and r.stdout.strip().endswith("ok"), r.stdout + r.stderr


@pytest.mark.parametrize("name", sorted(DATA))
def test_batched_sampler_writes_the_reference_chain(built, tmp_path, name):
    """-t 4 -s 43: 4 lanes in two groups, every step's proposals scored in one call per group."""
    ref = _run(os.path.join(HOST_BUILD, "kima_cpu_%s" % name), name, tmp_path / "ref", 4, CI_OPTIONS, one_core=True)
    bat = _run(os.path.join(built, "kima_batched_cpu_%s" % name), name, tmp_path / "bat", 4, CI_OPTIONS,
               {"KIMA_HOST_THREADS": "2"})
    _same(ref, bat, name)


def test_lanes_on_pool_threads_write_the_reference_chain(built, tmp_path):
    """32 lanes on 4 pool threads (the pool only splits ranges of at least two lanes per thread), hyperpriors
    and up to 10 planets, enough steps for levels to be created and lagging particles replaced across lanes."""
    name = "many_planets"
    opts = "2\n2000\n1600\n50\n0\n10\n100\n60\n"
    ref = _run(os.path.join(HOST_BUILD, "kima_cpu_%s" % name), name, tmp_path / "ref", 32, opts, one_core=True)
    bat = _run(os.path.join(built, "kima_batched_cpu_%s" % name), name, tmp_path / "bat", 32, opts,
               {"KIMA_HOST_THREADS": "4"})
    _same(ref, bat, "32 lanes")
    one = _run(os.path.join(built, "kima_batched_cpu_%s" % name), name, tmp_path / "one", 32, opts,
               {"KIMA_HOST_THREADS": "1"})
    _same(bat, one, "1 thread against 4")


def _run_ranks(exe, name, d, world, lanes, options, port, seed=43, extra_env=None):
    """`world` processes in ONE run directory, as a torchrun-style launcher starts the GPU ranks
    (RANK / WORLD_SIZE / MASTER_PORT); tests/cpp/RefEvaluator.h carries the collectives over TCP on 127.0.0.1."""
    d.mkdir()
    for f in DATA[name]:
        shutil.copy(os.path.join(E2E, f), d)
    (d / "OPTIONS").write_text(options)
    procs = []
    for r in range(world):
        # KIMA_CHECK_SHARDED_KILL: every rank also runs the serial flagging loop on a copy of the engine and compares
        env = dict(os.environ, RANK=str(r), WORLD_SIZE=str(world), MASTER_PORT=str(port), KIMA_HOST_THREADS="2",
                   KIMA_CHECK_SHARDED_KILL="1")
        env.update(extra_env or {})
        procs.append(subprocess.Popen([exe, "-t", str(lanes), "-s", str(seed)], cwd=d, env=env, stdout=subprocess.PIPE,
                                      stderr=subprocess.PIPE, text=True))
    outs = [p.communicate(timeout=600) for p in procs]
    for p, (so, se) in zip(procs, outs):
        assert p.returncode == 0, so[-1500:] + se[-1500:]
    return d, outs[0][0]


@pytest.mark.parametrize("name,world,port", [("BL2009", 2, 29621), ("BL2009", 4, 29622), ("CoRoT7", 2, 29623),
                                             ("multi_instrument", 2, 29624)])
def test_ranks_write_the_single_process_chain(built, tmp_path, name, world, port):
    """One process per rank, lanes sharded, BatchedSampler::exchange at every barrier (likelihoods, level
    assignments and rank 0's bookkeeping RNG to everybody; `above` lists concatenated in lane order; level counters
    summed), particles fetched from their owners where the bookkeeping reads one (bookkeeping_ranks): `world` ranks
    write the files of one process with the same -t, which are the reference sampler's.  The multi-GPU runs use exactly this code with NCCL underneath
    (tests/test_host_cpp.py::test_two_ranks_reproduce_the_single_process_chain, on 2 GPUs)."""
    opts = "2\n50\n10\n5\n0\n1\n100\n300\n"   # lambda = 1: lagging particles get replaced early and often
    exe = os.path.join(built, "kima_batched_cpu_%s" % name)
    ref = _run(os.path.join(HOST_BUILD, "kima_cpu_%s" % name), name, tmp_path / "ref", 8, opts, one_core=True)
    one = _run(exe, name, tmp_path / "one", 8, opts, {"KIMA_HOST_THREADS": "2"})
    ranks, out0 = _run_ranks(exe, name, tmp_path / "ranks", world, 8, opts, port)
    assert "Replacing lagging particle" in out0     # copies between lanes (and ranks) happened
    _same(ref, one, "reference against one process")
    _same(one, ranks, "%d ranks against one process" % world)
    if name == "BL2009" and world == 2:
        # the first version of the barrier (every particle to everybody, the inherited do_bookkeeping): same files
        full, _ = _run_ranks(exe, name, tmp_path / "full", world, 8, opts, port + 50, extra_env={"KIMA_FULL_EXCHANGE": "1"})
        _same(one, full, "2 ranks, full exchange")


def test_the_reference_itself_is_not_reproducible_with_hyperpriors_on_several_cores(built, tmp_path):
    """Not a requirement on this repository: a record of why the comparisons above pin the reference to one core.
    examples/CoRoT7 (hyperpriors) with -t 8 and frequent copies between threads: the batched sampler writes the
    same files on 1 and on 3 host threads; the reference, free to use all cores, usually does not repeat itself."""
    if len(os.sched_getaffinity(0)) < 4:
        pytest.skip("needs several cores")
    opts = "2\n50\n10\n5\n0\n1\n100\n300\n"
    exe = os.path.join(built, "kima_batched_cpu_CoRoT7")
    a = _run(exe, "CoRoT7", tmp_path / "b1", 8, opts, {"KIMA_HOST_THREADS": "1"})
    b = _run(exe, "CoRoT7", tmp_path / "b3", 8, opts, {"KIMA_HOST_THREADS": "3"})
    _same(a, b, "batched, 1 thread against 3")
    refexe = os.path.join(HOST_BUILD, "kima_cpu_CoRoT7")
    states = {_run(refexe, "CoRoT7", tmp_path / ("r%d" % i), 8, opts).joinpath("sampler_state.txt").read_bytes()
              for i in range(4)}
    if len(states) == 1:
        pytest.skip("the reference repeated itself four times on this machine")
    assert len(states) > 1
