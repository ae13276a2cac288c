"""GPU, end to end: the reference's unmodified examples with KIMA_DEVICE_PROPOSALS=1 — particles in
HBM, proposals / prior-ratio test / likelihood / accept / level counters on the GPU
(include/kima_sampler.h), DNest4's bookkeeping on the host (host/BatchedSampler.h, run_device).

The lanes draw from counter-based streams, so these chains are not the reference's mt19937
chains (tests/test_host_cpp.py pins the default, byte-identical path).  What must hold instead
(BASELINE north_star: "full runs must also reproduce log Z and the Np posterior within sampling
error"): the run is reproducible, its files have the reference's format, and its evidence,
information and best likelihood agree with the reference's CPU sampler on the same OPTIONS
within the scatter the reference shows between two of its own seeds.
"""
import os
import shutil
import subprocess

import numpy as np
import pytest

from dnest_post import log_evidence
from test_host_cpp import BUILD, CASES, E2E, load_sample

pytestmark = pytest.mark.gpu

# particles, new level interval, save interval, thread steps, max levels (0 = automatic),
# lambda, beta, max saves: 2M MCMC steps
OPTIONS = "2\n1000\n500\n100\n0\n10\n100\n4000\n"


def run(kind, name, tmp_path, tag, lanes, seed, device, options=OPTIONS, extra_env=None):
    exe = os.path.join(BUILD, "kima_%s_%s" % (kind, name))
    if not os.path.exists(exe):
        pytest.fail("%s is missing: build it here with `make -C host` (needs /root/reference)" % exe)
    d = tmp_path / tag
    d.mkdir()
    for f in CASES[name][0]:
        shutil.copy(os.path.join(E2E, f), d)
    (d / "OPTIONS").write_text(options)
    env = dict(os.environ)
    env.pop("KIMA_DEVICE_PROPOSALS", None)
    if device:
        env["KIMA_DEVICE_PROPOSALS"] = "1"
    env.update(extra_env or {})
    r = subprocess.run([exe, "-t", str(lanes), "-s", str(seed)], cwd=d, capture_output=True, text=True,
                       timeout=900, env=env)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    return d


def test_device_run_is_reproducible_and_keeps_the_file_formats(gpu, tmp_path):
    small = "2\n200\n100\n50\n0\n10\n100\n300\n"
    a = run("gpu", "BL2009", tmp_path, "a", 16, 43, True, small)
    b = run("gpu", "BL2009", tmp_path, "b", 16, 43, True, small)
    h = run("gpu", "BL2009", tmp_path, "h", 16, 43, False, small)
    for f in ("sample.txt", "sample_info.txt", "levels.txt"):
        assert (a / f).read_bytes() == (b / f).read_bytes(), f
    # all_above ordered and cut in HBM (the default in one process) or gathered and selected on the host:
    # the same levels, hence the same chain
    c = run("gpu", "BL2009", tmp_path, "c", 16, 43, True, small, {"KIMA_HOST_ABOVE": "1"})
    for f in ("sample.txt", "sample_info.txt", "levels.txt", "sampler_state.txt"):
        assert (a / f).read_bytes() == (c / f).read_bytes(), f
    ha, ra = load_sample(a / "sample.txt")
    hh, rh = load_sample(h / "sample.txt")
    assert ha == hh and ra.shape == rh.shape       # same columns, same number of saves
    assert not np.array_equal(ra, rh)              # ... of a different chain
    ia = np.loadtxt(a / "sample_info.txt", ndmin=2)
    assert ia.shape[1] == 4 and np.all(np.isfinite(ia[:, 1]))
    # a saved particle's columns are a consistent RVmodel: Np planets, the others printed as 0
    cols = ha.lstrip("# ").split()
    i_np, i_p0 = cols.index("Np"), cols.index("P0")
    assert set(np.unique(ra[:, i_np])) <= {0.0, 1.0}
    assert np.all((ra[:, i_p0] > 0) == (ra[:, i_np] == 1))
    st = (a / "sampler_state.txt").read_text().split()
    assert int(st[-1]) == 300 * 16 * 50  # count_mcmc_steps: one save per barrier of lanes x thread steps


def test_many_lanes_sorted_kernels_and_chunked_state_file_change_no_byte(gpu, tmp_path):
    """8192 lanes x 2 particles: the proposal kernels sort each CTA's lanes by move (the default from 2048
    lanes on) and sampler_state.txt is formatted in chunks by the host threads (BatchedSampler::print_state).
    Against the same run with one warp per CTA and no sorting (KC_SMP_BLOCK=0) and the state file through
    Sampler::print (SamplerImpl.h:590-612, KIMA_SERIAL_STATE=1): every output file equal byte for byte."""
    opts = "2\n20000\n10000\n20\n0\n10\n100\n20\n"
    a = run("gpu", "51Peg", tmp_path, "sorted", 8192, 43, True, opts)
    b = run("gpu", "51Peg", tmp_path, "plain", 8192, 43, True, opts, {"KIMA_SERIAL_STATE": "1", "KC_SMP_BLOCK": "0"})
    sa, sb = (a / "sampler_state.txt").read_bytes(), (b / "sampler_state.txt").read_bytes()
    assert len(sa) > 16384 * 40
    assert sa == sb
    for f in ("sample.txt", "sample_info.txt", "levels.txt"):
        assert (a / f).read_bytes() == (b / f).read_bytes(), f
    assert len(load_sample(a / "sample.txt")[1]) >= 15


def test_evidence_matches_the_reference_sampler(gpu, tmp_path):
    """log Z, information H, the best likelihood found and the posterior of Np: device proposals
    vs the reference's CPU sampler at the same -t and OPTIONS (two seeds of the reference give
    the scatter: 0.3 in log Z at this run length)."""
    name, lanes = "BL2009", 8
    ref1 = run("cpu", name, tmp_path, "ref1", lanes, 43, False)
    ref2 = run("cpu", name, tmp_path, "ref2", lanes, 99, False)
    dev = run("gpu", name, tmp_path, "dev", lanes, 43, True)
    z1, _, H1 = log_evidence(ref1)
    z2, _, H2 = log_evidence(ref2)
    zd, wd, Hd = log_evidence(dev)
    zr = 0.5 * (z1 + z2)
    tol = max(1.5, 4.0 * abs(z1 - z2))
    assert abs(zd - zr) < tol, (zd, z1, z2)
    assert abs(Hd - 0.5 * (H1 + H2)) < max(5.0, 4.0 * abs(H1 - H2)), (Hd, H1, H2)
    best = [np.loadtxt(d / "sample_info.txt", ndmin=2)[:, 1].max() for d in (ref1, ref2, dev)]
    assert abs(best[2] - max(best[:2])) < 1.0, best
    # posterior of the number of planets: BL2009 has one clear signal
    hdr, rows = load_sample(dev / "sample.txt")
    i_np = hdr.lstrip("# ").split().index("Np")
    assert np.sum(wd * (rows[:, i_np] == 1)) > 0.95


def test_evidence_at_many_lanes_matches_host_proposals(gpu, tmp_path):
    """256 lanes: against the host-proposal GPU path at the same -t (which is the reference's chain,
    tests/test_host_cpp.py; the CPU binary would need hours for these 38M steps per run).  A single
    DNest4 run of this length scatters by about 1.5 in log Z on either path (one saved particle that
    the beta term has just pushed many levels down can carry the whole sum): means over 3 seeds against
    their own standard error."""
    opts = "2\n1000\n500\n100\n0\n10\n100\n1500\n"
    z = {True: [], False: []}
    for seed in (7, 8, 9):
        for device in (True, False):
            d = run("gpu", "BL2009", tmp_path, "%s%d" % ("dev" if device else "host", seed), 256, seed, device, opts)
            z[device].append(log_evidence(d)[0])
    dev, host = np.array(z[True]), np.array(z[False])
    # (r01e on a B200: device -149.6 -146.9 -144.0, host -149.8 -148.4 -151.3)
    se = np.sqrt(dev.var(ddof=1) / 3 + host.var(ddof=1) / 3)
    assert abs(dev.mean() - host.mean()) < 1.0 + 3.0 * se, z
    assert abs(dev.mean() - (-146.0)) < 5.0, z   # the reference's CPU runs give -143.6 ... -147


def test_evidence_corot7_gp_matches_host_proposals(gpu, tmp_path):
    """The GP example (CoRoT7: N=177, hyperpriors, Np <= 5) with device proposals against the
    host-proposal GPU path at the same -t, which is the reference's own chain
    (tests/test_host_cpp.py): log Z and the information over two seeds each."""
    opts = "2\n1000\n500\n100\n0\n10\n100\n600\n"
    z, H = {True: [], False: []}, {True: [], False: []}
    for seed in (7, 8):
        for device in (True, False):
            d = run("gpu", "CoRoT7", tmp_path, "%s%d" % ("dev" if device else "host", seed), 128, seed, device, opts)
            zz, _, hh = log_evidence(d)
            z[device].append(zz)
            H[device].append(hh)
    print("CoRoT7 log Z device %s host %s; H device %s host %s" % (z[True], z[False], H[True], H[False]))
    dev, host = np.array(z[True]), np.array(z[False])
    spread = max(np.ptp(dev), np.ptp(host))
    assert abs(dev.mean() - host.mean()) < 2.0 + 2.0 * spread, z
    assert abs(np.mean(H[True]) - np.mean(H[False])) < 5.0 + 2.0 * max(np.ptp(H[True]), np.ptp(H[False])), H


@pytest.mark.parametrize("host_above", ["0", "1"])
def test_two_ranks_reproduce_the_single_process_device_chain(gpu, tmp_path, host_above):
    """One process per GPU with device proposals: each rank keeps its lanes' particles in its own
    HBM, lane k draws from stream seed + k whoever steps it, and the barrier (counters summed,
    `above` lists concatenated, likelihoods / level assignments gathered; a lagging particle's
    replacement travels as one column) leaves every rank with the state a single process has:
    2 ranks x 8 lanes write, byte for byte, the files of 1 process x 16 lanes — with all_above replicated in
    the GPUs' HBM (lists and counters exchanged device to device by kc_sampler_collect_pool) and with
    KIMA_HOST_ABOVE=1 (lists meeting on the hosts)."""
    from test_host_cpp import _gpu_count
    if _gpu_count() < 2:
        pytest.skip("needs 2 GPUs")
    opts = "2\n50\n10\n5\n0\n1\n100\n600\n"  # lambda = 1: lagging particles get replaced early and often
    one = run("gpu", "BL2009", tmp_path, "one", 16, 43, True, opts)
    exe = os.path.join(BUILD, "kima_gpu_BL2009")
    d = tmp_path / "two"
    d.mkdir()
    for f in CASES["BL2009"][0]:
        shutil.copy(os.path.join(E2E, f), d)
    (d / "OPTIONS").write_text(opts)
    procs = []
    for r in range(2):
        env = dict(os.environ, RANK=str(r), WORLD_SIZE="2", LOCAL_RANK=str(r), MASTER_PORT=str(29577 + int(host_above)),
                   KIMA_DEVICE_PROPOSALS="1", KIMA_HOST_ABOVE=host_above)
        procs.append(subprocess.Popen([exe, "-t", "16", "-s", "43"], cwd=d, env=env, stdout=subprocess.PIPE,
                                      stderr=subprocess.PIPE, text=True))
    outs = [p.communicate(timeout=600) for p in procs]
    for p, (so, se) in zip(procs, outs):
        assert p.returncode == 0, so[-1500:] + se[-1500:]
    assert "Replacing lagging particle" in outs[0][0]  # replacements happened (half of them cross ranks on average)
    for f in ("sample.txt", "sample_info.txt", "levels.txt", "sampler_state.txt"):
        assert (one / f).read_bytes() == (d / f).read_bytes(), f


@pytest.mark.parametrize("name", sorted(CASES))
def test_every_shipped_example_runs_with_device_proposals(gpu, tmp_path, name):
    """The reference's CI set-up (tests/test_suite/OPTIONS, -t 4 -s 43) of every example — GP with
    hyperpriors, Student-t, several instruments, free and fixed Np — through run_device: no fallback to
    host proposals, the reference's columns, finite likelihoods, levels being built."""
    ci = open(os.path.join(E2E, "OPTIONS")).read()
    exe = os.path.join(BUILD, "kima_gpu_%s" % name)
    d = tmp_path / name
    d.mkdir()
    for f in CASES[name][0]:
        shutil.copy(os.path.join(E2E, f), d)
    (d / "OPTIONS").write_text(ci)
    env = dict(os.environ, KIMA_DEVICE_PROPOSALS="1", KIMA_PROFILE="1")
    r = subprocess.run([exe, "-t", "4", "-s", "43"], cwd=d, capture_output=True, text=True, timeout=600, env=env)
    assert r.returncode == 0, r.stdout[-1500:] + r.stderr[-1500:]
    assert "(device proposals)" in r.stderr and "proposals stay on the host" not in r.stderr
    hdr, rows = load_sample(d / "sample.txt")
    hg, rg = load_sample(os.path.join(E2E, "%s_sample_seed43.txt" % name))
    assert hdr == hg and rows.shape == rg.shape
    info = np.loadtxt(d / "sample_info.txt", ndmin=2)
    assert np.all(np.isfinite(info[:, 1])) and np.all(info[:, 1] > -1e299)
    levels = np.loadtxt(d / "levels.txt", ndmin=2)
    assert len(levels) >= 3 and np.all(np.diff(levels[1:, 1]) > 0)
