"""End to end through the C++ host side: the reference's own example binaries, CPU sampler vs
batched GPU sampler, MCMC steps per second.

    steps/s = count_mcmc_steps (last field of sampler_state.txt, SamplerImpl.h:610-611) / wall time

Both run the same unmodified examples/<name>/kima_setup.cpp with the same OPTIONS; the CPU
binary with one DNest4 thread per host core, the GPU binary with `lanes` DNest4 threads (= lanes).
"""
import json
import os
import shutil
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
E2E = os.path.join(ROOT, "tests", "golden", "e2e")
BUILD = os.path.join(ROOT, "host", "_build")
DATA = {"51Peg": ["51Peg.rv"], "BL2009": ["BL2009_dataset2.kms.rv"], "CoRoT7": ["corot7.txt"],
        "studentT": ["d1.txt"],
        "multi_instrument": ["HD106252_ELODIE.txt", "HD106252_HET.txt", "HD106252_HJS.txt", "HD106252_Lick.txt"]}

OPTIONS = """# DNest4 options for the throughput run
2	# Number of particles
{nli}	# new level interval
1	# save interval
{ts}	# threadSteps
0	# maximum number of levels
10	# lambda
100	# beta
{saves}	# Maximum number of saves
"""


def _loop_seconds(stderr):
    for l in stderr.splitlines():
        if l.startswith("# sampling loop "):
            return float(l.split()[3])
    return None


def _run_seconds(stderr):
    for l in stderr.splitlines():
        if l.startswith("# run() "):
            return float(l.split()[2])
    return None


def run(kind, name, threads, thread_steps, saves, seed=43, env=None, timeout=1200):
    exe = os.path.join(BUILD, "kima_%s_%s" % (kind, name))
    if not os.path.exists(exe):
        return None
    d = tempfile.mkdtemp(prefix="kima_%s_%s_" % (kind, name))
    try:
        for f in DATA[name]:
            shutil.copy(os.path.join(E2E, f), d)
        with open(os.path.join(d, "OPTIONS"), "w") as fh:
            fh.write(OPTIONS.format(nli=max(2000, threads * thread_steps // 4), ts=thread_steps, saves=saves))
        e = dict(os.environ, KIMA_PROFILE="1")
        e.update(env or {})
        t0 = time.perf_counter()
        r = subprocess.run([exe, "-t", str(threads), "-s", str(seed)], cwd=d, capture_output=True, text=True,
                           timeout=timeout, env=e)
        wall = time.perf_counter() - t0
        if r.returncode != 0:
            return {"error": (r.stdout[-500:] + r.stderr[-500:])}
        with open(os.path.join(d, "sampler_state.txt")) as fh:
            steps = int(fh.read().split()[-1])
        prof = [l for l in r.stderr.splitlines() if l.startswith("# host threads") or l.startswith("# device proposals")]
        extra = {}
        run_s = _run_seconds(r.stderr)
        if run_s:
            # the sampling loop alone (Sampler::run), without process start, CUDA context and
            # the prior draws: what a long run converges to
            extra = {"run_s": run_s, "steps_per_s_run": steps / run_s}
        loop_s = _loop_seconds(r.stderr)
        if loop_s:
            # device proposals: barriers only (steps on the GPU + bookkeeping), without the one-off upload
            # of the particles before and the sampler_state.txt of all particles after
            extra.update({"loop_s": loop_s, "steps_per_s_loop": steps / loop_s})
            extra["one_off"] = [l for l in r.stderr.splitlines() if l.startswith(("# sampling loop", "# after the loop", "# kc_sampler_create"))]
        return {"kind": kind, "example": name, "profile": prof[-1] if prof else None, **extra, "threads": threads, "thread_steps": thread_steps, "saves": saves,
                "mcmc_steps": steps, "wall_s": wall, "steps_per_s": steps / wall}
    finally:
        shutil.rmtree(d, ignore_errors=True)


def prepare_dir(name, threads, thread_steps, saves):
    d = tempfile.mkdtemp(prefix="kima_ranks_%s_" % name)
    for f in DATA[name]:
        shutil.copy(os.path.join(E2E, f), d)
    with open(os.path.join(d, "OPTIONS"), "w") as fh:
        fh.write(OPTIONS.format(nli=max(2000, threads * thread_steps // 4), ts=thread_steps, saves=saves))
    return d


def run_rank(name, d, threads, rank, world, local, port, host_threads, seed=43, timeout=420, env=None):
    """This process's rank of a one-process-per-GPU run of the batched GPU sampler in the shared
    run directory `d` (every rank calls this at the same time)."""
    exe = os.path.join(BUILD, "kima_gpu_%s" % name)
    e = dict(os.environ, KIMA_PROFILE="1", RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(local), MASTER_PORT=str(port),
             KIMA_HOST_THREADS=str(host_threads))
    e.pop("KIMA_CUDA_DEVICE", None)
    e.update(env or {})
    t0 = time.perf_counter()
    try:
        r = subprocess.run([exe, "-t", str(threads), "-s", str(seed)], cwd=d, capture_output=True, text=True,
                           timeout=timeout, env=e)
    except subprocess.TimeoutExpired:
        return {"error": "rank %d timed out after %d s" % (rank, timeout)}
    wall = time.perf_counter() - t0
    if r.returncode != 0:
        return {"error": (r.stdout[-500:] + r.stderr[-500:])}
    steps = None
    if rank == 0:
        with open(os.path.join(d, "sampler_state.txt")) as fh:
            steps = int(fh.read().split()[-1])
    return {"wall_s": wall, "mcmc_steps": steps, "run_s": _run_seconds(r.stderr)}


PARITY_OPTS = "2\n50\n10\n5\n0\n1\n100\n600\n"  # lambda = 1: lagging particles get replaced early and often
PARITY_FILES = ("sample.txt", "sample_info.txt", "levels.txt", "sampler_state.txt")


def chains_equal_check(rank, world, local, port, exch, name="BL2009", lanes=16, seed=43):
    """The parity bit of the multi-GPU sampler leg (every rank calls this): `world` ranks x lanes/world
    lanes must write, byte for byte, the files of one process x `lanes` lanes, with host proposals
    (the reference's mt19937 chains) and with device proposals (per-lane Philox streams).  Rank 0
    runs the single-process twin and compares."""
    import torch.distributed as dist
    exe = os.path.join(BUILD, "kima_gpu_%s" % name)
    out = {"example": name, "lanes": lanes, "ranks": world, "seed": seed, "files": list(PARITY_FILES)}
    for mode, extra in (("host_proposals", {}), ("device_proposals", {"KIMA_DEVICE_PROPOSALS": "1"})):
        box = [None]
        if rank == 0:
            dirs = []
            for tag in ("one", "ranks"):
                d = tempfile.mkdtemp(prefix="kima_parity_%s_%s_" % (mode, tag))
                for f in DATA[name]:
                    shutil.copy(os.path.join(E2E, f), d)
                with open(os.path.join(d, "OPTIONS"), "w") as fh:
                    fh.write(PARITY_OPTS)
                dirs.append(d)
            e = dict(os.environ, KIMA_CUDA_DEVICE=str(local))
            for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK"):
                e.pop(k, None)
            e.update(extra)
            r = subprocess.run([exe, "-t", str(lanes), "-s", str(seed)], cwd=dirs[0], capture_output=True,
                               text=True, timeout=300, env=e)
            if r.returncode != 0:
                out[mode] = {"error": "single process: " + (r.stdout[-300:] + r.stderr[-300:])}
                dirs = None
            box = [dirs]
        dist.broadcast_object_list(box, src=0)
        dirs = box[0]
        if dirs is None:
            continue
        dist.barrier()
        e = dict(os.environ, RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(local),
                 MASTER_PORT=str(port + (1 if extra else 0)))
        e.pop("KIMA_CUDA_DEVICE", None)
        e.update(extra)
        bad = 0.0
        try:
            r = subprocess.run([exe, "-t", str(lanes), "-s", str(seed)], cwd=dirs[1], capture_output=True,
                               text=True, timeout=300, env=e)
            bad = 0.0 if r.returncode == 0 else 1.0
        except subprocess.TimeoutExpired:
            bad = 1.0
        bad = exch.max_float(bad)
        if rank == 0:
            if bad:
                out[mode] = {"error": "a rank failed"}
            else:
                eq = {f: open(os.path.join(dirs[0], f), "rb").read() == open(os.path.join(dirs[1], f), "rb").read()
                      for f in PARITY_FILES}
                rows = sum(1 for _ in open(os.path.join(dirs[1], "sample.txt"))) - 1
                out[mode] = {"chains_equal": all(eq.values()), "per_file": eq, "saved_particles": rows}
            for d in dirs:
                shutil.rmtree(d, ignore_errors=True)
    if rank == 0:
        out["chains_equal"] = all(isinstance(out.get(m), dict) and out[m].get("chains_equal") is True
                                  for m in ("host_proposals", "device_proposals"))
    return out


def compare(name, lanes, cores, gpu_saves=20, cpu_saves=20, thread_steps=100, env=None):
    cpu = run("cpu", name, cores, thread_steps, cpu_saves)
    gpu = run("gpu", name, lanes, thread_steps, gpu_saves, env=env)
    out = {"example": name, "cpu": cpu, "gpu": gpu}
    if cpu and gpu and "steps_per_s" in cpu and "steps_per_s" in gpu:
        out["speedup"] = gpu["steps_per_s"] / cpu["steps_per_s"]
    return out


if __name__ == "__main__":
    names = sys.argv[1].split(",") if len(sys.argv) > 1 else ["CoRoT7", "51Peg"]
    lanes = int(sys.argv[2]) if len(sys.argv) > 2 else 4096
    cores = os.cpu_count() or 8
    os.environ["KIMA_PROFILE"] = "1"
    for n in names:
        print(json.dumps(compare(n, lanes, cores, gpu_saves=40, cpu_saves=100)), flush=True)
