"""Developer check on a GPU box: the CoRoT7 example binary with and without the factor cache, host and device
proposals: MCMC steps/s over run(), and the output files byte for byte (chains must not depend on the cache)."""
import filecmp, json, os, shutil, subprocess, sys, tempfile
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import sampler_bench as sb

FILES = ("sample.txt", "sample_info.txt", "levels.txt", "sampler_state.txt")


def one(name, threads, ts, saves, env):
    exe = os.path.join(sb.BUILD, "kima_gpu_%s" % name)
    d = tempfile.mkdtemp(prefix="kima_cache_")
    for f in sb.DATA[name]:
        shutil.copy(os.path.join(sb.E2E, f), d)
    with open(os.path.join(d, "OPTIONS"), "w") as fh:
        fh.write(sb.OPTIONS.format(nli=max(2000, threads * ts // 4), ts=ts, saves=saves))
    e = dict(os.environ, KIMA_PROFILE="1")
    e.update(env)
    r = subprocess.run([exe, "-t", str(threads), "-s", "43"], cwd=d, capture_output=True, text=True, timeout=900, env=e)
    if r.returncode != 0:
        return d, {"error": r.stderr[-800:]}
    for ln in r.stderr.split("\n"):
        if "MISMATCH" in ln:
            print(ln[:300], flush=True)
    steps = int(open(os.path.join(d, "sampler_state.txt")).read().split()[-1])
    run_s = sb._run_seconds(r.stderr)
    return d, {"steps": steps, "run_s": run_s, "steps_per_s_run": steps / run_s if run_s else None}


if __name__ == "__main__":
    for mode, env, threads, ts, saves in (("host_proposals", {}, 4096, 100, 40),
                                          ("device_proposals", {"KIMA_DEVICE_PROPOSALS": "1"}, 16384, 100, 10)):
        res = {}
        dirs = {}
        for tag, cache in (("cache_off", "0"), ("cache_on", "1")):
            dirs[tag], res[tag] = one("CoRoT7", threads, ts, saves, dict(env, KIMA_CUDA_GP_CACHE=cache))
        same = all(os.path.exists(os.path.join(dirs["cache_on"], f)) and
                   filecmp.cmp(os.path.join(dirs["cache_off"], f), os.path.join(dirs["cache_on"], f), shallow=False) for f in FILES)
        sp = None
        if "error" not in res["cache_on"] and "error" not in res["cache_off"]:
            sp = res["cache_on"]["steps_per_s_run"] / res["cache_off"]["steps_per_s_run"]
        print(json.dumps(dict(example="CoRoT7", mode=mode, lanes=threads, thread_steps=ts, saves=saves, files_equal=same, speedup=sp, **res)), flush=True)
        for d in dirs.values():
            shutil.rmtree(d, ignore_errors=True)
