"""Developer check: `world` ranks of the batched GPU sampler (one process per GPU) in one run
directory, MCMC steps/s and the host-side time split of rank 0."""
import json, os, subprocess, sys, time
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import sampler_bench as sb

name = sys.argv[1] if len(sys.argv) > 1 else "CoRoT7"
world = int(sys.argv[2]) if len(sys.argv) > 2 else 2
lanes = int(sys.argv[3]) if len(sys.argv) > 3 else 4096 * world
saves = int(sys.argv[4]) if len(sys.argv) > 4 else 20
d = sb.prepare_dir(name, lanes, 100, saves)
exe = os.path.join(sb.BUILD, "kima_gpu_%s" % name)
cores = os.cpu_count() or 8
procs = []
t0 = time.perf_counter()
for r in range(world):
    env = dict(os.environ, RANK=str(r), WORLD_SIZE=str(world), LOCAL_RANK=str(r), MASTER_PORT="29577",
               KIMA_HOST_THREADS=str(max(1, cores // world)), KIMA_PROFILE="1")
    procs.append(subprocess.Popen([exe, "-t", str(lanes), "-s", "43"], cwd=d, env=env, stdout=subprocess.PIPE,
                                  stderr=subprocess.PIPE, text=True))
outs = [p.communicate(timeout=900) for p in procs]
wall = time.perf_counter() - t0
steps = int(open(os.path.join(d, "sampler_state.txt")).read().split()[-1])
prof = [l for l in outs[0][1].splitlines() if l.startswith("# ")]
print(json.dumps(dict(example=name, ranks=world, lanes=lanes, steps=steps, wall_s=wall, steps_per_s=steps / wall,
                      rank0_profile=prof if prof else outs[0][1][-300:])))
