"""Copy what one tools/gpu_session_e.sh run left under gpurun_out/<tag>/ into profiles/ (tracked):
the bench lines, the tail of the GPU test log, the smoke log, the sampler launch list and its summary,
the device-sampler timings and the ncu summary of the step kernel."""
import collections
import csv
import json
import os
import shutil
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def last_line(path):
    with open(path) as fh:
        lines = [l for l in fh.read().splitlines() if l.strip()]
    return lines[-1] + "\n"


def launch_summary(csv_path, out_path, header):
    rows = [l for l in open(csv_path) if l.startswith('"')]
    acc = collections.OrderedDict()
    for x in csv.DictReader(rows):
        if x.get("Metric Name") != "gpu__time_duration.sum":
            continue
        k = x["Kernel Name"].split("(")[0].replace("<unnamed>::", "")
        v = float(x["Metric Value"].replace(",", ""))
        if x["Metric Unit"] in ("ns", "nsecond"):
            v /= 1000.0
        a = acc.setdefault(k, [0, 0.0])
        a[0] += 1
        a[1] += v
    tot = sum(a[1] for a in acc.values())
    out = list(header)
    for k, a in sorted(acc.items(), key=lambda kv: -kv[1][1]):
        out.append("%-46s n=%3d total %9.1f us  mean %8.1f us  share %5.1f%%" % (k[:46], a[0], a[1], a[1] / a[0], 100 * a[1] / tot))
    with open(out_path, "w") as fh:
        fh.write("\n".join(out) + "\n")
    return out


def main():
    tag = sys.argv[1] if len(sys.argv) > 1 else "r01e"
    src = os.path.join(ROOT, "gpurun_out", tag)
    dst = os.path.join(ROOT, "profiles")
    open(os.path.join(dst, "bench_%s.json" % tag), "w").write(last_line(os.path.join(src, "bench.json")))
    open(os.path.join(dst, "bench_ref_%s.json" % tag), "w").write(last_line(os.path.join(src, "bench_ref.json")))
    with open(os.path.join(src, "pytest_gpu.log")) as fh:
        open(os.path.join(dst, "pytest_gpu_%s.txt" % tag), "w").write("".join(fh.readlines()[-25:]))
    with open(os.path.join(src, "smoke.log")) as fh:
        open(os.path.join(dst, "smoke_%s.txt" % tag), "w").write("".join(fh.readlines()[-9:]))
    for a, b in (("launches_sampler.csv", "launches_sampler_%s.csv"), ("devprop_e2e.jsonl", "devprop_e2e_%s.jsonl"),
                 ("dev_sampler_timings.jsonl", "dev_sampler_timings_%s.jsonl")):
        if os.path.exists(os.path.join(src, a)):
            shutil.copy(os.path.join(src, a), os.path.join(dst, b % tag))
    out = launch_summary(os.path.join(dst, "launches_sampler_%s.csv" % tag),
                         os.path.join(dst, "launches_sampler_%s_summary.txt" % tag),
                         ["ncu --metrics gpu__time_duration.sum --clock-control none, python tools/dev_sampler_prof.py 65536 20 1",
                          "(51Peg shape: N=256, Np<=1, 65536 lanes x 1 particle, 20 lock-step MCMC steps; cold-cache serialised launches)"])
    print("\n".join(out[:6]))
    rep = os.path.join(src, "smp_full.ncu-rep")
    if os.path.exists(rep):
        r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "ncu_summary.py"), rep], capture_output=True, text=True)
        if r.returncode == 0:
            open(os.path.join(dst, "ncu_smp_%s.json" % tag), "w").write(r.stdout)
    d = json.loads(last_line(os.path.join(src, "bench.json")))
    print("kep %.4g evals/s (frac %.3f), e2e %.4g; gp %.4g (frac %.3f)" % (d["value"], d["roofline"]["frac"], d["e2e"]["value"],
                                                                      d["gp"]["value"], d["gp"]["roofline"]["frac"]))
    s = d.get("sampler", {})
    for r in s.get("device_proposals", []):
        print("device proposals", r["example"], r["threads"], "run %.4g steps/s" % r["steps_per_s_run"], "loop %.4g" % r.get("steps_per_s_loop", 0))
    for r in s.get("runs", []):
        print("host proposals", r["example"], "cpu %.4g" % r["cpu"]["steps_per_s"], "gpu %.4g" % r["gpu"]["steps_per_s_run"])
    if "device_steps" in s:
        print("device steps %.4g steps/s, %.4f ms/step" % (s["device_steps"]["value"], s["device_steps"]["ms_per_step"]))


if __name__ == "__main__":
    main()
