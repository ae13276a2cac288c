#!/usr/bin/env python
"""bench.py — log-likelihood evaluations per second on the BASELINE.json workloads.

    python bench.py --gpus N --steps K --warmup W            # the CUDA path
    python bench.py --impl reference --gpus N --steps K ...  # the reference's CPU path

Workloads (BASELINE.json `metric` and `configs`):
  kep     Keplerian, N=1000 epochs, Np=10 planets in every lane, Gaussian white noise  (headline K)
  gp      quasi-periodic GP (dense Cholesky), N=1000 epochs, Np=10                     (headline G)
  gp4096  the scale point: QP-GP, N=4096, Np=10, 1024 lanes per GPU (8192 over 8 GPUs, configs[4])
plus the same three with Np ~ U{0..10} (`np_mix`) and the four example shapes (`shapes`).
A "step" scores one batch of lanes (particles) once, with one call of the library; batches are
sized so that the timed region of every workload lasts about a second or more.  The top-level
JSON keys are workload kep (value, e2e, ms_per_step); workload gp sits in `value_gp` / `e2e_gp` /
`ms_per_step_gp`, the scale point in `gp4096`; `roofline` is the workload FURTHEST from its
roofline (with all of them under `roofline.workloads`).  One rank per GPU; lanes shard with no
data-path collective (weak scaling: lanes per GPU fixed).

Timing: per step, the L2 is flushed (untimed), then the step is timed on the device with CUDA
events on the stream the kernels run on (kc_last_kernel_ms); the K steps are bracketed by a
barrier + device synchronise and the max over ranks is taken.  `e2e` times the public call with
host buffers (H2D + kernels + D2H) by wall clock.
"""
from __future__ import annotations

import argparse
import json
import os
import platform
import statistics
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

from kima_b200.flags import ModelFlags  # noqa: E402
from kima_b200.synthetic import make_data, make_params  # noqa: E402

METRIC = "log-likelihood evals/sec: Keplerian (N=1k, Np=10) and QP-GP Cholesky (N=1k)"
NP = 10


def flops_keplerian(N, Np, studentt=False):
    """F_K = N (330 Np + 45) (Student-t: 87 per residual term)."""
    return N * (330.0 * Np + (87.0 if studentt else 45.0))


def flops_gp(N, Np):
    """F_G = N^3/3 + 37 N^2 + F_K."""
    return N ** 3 / 3.0 + 37.0 * N * N + flops_keplerian(N, Np)


# name -> what one step is.  lanes = lanes per GPU and step; in_flight = lanes the library has in
# flight at once (GP: factor-storage slots; the rest of the step follows in waves)
WORKLOADS = {
    "kep": dict(N=1000, flags=ModelFlags(npmax=NP), lanes=64 * 8192, ref_lanes=4096, max_steps=None,
                text="Keplerian N=1000 Np=10 Gaussian white noise, 524288 lanes/GPU per step (64 distinct "
                     "parameter blocks of 8192 lanes, 218 MB of parameters: larger than L2)"),
    "gp": dict(N=1000, flags=ModelFlags(GP=True, npmax=NP), lanes=3 * 1184, gp_slots=1184, ref_lanes=48,
               max_steps=None,
               text="QP-GP dense Cholesky N=1000 Np=10, 3552 lanes/GPU per step as 3 waves of 1184 lanes in "
                    "flight (8 per SM, 9.5 GB of factor storage: larger than L2)"),
    "gp4096": dict(N=4096, flags=ModelFlags(GP=True, npmax=NP), lanes=1024, ref_lanes=None, max_steps=3,
                   text="scale point (BASELINE.json configs[4]): QP-GP N=4096 Np=10, 1024 lanes/GPU per step "
                        "(8192 over 8 GPUs), 134 MB of factor storage per lane in flight, waves within 64 GB"),
}

# the shapes of the BASELINE configs, evaluator level (synthetic data of the shape)
SHAPES = [
    dict(name="51Peg", N=256, n_inst=1, flags=ModelFlags(npmax=1), lanes=262144,
         text="examples/51Peg: N=256, Np<=1, white noise"),
    dict(name="BL2009", N=34, n_inst=1, flags=ModelFlags(npmax=5), lanes=524288,
         text="examples/BL2009: N=34, Np<=5, white noise"),
    dict(name="many_planets", N=190, n_inst=3,
         flags=ModelFlags(npmax=10, hyperpriors=True, studentt=True, multi_instrument=True), lanes=131072,
         text="examples/many_planets, BASELINE variant: N=190, Np<=10, hyperpriors, Student-t, 3 instruments"),
    dict(name="CoRoT7", N=177, n_inst=1, flags=ModelFlags(GP=True, npmax=5, hyperpriors=True), lanes=16384,
         text="examples/CoRoT7: N=177, QP-GP (shared-memory-resident Cholesky), Np<=5, hyperpriors"),
]


def workload_flops(w, np_mean=None):
    N, f = w["N"], w["flags"]
    npl = f.npmax if np_mean is None else np_mean
    return flops_gp(N, npl) if f.GP else flops_keplerian(N, npl, bool(f.studentt))


def config_dict(kepler_mode: int = 0):
    """Identical for both arms (the reference arm times bounded samples of the same workloads)."""
    return {"workload": "kep: " + WORKLOADS["kep"]["text"] + "; gp: " + WORKLOADS["gp"]["text"] +
                        "; gp4096: " + WORKLOADS["gp4096"]["text"],
            "lanes_per_gpu": WORKLOADS["kep"]["lanes"], "gp_lanes_per_gpu": WORKLOADS["gp"]["lanes"],
            "gp4096_lanes_per_gpu": WORKLOADS["gp4096"]["lanes"], "np": "fixed at npmax=10 (np_mix: U{0..10})",
            "kepler_mode": kepler_mode,
            "parallelism": "lanes sharded over the GPUs, no data-path collective",
            "l2": "flushed between timed iterations (256 MiB memset on the kernel stream); every step's "
                  "inputs / working set also exceed L2"}


class ClockSampler:
    """nvidia-smi clocks + throttle reasons while the timed region runs."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.idx = gpu_index
        self.rows = []
        self.proc = None
        self.th = None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.idx), "--query-gpu=" + self.Q,
                 "--format=csv,noheader,nounits", "-lms", "100"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except OSError:
            return

        def pump():
            for line in self.proc.stdout:
                self.rows.append([c.strip() for c in line.split(",")])
        self.th = threading.Thread(target=pump, daemon=True)
        self.th.start()

    def stop(self) -> dict:
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, smax, power, reasons = [], None, [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            if len(r) < 9:
                continue
            try:
                sm.append(float(r[1]))
                smax = float(r[2])
                power.append(float(r[3]))
            except ValueError:
                continue
            for k, nm in enumerate(names):
                if r[5 + k].lower().startswith("active"):
                    reasons.add(nm)
        sm_sorted = sorted(sm)
        # "under load": the upper half of the samples (idle samples bracket the region)
        load = sm_sorted[len(sm_sorted) // 2:] if sm_sorted else []
        return {"sm_mhz": statistics.median(load) if load else None, "sm_max_mhz": smax,
                "reasons": sorted(reasons), "samples": len(sm), "power_w_max": max(power) if power else None}


def merge_clocks(cs):
    cs = [c for c in cs if c and c.get("sm_mhz") is not None]
    if not cs:
        return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
    return {"sm_mhz": statistics.median([c["sm_mhz"] for c in cs]), "sm_max_mhz": cs[0]["sm_max_mhz"],
            "reasons": sorted(set(sum((c["reasons"] for c in cs), []))), "samples": sum(c["samples"] for c in cs),
            "power_w_max": max(c["power_w_max"] or 0 for c in cs)}


def init_dist(n_gpus: int):
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1:
        if os.environ.get("NCCL_DEBUG", "VERSION").upper() == "VERSION":
            os.environ["NCCL_DEBUG"] = "WARN"   # keep NCCL's version banner off stdout (one JSON line)
        import torch
        import torch.distributed as dist
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    return rank, world, local


def cpu_model() -> str:
    try:
        for line in open("/proc/cpuinfo"):
            if line.startswith("model name"):
                return line.split(":", 1)[1].strip()
    except OSError:
        pass
    return platform.processor() or "unknown"


def cpu_arm(flags: ModelFlags, data, lanes: int, threads: int, np_mode: str = "max"):
    """Time the reference's own CPU implementation (oracle/_ref, kind "reference") or, where it was
    not built, the plain-C restatement (kind "port") on `lanes` lanes with `threads` host threads."""
    from oracle.oracle import Oracle, have_reference
    kind = "reference" if have_reference() else "port"
    o = Oracle(kind)
    o.set_rvdata(data)
    o.set_flags(flags)
    P, nps = make_params(flags, data, lanes, np_mode)
    _, secs = o.eval_batch(P, nps, threads)
    return kind, lanes / secs, secs


def load_ncu_profile():
    """Per-launch ncu counters of the HEAD build (tools/ncu_exec_flops.py writes the file from the
    `ncu --metrics` pass of this same script; profiles/ is committed).  None when absent."""
    path = os.path.join(ROOT, "profiles", "ncu_exec_r02.json")
    try:
        return json.load(open(path))
    except (OSError, ValueError):
        return None


def run_workload(eng, lanes: int, steps: int, warmup: int, flops_per_eval: float, peak_tflops: float,
                 exch, sampler_idx: int, e2e_steps: int = None):
    """W warm-up + K timed steps of one workload on this rank's engine."""
    for _ in range(warmup):
        eng.flush_l2()
        eng.log_likelihood_device(lanes)
    eng.sync()
    sampler = ClockSampler(sampler_idx)
    sampler.start()
    if exch:
        exch.barrier()
    n0 = eng.launch_count()
    dev_ms = []
    t0 = time.perf_counter()
    for _ in range(steps):
        eng.flush_l2()
        eng.log_likelihood_device(lanes)
        eng.sync()
        dev_ms.append(eng.last_kernel_ms())
    eng.sync()
    if exch:
        exch.barrier()
    wall = time.perf_counter() - t0
    launches = eng.launch_count() - n0
    clocks = sampler.stop()
    total_ms = float(sum(dev_ms))
    # end to end through the public call, host buffers: H2D + kernels + D2H inside the clock
    out = np.empty(lanes)
    eng.log_likelihood(lanes, out)
    e2e_t = []
    for _ in range(e2e_steps or steps):
        eng.flush_l2()
        eng.sync()
        t1 = time.perf_counter()
        eng.log_likelihood(lanes, out)
        e2e_t.append(time.perf_counter() - t1)
    e2e_s = float(sum(e2e_t)) / len(e2e_t)
    if exch:
        total_ms = exch.max_float(total_ms)
        e2e_s = exch.max_float(e2e_s)
        launches = int(exch.sum_float(float(launches)))
    world = exch.world if exch else 1
    ms_per_step = total_ms / steps
    value = world * lanes / (ms_per_step * 1e-3)
    per_gpu = lanes / (ms_per_step * 1e-3)
    achieved = per_gpu * flops_per_eval * 1e-12
    n_rows = eng.layout["n_rows"]
    return {
        "value": value, "unit": "evals/s", "ms_per_step": ms_per_step, "lanes_per_step_per_gpu": lanes,
        "steps": steps, "gpu_launches": launches, "timed_region_s": total_ms * 1e-3,
        "e2e": {"value": world * lanes / e2e_s, "unit": "evals/s",
                "h2d_bytes_per_step": int(n_rows * lanes * 8 + lanes * 4),
                "d2h_bytes_per_step": int(lanes * 8)},
        "roofline": {"achieved": achieved, "peak": peak_tflops, "unit": "TFLOP/s", "frac": achieved / peak_tflops,
                     "flops_per_eval": flops_per_eval},
        "clocks": clocks, "wall_s_timed_region": wall, "finite": bool(np.all(np.isfinite(out))),
    }


def algorithmic_bytes(key: str, n_rows: int):
    """SURVEY.md 8(d) / DESIGN.md 3: bytes one evaluation has to move through HBM."""
    par = n_rows * 8 + 4 + 8                        # the lane's parameter column + np in, logL out
    if key in ("gp", "gp4096"):
        N = WORKLOADS[key]["N"]
        return par + 4.0 * N * N + 8.0 * N ** 3 / (6 * 64)   # L written once + left-looking panel re-reads, NB=64
    return par


def annotate_roofline(key: str, r: dict, prof, n_rows: int = 0):
    """bound / pipe / dominant kernel, and the ncu counters of the HEAD build (profiles/) where present."""
    rf = r["roofline"]
    if key.startswith("kep") or key in ("51Peg", "BL2009", "many_planets"):
        rf["bound"] = "fp64"
        rf["pipe"] = "FP64 vector (DFMA); no tensor-core instructions: the peak is the FP64 rate of the SM, " \
                     "measured as max(DFMA, DMMA) in this run"
        rf["kernel"] = "keplerian_warp_kernel (1 launch per step)"
    elif key == "CoRoT7":
        rf["bound"] = "tensor"
        rf["pipe"] = "FP64 tensor (DMMA.8x8x4) + FP64 vector; latency-bound on the diagonal-block chain"
        rf["kernel"] = "gps_kernel (1 launch per step, one CTA per lane, factor in shared memory)"
    else:
        rf["bound"] = "tensor"
        rf["pipe"] = "FP64 tensor (DMMA.8x8x4)"
        rf["kernel"] = "gpb_update_kernel dominates (per wave and lane group: prep + N/64 x (diag + update) + finish)"
    rf["peak_source"] = "kc_measure_fp64_peak in this run (max of DFMA / DMMA m8n8k4); MEASURED_PEAKS.json has no FP64 entry"
    rf["traffic"] = None
    p = (prof or {}).get(key)
    if p:
        # counters of one evaluation of p["lanes"] lanes under ncu (HEAD build); per eval, then x the live rate
        per_eval = p["executed_flops"] / p["lanes"]
        rf["executed_flops_per_eval"] = per_eval
        rf["executed_tflops"] = per_eval * (rf["achieved"] / rf["flops_per_eval"])
        rf["executed_over_algorithmic"] = per_eval / rf["flops_per_eval"]
        rf["traffic"] = p["dram_bytes"] / p["lanes"] * r["lanes_per_step_per_gpu"]
        rf["traffic_detail"] = {"dram_bytes_per_eval": p["dram_bytes"] / p["lanes"],
                                "algorithmic_bytes_per_eval": algorithmic_bytes(key, n_rows),
                                "source": p.get("source")}


def predictive_leg(data, flags, device: int, threads: int, B: int = 1184, nt: int = 1000, cpu_B: int = 4):
    """posterior-predictive curves (deterministic model + GP mean on a time
    grid) per second through kc_predict with host buffers, beside the CPU restatement of
    pykima's per-sample loop (oracle/predictive.py) on a few samples."""
    from kima_b200 import LikelihoodEngine
    from oracle import predictive as opred
    from oracle.oracle import Oracle
    P, nps = make_params(flags, data, B, "max", seed=47)
    tg = np.linspace(data.t.min(), data.t.max(), nt)
    with LikelihoodEngine(data, flags, B, device=device) as eng:
        eng.predict(P, nps, tg)
        secs, kms = [], []
        for _ in range(3):
            t0 = time.perf_counter()
            m, g = eng.predict(P, nps, tg)
            secs.append(time.perf_counter() - t0)
            kms.append(eng.last_kernel_ms())
        launches = eng.launch_count()
    orc = Oracle("port")
    orc.set_rvdata(data)
    orc.set_flags(flags)
    t0 = time.perf_counter()
    err = 0.0
    for b in range(cpu_B):
        want = opred.model(orc, flags, data, P, b, nps[b], tg) + opred.gp_mean(orc, flags, data, P, b, nps[b], tg)
        err = max(err, float(np.max(np.abs(m[b] + g[b] - want)) / (np.abs(want).max() + 1.0)))
    cpu_s = time.perf_counter() - t0
    return {"unit": "predictive curves/s", "workload": "QP-GP N=%d Np=%d, %d posterior samples, %d grid times" % (data.N(), NP, B, nt),
            "value": B / min(secs), "ms_per_call": 1e3 * min(secs), "kernels_ms": min(kms), "gpu_launches": launches,
            "d2h_bytes_per_call": 2 * 8 * B * nt,
            "cpu_baseline": {"value": cpu_B / cpu_s, "unit": "predictive curves/s", "cores": threads, "kind": "port",
                             "sample": "%d samples, numpy/LAPACK restatement of pykima's per-sample loop, %.1f s" % (cpu_B, cpu_s)},
            "max_rel_err_vs_cpu": err}


def sampler_device_leg(local: int, lanes: int = 65536, steps: int = 50, reps: int = 3):
    """SURVEY.md 8 row f1, device-timed: whole MCMC steps (proposal, prior test, likelihood, accept, level
    move, counters) of `lanes` lanes on the GPU through kc_sampler_run, synthetic white-noise data of the
    51Peg shape (N=256, Np in {0,1}); CUDA events on the library's stream around each block of steps."""
    from kima_b200 import LikelihoodEngine
    from kima_b200.sampler import DeviceSampler, SamplerOptions, default_prior_table, draw_from_prior
    flags = ModelFlags(npmax=1)
    data = make_data(256)
    pri = default_prior_table(flags, data)
    with LikelihoodEngine(data, flags, lanes, device=local) as eng:
        ds = DeviceSampler(eng, pri, SamplerOptions(lanes=lanes, num_particles=1, thread_steps=steps, max_levels=64))
        S, nps = draw_from_prior(ds, flags, pri, lanes)
        logl = eng.evaluate(S[:eng.layout["n_rows"]], nps)
        logl = np.where(np.isfinite(logl), logl, -1e300)
        tb = np.random.default_rng(1).uniform(0.01, 0.99, lanes)
        lv_logl = np.concatenate([[-1e300], np.quantile(logl, np.linspace(0.1, 0.8, 7))])
        ds.upload(S, nps, logl, tb, np.zeros(lanes, np.int32))
        tries = np.full(8, 1000, np.uint64)
        ms = []
        launches0 = eng.launch_count()
        for _ in range(reps + 1):  # the first block is the warm-up
            ds.set_levels(lv_logl, np.full(8, 0.5), -np.arange(8.0), tries, True, True)
            ds.run(steps)
            cnt, _ = ds.collect()
            ms.append(ds.last_run_ms())
            tries = tries + cnt[:, 1]
        launches = eng.launch_count() - launches0
        eng.log_likelihood_device(lanes)
        eng.sync()
        lk = eng.last_kernel_ms()
        ds.close()
    best = min(ms[1:])
    return {"unit": "MCMC steps/s", "workload": "white noise N=256, Np<=1 (51Peg shape), %d lanes x 1 particle, "
            "8 levels, blocks of %d lock-step steps" % (lanes, steps), "value": lanes * steps / best * 1e3,
            "ms_per_step": best / steps, "likelihood_kernel_ms_per_step": lk, "gpu_launches": launches,
            "accept_rate": float(cnt[:, 0].sum() / max(1, cnt[:, 1].sum())),
            "kernels": "smp_sorted_propose_kernel<256> (accept of step i + move-sorted proposal of step i+1) + keplerian kernel per step"}


def reference_arm(a, base, threads):
    """The reference's own CPU implementation of the path (oracle/_ref/libkima_ref.so: the unmodified
    RVmodel::calculate_mu / calculate_C / log_likelihood) on all host threads; every step is a bounded
    sample of the workload."""
    out = dict(base, impl="reference")
    cm = cpu_model()
    res = {}
    for key in ("kep", "gp", "gp4096"):
        w = WORKLOADS[key]
        data = make_data(w["N"])
        lanes = w["ref_lanes"] or threads      # gp4096: one lane per host thread (about 4-8 s per step)
        steps = a.steps if w["max_steps"] is None else min(a.steps, 2)
        if key != "gp4096":
            cpu_arm(w["flags"], data, max(lanes // 8, threads), threads)   # untimed: caches, thread pool
        secs, kind = [], None
        for _ in range(steps):
            kind, _, s = cpu_arm(w["flags"], data, lanes, threads)
            secs.append(s)
        v = lanes * len(secs) / sum(secs)
        res[key] = {"value": v, "unit": "evals/s", "ms_per_step": 1e3 * sum(secs) / len(secs), "steps": len(secs),
                    "cpu_baseline": {"value": v, "unit": "evals/s", "cores": threads, "cpu_model": cm, "kind": kind,
                                     "sample": "%d lanes of the workload per step, %d steps, %d host threads" % (lanes, len(secs), threads)},
                    "e2e": {"value": v, "unit": "evals/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    out.update(res["kep"])
    out["value_gp"] = res["gp"]["value"]
    out["e2e_gp"] = res["gp"]["e2e"]
    out["ms_per_step_gp"] = res["gp"]["ms_per_step"]
    out["gp"] = res["gp"]
    out["gp4096"] = res["gp4096"]
    out["config"] = config_dict(a.kepler_mode)
    print(json.dumps(out))
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="kima_b200", choices=["kima_b200", "reference"])
    ap.add_argument("--kepler-mode", type=int, default=0)
    ap.add_argument("--skip-cpu-baseline", action="store_true")
    ap.add_argument("--workloads", default="kep,gp,gp4096")
    ap.add_argument("--skip-variants", action="store_true", help="no np_mix / shapes legs")
    ap.add_argument("--skip-sampler", action="store_true")
    ap.add_argument("--skip-predictive", action="store_true")
    ap.add_argument("--profile-one", default="", help="ncu helper: run ONE evaluation of this workload / shape and exit")
    a = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    threads = os.cpu_count() or 1
    warmup = max(a.warmup, 3)
    base = {"metric": METRIC, "unit": "evals/s", "n_gpus": a.gpus, "steps": a.steps,
            "warmup": warmup, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic"}

    if a.impl == "reference":
        if rank != 0:
            return 0
        return reference_arm(a, base, threads)

    # stdout carries exactly one JSON line: whatever libraries write there meanwhile (NCCL's version
    # banner at the first collective, ...) goes to stderr
    sys.stdout.flush()
    saved_stdout = os.dup(1)
    os.dup2(2, 1)
    rank, world, local = init_dist(a.gpus)
    from kima_b200 import LikelihoodEngine
    from kima_b200.dist import TorchBarrierExchange
    exch = None
    if world > 1:
        import torch
        exch = TorchBarrierExchange(torch.device("cuda", local))
    prof = load_ncu_profile()
    out = dict(base)
    results, np_mix, shapes = {}, {}, []
    peak = None
    wanted = a.workloads.split(",")

    def one(key, w, np_mode, lanes, steps, e2e_steps=None, cpu=False, data=None):
        nonlocal peak
        data = data if data is not None else make_data(w["N"], w.get("n_inst", 1))
        eng = LikelihoodEngine(data, w["flags"], lanes, device=local, kepler_mode=a.kepler_mode,
                               gp_slots=w.get("gp_slots", 0))
        try:
            if peak is None:
                dfma, dmma = eng.measure_fp64_peak(5)
                peak = max(dfma, dmma)
            # each rank scores its own shard of the global lane set (different seeds per rank)
            P, nps = make_params(w["flags"], data, lanes, np_mode, seed=43 + rank)
            eng.set_params(P, nps)
            eng.log_likelihood(lanes)  # puts the block on the device
            fl = workload_flops(w, None if np_mode == "max" else float(np.mean(nps)))
            r = run_workload(eng, lanes, steps, warmup, fl, peak, exch, local, e2e_steps)
            r["np"] = "fixed at npmax" if np_mode == "max" else "U{0..npmax}, mean %.2f" % float(np.mean(nps))
        finally:
            eng.close()
        if cpu and rank == 0 and world == 1 and not a.skip_cpu_baseline:
            ref_lanes = (w.get("ref_lanes") or threads)
            ref_lanes = ref_lanes * 4 if key == "kep" else (ref_lanes * 8 if key == "gp" else ref_lanes)
            kind, v, secs = cpu_arm(w["flags"], data, ref_lanes, threads, np_mode)
            r["cpu_baseline"] = {"value": v, "unit": "evals/s", "cores": threads, "cpu_model": cpu_model(), "kind": kind,
                                 "sample": "%d lanes of the same workload, %d host threads, %.1f s" % (ref_lanes, threads, secs)}
        return r

    if a.profile_one:
        # one evaluation, no timing loop: the command ncu wraps (tools/r02_*.sh)
        key = a.profile_one
        w = WORKLOADS.get(key) or next(s for s in SHAPES if s["name"] == key)
        lanes = {"kep": 8192, "gp": 1184, "gp4096": 148}.get(key, 16384)   # tools/ncu_exec_flops.py: LANES
        data = make_data(w["N"], w.get("n_inst", 1))
        with LikelihoodEngine(data, w["flags"], lanes, device=local, kepler_mode=a.kepler_mode) as eng:
            P, nps = make_params(w["flags"], data, lanes, "max", seed=43)
            for _ in range(2):
                eng.evaluate(P, nps)
        os.dup2(saved_stdout, 1)
        print(json.dumps({"profile_one": key, "lanes": lanes}))
        return 0

    for key in ("kep", "gp", "gp4096"):
        if key not in wanted:
            continue
        w = WORKLOADS[key]
        steps = a.steps if w["max_steps"] is None else min(a.steps, w["max_steps"])
        results[key] = one(key, w, "max", w["lanes"], steps, e2e_steps=min(steps, 5), cpu=True)
        annotate_roofline(key, results[key], prof, results[key]["e2e"]["h2d_bytes_per_step"] // (8 * w["lanes"]))
        if not a.skip_variants:
            # Np ~ U{0..npmax}, the DNest4-realistic mix (fewer steps: a side figure)
            np_mix[key] = one(key + "_mix", w, "mix", w["lanes"], min(steps, 5), e2e_steps=2)
            annotate_roofline(key, np_mix[key], None)
    if not a.skip_variants:
        for s in SHAPES:
            r = one(s["name"], s, "mix", s["lanes"], min(a.steps, 10), e2e_steps=2)
            annotate_roofline(s["name"], r, prof)
            shapes.append({"name": s["name"], "workload": s["text"] + "; %d lanes/GPU per step, Np ~ U{0..npmax}" % s["lanes"],
                           "value": r["value"], "unit": "evals/s", "ms_per_step": r["ms_per_step"],
                           "e2e": r["e2e"], "roofline": r["roofline"], "gpu_launches": r["gpu_launches"]})

    first = next(k for k in ("kep", "gp", "gp4096") if k in results)
    top = dict(results[first])
    top_roof = top.pop("roofline")
    out.update(top)
    if "gp" in results and first != "gp":
        g = results["gp"]
        out["value_gp"], out["e2e_gp"], out["ms_per_step_gp"] = g["value"], g["e2e"], g["ms_per_step"]
        out["gp"] = g
    if "gp4096" in results and first != "gp4096":
        out["gp4096"] = results["gp4096"]
    # the headline roofline is the workload furthest from its roofline; all of them beneath it
    worst = min(results, key=lambda k: results[k]["roofline"]["frac"])
    out["roofline"] = dict(results[worst]["roofline"], workload=worst,
                           workloads={k: results[k]["roofline"] for k in results})
    out["gpu_launches"] = int(sum(r["gpu_launches"] for r in results.values()))
    out["clocks"] = merge_clocks([r["clocks"] for r in results.values()])
    if np_mix:
        out["np_mix"] = {k: {kk: r[kk] for kk in ("value", "unit", "ms_per_step", "np", "e2e", "roofline")}
                         for k, r in np_mix.items()}
    if shapes:
        out["shapes"] = shapes
    out["config"] = config_dict(a.kepler_mode)   # identical in both arms
    out["fp64_peak_tflops_measured"] = peak

    if rank == 0 and world == 1 and not a.skip_sampler:
        # the reference's unmodified examples end to end, MCMC steps/s of
        # the batched GPU sampler (host/BatchedSampler.h) next to the reference's CPU sampler
        sys.path.insert(0, os.path.join(ROOT, "tools"))
        import sampler_bench
        os.environ.setdefault("KIMA_CUDA_DEVICE", str(local))
        out["sampler"] = {"unit": "MCMC steps/s", "lanes": 4096, "cpu_threads": threads,
                          "runs": [sampler_bench.compare(n, 4096, threads, gpu_saves=80, cpu_saves=100)
                                   for n in ("CoRoT7", "51Peg")]}
        # CoRoT7 (GP): one factor-cache slot per particle is the default on this path too
        # (host/RVmodelDevice.h::enable_factor_cache); the same run without it beside it
        out["sampler"]["runs"][0]["factor_cache"] = True
        out["sampler"]["runs"].append(dict(example="CoRoT7", factor_cache=False,
                                           gpu=sampler_bench.run("gpu", "CoRoT7", 4096, 100, 80,
                                                                 env={"KIMA_CUDA_GP_CACHE": "0"})))
        # the same examples with proposals, prior test, accept logic and level
        # counters on the GPU as well (KIMA_DEVICE_PROPOSALS=1, include/kima_sampler.h): particles stay
        # in HBM, the host only does DNest4's bookkeeping between blocks of thread_steps steps
        dev_env = {"KIMA_DEVICE_PROPOSALS": "1"}
        # CoRoT7 (GP): the sampler's factor cache (kima_cuda.h: kc_gp_cache_enable) is on by default with device
        # proposals; the same run without it is reported beside it (the chains are the same bytes)
        out["sampler"]["device_proposals"] = [
            sampler_bench.run("gpu", "51Peg", 65536, 100, 20, env=dev_env),
            dict(sampler_bench.run("gpu", "CoRoT7", 16384, 100, 10, env=dev_env) or {}, factor_cache=True),
            dict(sampler_bench.run("gpu", "CoRoT7", 16384, 100, 10, env=dict(dev_env, KIMA_CUDA_GP_CACHE="0")) or {},
                 factor_cache=False)]
        out["sampler"]["device_steps"] = sampler_device_leg(local)
    if rank == 0 and world == 1 and not a.skip_predictive and not a.skip_cpu_baseline:
        out["predictive"] = predictive_leg(make_data(1000), WORKLOADS["gp"]["flags"], local, threads)
    if world > 1 and not a.skip_sampler and os.path.exists(os.path.join(ROOT, "host", "_build", "kima_gpu_CoRoT7")):
        out["sampler"] = multi_gpu_sampler_leg(rank, world, local, threads, exch)
    if exch:
        import torch.distributed as dist
        dist.barrier()
        dist.destroy_process_group()
    sys.stdout.flush()
    os.dup2(saved_stdout, 1)
    os.close(saved_stdout)
    if rank == 0:
        print(json.dumps(out))
    return 0


def multi_gpu_sampler_leg(rank, world, local, threads, exch):
    """The reference's examples end to end at N GPUs: one C++ process per GPU (host/BatchedSampler.h),
    lanes sharded, the level-creation barrier over NCCL (weak scaling: lanes per GPU fixed).  Rank 0
    also reports the parity bit: a small run at N ranks writes the files of the same run in one process."""
    import shutil
    import torch.distributed as dist
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    import sampler_bench
    runs = []
    port0 = int(os.environ.get("MASTER_PORT", "29500")) + 11
    # (example, lanes per GPU, saves, device proposals?): host proposals as in the 1-GPU leg (the
    # reference's byte-identical chains: perturb on the host cores, which the ranks share), and the
    # whole step on the GPUs (each rank keeps its lanes' particles in HBM) for both examples.
    # `scaling` = steps/s over run() at `world` ranks / (world x the same binary at the same lanes
    # per GPU in ONE process, which rank 0 runs first while the others wait).
    # (51Peg: 200 barriers, so that run() lasts seconds and NCCL's lazy connection set-up inside the first
    # barrier does not decide the figure)
    legs = (("CoRoT7", 4096, 80, False), ("CoRoT7", 16384, 10, True), ("51Peg", 16384, 200, True))
    for i, (name, lpg, saves, dev) in enumerate(legs):
        env = {"KIMA_DEVICE_PROPOSALS": "1"} if dev else None
        one = None
        if rank == 0:
            one = sampler_bench.run("gpu", name, lpg, 100, saves,
                                    env=dict(env or {}, KIMA_CUDA_DEVICE=str(local), RANK="0", WORLD_SIZE="1",
                                             LOCAL_RANK=str(local), KIMA_HOST_THREADS=str(max(1, threads // world))))
        box = [sampler_bench.prepare_dir(name, lpg * world, 100, saves) if rank == 0 else None]
        dist.broadcast_object_list(box, src=0)
        dist.barrier()
        r = sampler_bench.run_rank(name, box[0], lpg * world, rank, world, local, port0 + i,
                                   max(1, threads // world), env=env)
        wall = exch.max_float(float(r.get("wall_s", 0.0)))
        bad = exch.max_float(1.0 if "error" in r else 0.0)
        if rank == 0:
            if bad or not r.get("mcmc_steps"):
                runs.append({"example": name, "device_proposals": dev, "error": r.get("error", "a rank failed")})
            else:
                sps_run = r["mcmc_steps"] / r["run_s"] if r.get("run_s") else None
                one_run = (one or {}).get("steps_per_s_run")
                runs.append({"example": name, "lanes": lpg * world, "lanes_per_gpu": lpg, "device_proposals": dev,
                             "mcmc_steps": r["mcmc_steps"], "wall_s": wall, "steps_per_s": r["mcmc_steps"] / wall,
                             "run_s": r.get("run_s"), "steps_per_s_run": sps_run,
                             "one_process_steps_per_s_run": one_run, "one_process_steps_per_s": (one or {}).get("steps_per_s"),
                             "scaling": sps_run / (world * one_run) if sps_run and one_run else None})
            shutil.rmtree(box[0], ignore_errors=True)
    out = {"unit": "MCMC steps/s", "ranks": world, "host_threads_per_rank": max(1, threads // world), "runs": runs}
    # parity bit: BL2009, 16 lanes over `world` ranks against the same run in one process (rank 0)
    try:
        out["chains_equal"] = sampler_bench.chains_equal_check(rank, world, local, port0 + 7, exch)
    except Exception as e:  # the timing legs stand on their own
        out["chains_equal"] = {"error": repr(e)}
    return out


if __name__ == "__main__":
    sys.exit(main())
