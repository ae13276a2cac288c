#!/usr/bin/env python
"""Executed FP64 FLOPs and DRAM bytes of ONE evaluation of each bench workload, from ncu.

    ncu --metrics <METRICS> --clock-control none --csv --log-file gpurun_out/<tag>/exec_<key>.csv \
        python bench.py --profile-one <key>            # on the GPU box, one per workload
    python tools/ncu_exec_flops.py gpurun_out/<tag> > profiles/ncu_exec_r02.json    # here

`bench.py --profile-one` evaluates the same batch twice; the launches of the second evaluation are
summed.  executed_flops = 2 x DFMA + DADD + DMUL thread-instructions + the FP64 tensor-core operations
(sm__ops_path_tensor_src_fp64: one multiply and one add per MAC of every DMMA), the figure SURVEY.md
8(d) asks for beside the algorithmic one.  bench.py reads the resulting JSON (roofline.executed_*,
roofline.traffic).
"""
import csv
import json
import os
import sys

METRICS = ("smsp__sass_thread_inst_executed_op_dfma_pred_on.sum,smsp__sass_thread_inst_executed_op_dadd_pred_on.sum,"
           "smsp__sass_thread_inst_executed_op_dmul_pred_on.sum,sm__ops_path_tensor_src_fp64.sum,"
           "dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum")
UNIT = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1e-9, "us": 1e-6, "usecond": 1e-6,
        "msecond": 1e-3, "ms": 1e-3, "nsecond": 1e-9, "second": 1.0, "inst": 1.0, "": 1.0}
LANES = {"kep": 8192, "gp": 1184, "gp4096": 148}


def parse(path):
    rows = list(csv.reader(open(path, errors="replace")))
    h = [i for i, r in enumerate(rows) if r and r[0] == "ID"]
    if not h:
        return None
    launches = {}
    for r in rows[h[0] + 1:]:
        if len(r) < 15:
            continue
        lid = int(r[0])
        d = launches.setdefault(lid, {"kernel": r[4].split("(")[0].replace("void ", "").replace("kc::", "")})
        try:
            d[r[12]] = float(r[14].replace(",", "")) * UNIT.get(r[13], 1.0)
        except ValueError:
            pass
    ids = sorted(launches)
    half = ids[len(ids) // 2:]  # the second of the two identical evaluations
    tot = {"dfma": 0.0, "dadd": 0.0, "dmul": 0.0, "dmma_ops": 0.0, "dram": 0.0, "time_s": 0.0}
    per_kernel = {}
    for i in half:
        d = launches[i]
        v = {"dfma": d.get("smsp__sass_thread_inst_executed_op_dfma_pred_on.sum", 0.0),
             "dadd": d.get("smsp__sass_thread_inst_executed_op_dadd_pred_on.sum", 0.0),
             "dmul": d.get("smsp__sass_thread_inst_executed_op_dmul_pred_on.sum", 0.0),
             "dmma_ops": d.get("sm__ops_path_tensor_src_fp64.sum", 0.0),
             "dram": d.get("dram__bytes_read.sum", 0.0) + d.get("dram__bytes_write.sum", 0.0),
             "time_s": d.get("gpu__time_duration.sum", 0.0)}
        k = per_kernel.setdefault(d["kernel"], dict.fromkeys(v, 0.0) | {"launches": 0})
        for a, b in v.items():
            tot[a] += b
            k[a] += b
        k["launches"] += 1
    tot["launches"] = len(half)
    return tot, per_kernel


def main(d):
    out = {}
    for f in sorted(os.listdir(d)):
        if not (f.startswith("exec_") and f.endswith(".csv")):
            continue
        key = f[5:-4]
        p = parse(os.path.join(d, f))
        if not p:
            continue
        tot, per_kernel = p
        lanes = LANES.get(key)
        if lanes is None:
            side = os.path.join(d, "exec_%s.lanes" % key)
            lanes = int(open(side).read()) if os.path.exists(side) else 16384
        flops = 2 * tot["dfma"] + tot["dadd"] + tot["dmul"] + tot["dmma_ops"]
        out[key] = {"lanes": lanes, "executed_flops": flops, "dram_bytes": tot["dram"],
                    "kernel_time_s_under_ncu": tot["time_s"], "launches": tot["launches"],
                    "thread_inst": {k: tot[k] for k in ("dfma", "dadd", "dmul")}, "dmma_ops": tot["dmma_ops"],
                    "per_kernel": {k: {"launches": v["launches"], "time_share": v["time_s"] / max(tot["time_s"], 1e-30),
                                       "flops": 2 * v["dfma"] + v["dadd"] + v["dmul"] + v["dmma_ops"],
                                       "dram_bytes": v["dram"]} for k, v in per_kernel.items()},
                    "source": "ncu --metrics pass over `bench.py --profile-one %s` (second of two evaluations of %d lanes), %s"
                              % (key, lanes, os.path.join(d, f))}
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main(sys.argv[1])
