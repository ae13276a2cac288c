#!/usr/bin/env python
"""Condense an .ncu-rep (ncu --set full) into the handful of counters DESIGN.md / bench.py cite.

    python tools/ncu_summary.py gpurun_out/<tag>/kep_full.ncu-rep > profiles/<name>.json
Runs here (no GPU needed): it only reads the report with `ncu -i ... --page raw --csv`.
"""
import csv
import io
import json
import subprocess
import sys

KEYS = {
    "gpu__time_duration.sum": "duration",
    "dram__bytes_read.sum": "dram_read",
    "dram__bytes_write.sum": "dram_write",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed": "dram_pct",
    "lts__t_bytes.sum": "l2_bytes",
    "launch__registers_per_thread": "regs",
    "launch__grid_size": "grid",
    "launch__block_size": "block",
    "launch__shared_mem_per_block_dynamic": "smem_dyn",
    "launch__shared_mem_per_block_static": "smem_static",
    "launch__occupancy_limit_registers": "occ_limit_regs",
    "launch__occupancy_limit_shared_mem": "occ_limit_smem",
    "sm__warps_active.avg.pct_of_peak_sustained_active": "warps_active_pct",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed": "sm_throughput_pct",
    "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_elapsed": "fp64_pipe_pct_elapsed",
    "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active": "fp64_pipe_pct_active",
    "sm__inst_executed_pipe_tensor_subpipe_dmma.avg.pct_of_peak_sustained_active": "dmma_pct_active",
    "sm__ops_path_tensor_src_fp64.avg.pct_of_peak_sustained_elapsed": "dmma_ops_pct_elapsed",
    "sm__issue_active.avg.pct_of_peak_sustained_elapsed": "issue_active_pct",
    "smsp__sass_thread_inst_executed_op_dfma_pred_on.sum": "dfma_thread_inst",
    "smsp__sass_thread_inst_executed_op_dadd_pred_on.sum": "dadd_thread_inst",
    "smsp__sass_thread_inst_executed_op_dmul_pred_on.sum": "dmul_thread_inst",
    "sm__ops_path_tensor_src_fp64.sum": "dmma_ops",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum": "smem_bank_conflicts",
    "smsp__cycles_elapsed.avg.per_second": "sm_clock_hz",
}
STALLS = "smsp__average_warps_issue_stalled_"


def summarise(path):
    raw = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True,
                         text=True, check=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units = rows[0], rows[1]
    out = []
    for vals in rows[2:]:
        d = dict(zip(hdr, vals))
        u = dict(zip(hdr, units))
        k = {"kernel": d.get("Kernel Name"), "id": d.get("ID")}
        for name, short in KEYS.items():
            if name in d and d[name] != "":
                k[short] = {"v": d[name], "unit": u[name]}
        st = {}
        for h in hdr:
            if h.startswith(STALLS) and h.endswith("_per_issue_active.ratio"):
                try:
                    st[h[len(STALLS):-len("_per_issue_active.ratio")]] = round(float(d[h]), 3)
                except ValueError:
                    pass
        k["stalls_per_issue"] = dict(sorted(st.items(), key=lambda kv: -kv[1])[:6])
        out.append(k)
    return out


if __name__ == "__main__":
    print(json.dumps({"report": sys.argv[1], "kernels": summarise(sys.argv[1])}, indent=1))
