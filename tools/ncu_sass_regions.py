#!/usr/bin/env python
"""Aggregate warp-stall samples of an ncu report's SASS page into regions split at
the kernel's barriers / DMMA loop, to see which phase of a kernel holds the time.
    python tools/ncu_sass_regions.py report.ncu-rep [chunk]
"""
import csv, io, subprocess, sys, collections
rep = sys.argv[1]
chunk = int(sys.argv[2]) if len(sys.argv) > 2 else 0
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr = rows[1]
ix = {h: i for i, h in enumerate(hdr)}
ins = rows[2:]
tot = sum(int(r[ix["# Samples"]] or 0) for r in ins)
stall_cols = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
def op(r): return r[ix["Source"]].split()[0] if not r[ix["Source"]].strip().startswith("@") else r[ix["Source"]].split()[1]
# regions: split at BAR.SYNC
regions = []; cur = []
for n, r in enumerate(ins):
    cur.append((n, r))
    if "BAR.SYNC" in r[ix["Source"]] or (chunk and len(cur) >= chunk):
        regions.append(cur); cur = []
if cur: regions.append(cur)
print("total samples", tot, "instructions", len(ins))
for reg in regions:
    s = sum(int(r[ix["# Samples"]] or 0) for _, r in reg)
    if s < 0.005 * tot: continue
    ops = collections.Counter(op(r).split(".")[0] for _, r in reg)
    st = collections.Counter()
    for _, r in reg:
        for c in stall_cols:
            st[c] += int(r[ix[c]] or 0)
    ex = sum(int(r[ix["Instructions Executed"]] or 0) for _, r in reg)
    print("insn %5d-%5d  samples %6d (%4.1f%%) exec %10d  ops %s  stalls %s" % (
        reg[0][0], reg[-1][0], s, 100.0 * s / tot, ex, dict(ops.most_common(5)), dict(st.most_common(4))))
