#!/usr/bin/env python
"""Warp-stall samples of an ncu report per CUDA source line (needs -lineinfo and --import-source on).
    python tools/ncu_lines.py report.ncu-rep [top]
"""
import collections
import csv
import io
import subprocess
import sys

rep = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
cur = None
hdr = None
agg = collections.defaultdict(lambda: [0, 0, collections.Counter(), ""])
tot = 0
for r in rows:
    if not r:
        continue
    if r[0] == "File Path":
        cur = r[1].split("/")[-1]
        continue
    if r[0] == "Line No":
        hdr = r
        ix = {}
        for i, h in enumerate(hdr):
            ix.setdefault(h, i)
        stall = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
        continue
    if hdr is None or r[0] in ("Function Name",):
        continue
    try:
        n = int(r[ix["# Samples"]] or 0)
        ex = int(r[ix["Instructions Executed"]] or 0)
    except (ValueError, IndexError):
        continue
    key = (cur, r[0])
    a = agg[key]
    a[0] += n
    a[1] += ex
    a[3] = r[1].strip()[:90]
    for h in stall:
        try:
            a[2][h[6:]] += int(r[ix[h]] or 0)
        except ValueError:
            pass
    tot += n
print("total samples", tot)
for (f, line), a in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
    print("%5.1f%% %9d exec  %-18s:%-4s %-28s | %s" % (100.0 * a[0] / max(tot, 1), a[1], f, line,
                                                   dict(a[2].most_common(3)), a[3]))
