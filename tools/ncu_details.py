#!/usr/bin/env python
"""Print the interesting rows of `ncu --page details` for kernels matching a pattern.
  python tools/ncu_details.py gpurun_out/X.ncu-rep <kernel-substring> [section-substring ...]"""
import csv
import subprocess
import sys

rep, pat = sys.argv[1], sys.argv[2]
secs = sys.argv[3:] or ["Speed Of Light", "Memory Workload", "Occupancy", "Warp State", "Scheduler", "Launch"]
raw = subprocess.run(["ncu", "-i", rep, "--page", "details", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
h = rows[0]
ki, idc, sn, mn, mv, mu = (h.index(x) for x in ("Kernel Name", "ID", "Section Name", "Metric Name", "Metric Value", "Metric Unit"))
seen = None
for r in rows[1:]:
    if pat in r[ki] and any(s in r[sn] for s in secs):
        if seen is None:
            seen = r[idc]
        if r[idc] != seen:
            continue
        print(f"{r[sn][:18]:18s} | {r[mn][:50]:50s} | {r[mv]} {r[mu]}")
