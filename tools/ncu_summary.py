#!/usr/bin/env python
"""Turn the ncu outputs brought back in gpurun_out/ into the tracked summaries under profiles/.

  python tools/ncu_summary.py <round-tag> [--launches gpurun_out/X.csv] [--rep gpurun_out/Y.ncu-rep]

Writes profiles/<tag>_launches.md (per-kernel launch count, mean and share of device time from the
`--metrics gpu__time_duration.sum` pass), profiles/<tag>_ncu_full.csv (selected metrics of the
`--set full` capture, one row per captured launch) and updates profiles/traffic.json
(dram read+write bytes per launch, per kernel) which bench.py reports as roofline.traffic.
"""
import argparse
import collections
import csv
import json
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
METRICS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
           "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
           "sm__throughput.avg.pct_of_peak_sustained_elapsed",
           "sm__warps_active.avg.pct_of_peak_sustained_active",
           "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
           "launch__shared_mem_per_block_static", "launch__shared_mem_per_block_dynamic",
           "lts__t_bytes.sum", "l1tex__t_bytes.sum", "sm__inst_executed.sum"]


def short(name):
    m = re.search(r"(?:yh::)?(\w+_kernel)", name)
    return m.group(1) if m else name[:70]


def launches(path, out):
    rows = [r for r in csv.reader(open(path)) if len(r) > 10]
    hdr = rows[0]
    ki, vi = hdr.index("Kernel Name"), hdr.index("Metric Value")
    d = collections.OrderedDict()
    for r in rows[1:]:
        d.setdefault(short(r[ki]), []).append(float(r[vi].replace(",", "")))
    ours = {k: v for k, v in d.items() if k.endswith("_kernel") and "elementwise" not in k}
    tot = sum(sum(v) for v in ours.values())
    with open(out, "w") as f:
        f.write(f"# ncu launch list summary ({os.path.basename(path)})\n\n")
        f.write("`ncu --metrics gpu__time_duration.sum --clock-control none` — cold-cache, serialised:"
                " compare SHARES.\n\n| kernel | launches | mean us | share of yolohot device time |\n|---|---|---|---|\n")
        for k, v in sorted(ours.items(), key=lambda kv: -sum(kv[1])):
            f.write(f"| {k} | {len(v)} | {sum(v) / len(v) / 1e3:.2f} | {sum(v) / tot:.3f} |\n")
        f.write("\nOther (torch, input generation / bench bookkeeping outside the timed region):\n\n")
        for k, v in d.items():
            if k not in ours:
                f.write(f"- {k}: {len(v)} launches, mean {sum(v) / len(v) / 1e3:.2f} us\n")


def full(rep, out_csv, traffic_path, tag, command, commit):
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, units = rows[0], rows[1]
    cols = ["Kernel Name"] + [m for m in METRICS if m in hdr]
    idx = [hdr.index(c) for c in cols]
    traffic = {}
    try:
        traffic = json.load(open(traffic_path))
    except Exception:
        pass
    per = collections.defaultdict(list)
    with open(out_csv, "w", newline="") as f:
        w = csv.writer(f)
        w.writerow(cols)
        w.writerow([units[i] for i in idx])
        for r in rows[2:]:
            w.writerow([short(r[i]) if c == "Kernel Name" else r[i] for c, i in zip(cols, idx)])
            def val(name):
                i = hdr.index(name)
                v = float(r[i].replace(",", ""))
                u = units[i].lower()
                return v * {"byte": 1, "kbyte": 1e3, "mbyte": 1e6, "gbyte": 1e9}.get(u, 1)
            per[short(r[hdr.index("Kernel Name")])].append(val("dram__bytes_read.sum") + val("dram__bytes_write.sum"))
    for k, v in per.items():
        traffic[k] = sum(v) / len(v)
    # provenance: bench.py prints it as roofline.traffic_source, so that the DRAM-byte figure of
    # the bench line can be traced to the capture (and the source state) it came from
    traffic["_source"] = {"capture": f"profiles/{tag}_ncu_full.csv", "commit": commit, "command": command,
                          "what": "dram__bytes_read.sum + dram__bytes_write.sum per launch, mean over the "
                                  "captured launches of each kernel"}
    json.dump(traffic, open(traffic_path, "w"), indent=1, sort_keys=True)


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("tag")
    ap.add_argument("--launches")
    ap.add_argument("--rep")
    ap.add_argument("--command", default="ncu --set full --clock-control none --import-source on -k regex:<kernels> "
                                         "python bench.py --steps 3 --warmup 3 --no-e2e --no-cpu-baseline --no-extras")
    ap.add_argument("--commit", default=None)
    a = ap.parse_args()
    if a.commit is None:
        a.commit = subprocess.run(["git", "-C", ROOT, "rev-parse", "--short", "HEAD"], capture_output=True,
                                  text=True).stdout.strip()
    os.makedirs(os.path.join(ROOT, "profiles"), exist_ok=True)
    if a.launches:
        launches(a.launches, os.path.join(ROOT, "profiles", f"{a.tag}_launches.md"))
    if a.rep:
        full(a.rep, os.path.join(ROOT, "profiles", f"{a.tag}_ncu_full.csv"),
             os.path.join(ROOT, "profiles", "traffic.json"), a.tag, a.command, a.commit)
