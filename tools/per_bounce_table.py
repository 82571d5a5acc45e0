#!/usr/bin/env python3
"""Formats the ncu CSV of tools/per_bounce.sh: one row per k_extend / k_nee launch of a warm batch, in launch order."""
import collections
import csv
import re
import sys

scene, path = sys.argv[1:3]
with open(path) as f:
    lines = [l for l in f if not l.startswith("==")]
rows = collections.OrderedDict()
for r in csv.DictReader(lines):
    d = rows.setdefault(int(r["ID"]), {"kernel": r["Kernel Name"]})
    v = float(r["Metric Value"].replace(",", ""))
    if r["Metric Name"] == "gpu__time_duration.sum":
        v *= {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}.get(r["Metric Unit"], 1.0)
    d[r["Metric Name"]] = v
print(f"# {scene}: k_extend / k_nee launches of one warm batch, in launch order (= bounce order), ncu metrics pass (no --set full)")
print(f"{'id':>4} {'kernel':26s} {'bounce':>6} {'time_us':>9} {'lanes/32':>8} {'issue%':>7} {'L1hit%':>7} {'L2hit%':>7} {'Mwarp-inst':>10}")
seen = collections.Counter()
for i, (k, d) in enumerate(rows.items()):
    name = re.sub(r"\(.*", "", d["kernel"]).replace("void ", "").replace("pt::<", "").strip()[:26]
    base = "ext" if "k_extend" in name else "nee"
    b = seen[base]
    seen[base] += 1
    print(f"{i:4d} {name:26s} {b:6d} {d['gpu__time_duration.sum']:9.1f} "
          f"{d['smsp__thread_inst_executed_per_inst_executed.ratio']:8.2f} "
          f"{d['smsp__issue_active.avg.pct_of_peak_sustained_active']:7.1f} {d['l1tex__t_sector_hit_rate.pct']:7.1f} "
          f"{d['lts__t_sector_hit_rate.pct']:7.1f} {d['smsp__inst_executed.sum'] / 1e6:10.1f}")
