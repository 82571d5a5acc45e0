#!/usr/bin/env python3
"""Formats the ncu CSV of tools/batch_profile.sh: one row per kernel launch of a warm wavefront batch in launch order,
then totals per kernel."""
import collections
import csv
import re
import sys

scene, path = sys.argv[1:3]
json_out = sys.argv[3] if len(sys.argv) > 3 else None
with open(path) as f:
    lines = [l for l in f if not l.startswith("==")]
rows = collections.OrderedDict()
for r in csv.DictReader(lines):
    d = rows.setdefault(int(r["ID"]), {"kernel": r["Kernel Name"]})
    try:
        v = float(r["Metric Value"].replace(",", ""))
    except ValueError:
        v = 0.0
    u = r["Metric Unit"]
    if r["Metric Name"] == "gpu__time_duration.sum":
        v *= {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}.get(u, 1.0)
    if r["Metric Name"].startswith("dram__bytes"):
        v *= {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(u, 1.0)
    d[r["Metric Name"]] = v


def short(name):
    name = re.sub(r"\(.*", "", name).replace("void ", "")
    name = re.sub(r"pt::\(anonymous namespace\)::|pt::<unnamed>::|<unnamed>::", "", name)
    return name.strip()


print(f"# {scene}: every kernel launch of one warm wavefront batch, in launch order; ncu metrics pass (cold-cache, serialised)")
print(f"{'id':>4} {'kernel':34s} {'time_us':>9} {'lanes/32':>8} {'issue%':>7} {'L2hit%':>7} {'Mwarp-inst':>10} {'dramMB':>9} {'GB/s':>7}")
agg = collections.OrderedDict()
for i, (k, d) in enumerate(rows.items()):
    name = short(d["kernel"])[:34]
    t = d.get("gpu__time_duration.sum", 0.0)
    by = d.get("dram__bytes_read.sum", 0.0) + d.get("dram__bytes_write.sum", 0.0)
    lanes = d.get("smsp__thread_inst_executed_per_inst_executed.ratio", 0.0)
    inst = d.get("smsp__inst_executed.sum", 0.0)
    print(f"{i:4d} {name:34s} {t:9.1f} {lanes:8.2f} {d.get('smsp__issue_active.avg.pct_of_peak_sustained_active', 0):7.1f} "
          f"{d.get('lts__t_sector_hit_rate.pct', 0):7.1f} {inst / 1e6:10.1f} {by / 1e6:9.1f} {by / max(t, 1e-9) / 1e3:7.0f}")
    a = agg.setdefault(re.sub(r"<.*", "", name), [0, 0.0, 0.0, 0.0, 0.0, 0.0])
    a[0] += 1
    a[1] += t
    a[2] += by
    a[3] += inst
    a[4] += inst * lanes
    a[5] += t * d.get("smsp__issue_active.avg.pct_of_peak_sustained_active", 0.0)
tot = sum(a[1] for a in agg.values())
print(f"\n{'kernel':20s} {'launches':>8s} {'total_us':>10s} {'share':>6s} {'dramMB':>9s} {'GB/s':>7s} {'lanes/32':>8s}")
for k, a in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"{k:20s} {a[0]:8d} {a[1]:10.1f} {a[1] / tot:6.3f} {a[2] / 1e6:9.1f} {a[2] / max(a[1], 1e-9) / 1e3:7.0f} {a[4] / max(a[3], 1e-9):8.2f}")
print(f"{'TOTAL':20s} {sum(a[0] for a in agg.values()):8d} {tot:10.1f}")
if json_out:
    import json
    out = {}
    for k, a in agg.items():
        name = k.split("::")[-1]
        out[name] = {"launches": a[0], "total_us": round(a[1], 1), "share_of_batch": round(a[1] / tot, 4),
                     "dram_bytes_per_launch": round(a[2] / a[0]), "dram_gbs": round(a[2] / max(a[1], 1e-9) / 1e3, 1),
                     "issue_active_pct": round(a[5] / max(a[1], 1e-9), 2),      # time-weighted over the launches
                     "lanes_per_warp_inst": round(a[4] / max(a[3], 1e-9), 2)}  # instruction-weighted
    with open(json_out, "w") as f:
        json.dump(out, f, indent=1)
