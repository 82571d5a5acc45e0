#!/usr/bin/env python3
"""Turns ncu outputs brought back in gpurun_out/ into the small text summaries committed under profiles/.

  launches <launches.csv>            per-kernel launch count, total time and share (gpu__time_duration.sum pass)
  full <report.ncu-rep> [kernel-id]  key metrics of a `--set full` capture (+ stall-reason and opcode mix from SASS)
"""
import collections
import csv
import re
import subprocess
import sys


def launches(path):
    with open(path) as f:
        lines = [l for l in f if not l.startswith("==")]
    agg = collections.defaultdict(lambda: [0, 0.0])
    for row in csv.DictReader(lines):
        name = re.sub(r"\(.*", "", row["Kernel Name"]).replace("void ", "").replace("unnamed>::", "")
        v = float(row["Metric Value"].replace(",", ""))
        v *= {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}.get(row["Metric Unit"], 1.0)
        agg[name][0] += 1
        agg[name][1] += v
    tot = sum(v[1] for v in agg.values())
    print(f"{'kernel':28s} {'launches':>8s} {'total_us':>12s} {'share':>7s} {'avg_us':>10s}")
    for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"{k[:28]:28s} {v[0]:8d} {v[1]:12.1f} {v[1] / tot:7.3f} {v[1] / v[0]:10.1f}")
    print(f"{'TOTAL':28s} {sum(v[0] for v in agg.values()):8d} {tot:12.1f}")


KEYS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "lts__t_sector_hit_rate.pct",
        "l1tex__t_sector_hit_rate.pct", "launch__registers_per_thread", "launch__grid_size",
        "launch__occupancy_limit_registers", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "smsp__thread_inst_executed_per_inst_executed.ratio", "smsp__inst_executed.sum",
        "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "lts__t_bytes.sum",
        "sass__inst_executed_local_loads", "sass__inst_executed_local_stores", "sm__cycles_elapsed.avg.per_second"]


def full(path):
    raw = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, units, data = rows[0], rows[1], rows[2:]
    names = [r[hdr.index("Kernel Name")][:60] for r in data]
    print("launches captured:", len(data))
    for n in names:
        print("   ", n)
    for key in KEYS:
        if key in hdr:
            i = hdr.index(key)
            print(f"{key:62s} {units[i]:14s} " + "  ".join(r[i] for r in data))
    sass = subprocess.run(["ncu", "-i", path, "--page", "source", "--csv", "--print-source", "sass"],
                          capture_output=True, text=True).stdout
    secs, cur = [], None
    for r in csv.reader(sass.splitlines()):
        if r and r[0] == "Kernel Name":
            cur = {"hdr": None, "rows": []}
            secs.append(cur)
        elif cur is not None and r and r[0] == "Address":
            cur["hdr"] = r
        elif cur is not None and cur["hdr"] and len(r) == len(cur["hdr"]):
            cur["rows"].append(r)
    if not secs:
        return
    s = secs[0]
    idx = {n: i for i, n in enumerate(s["hdr"])}
    tot = sum(int(r[idx["# Samples"]]) for r in s["rows"])
    inst = sum(int(r[idx["Instructions Executed"]]) for r in s["rows"])
    print(f"\nfirst captured launch: {len(s['rows'])} SASS instructions, {inst} warp instructions executed, "
          f"{tot} stall samples")
    stalls = {n: sum(int(r[idx[n]]) for r in s["rows"]) for n in s["hdr"]
              if n.startswith("stall_") and "Not Issued" not in n}
    print("stall reasons (share of samples): " +
          ", ".join(f"{k[6:]} {v / tot:.3f}" for k, v in sorted(stalls.items(), key=lambda kv: -kv[1])[:8]))
    op = collections.Counter()
    for r in s["rows"]:
        m = re.match(r"\s*(@!?U?P\d+\s+)?([A-Z0-9_.]+)", r[idx["Source"]])
        op[m.group(2).split(".")[0] if m else "?"] += int(r[idx["Instructions Executed"]])
    print("opcode mix (share of executed warp instructions): " + ", ".join(f"{k} {v / inst:.3f}" for k, v in op.most_common(14)))


if __name__ == "__main__":
    {"launches": launches, "full": full}[sys.argv[1]](sys.argv[2])
