#!/usr/bin/env python3
"""profiles/ncu_metrics.json from the per-scene JSONs of tools/batch_profile.sh (ncu metrics pass over every launch of one warm
wavefront batch): per scene and kernel class the measured DRAM bytes per launch, issue-slot use and active lanes, with the
commit they were taken at.  bench.py cites these in `roofline` (traffic, issue_bound).

usage: python tools/make_ncu_metrics.py <tag> scene=spp_per_batch [scene=spp ...]"""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
tag = sys.argv[1]
scenes = {}
for arg in sys.argv[2:]:
    scene, spp = arg.split("=")
    with open(os.path.join(ROOT, "gpurun_out", f"{tag}_batch_{scene}.json")) as f:
        scenes[scene] = dict(json.load(f), batch_spp=int(spp))
commit = subprocess.run(["git", "rev-parse", "--short", "HEAD"], cwd=ROOT, capture_output=True, text=True).stdout.strip()
dirty = bool(subprocess.run(["git", "status", "--porcelain", "--", "pathtrace_b200"], cwd=ROOT, capture_output=True,
                            text=True).stdout.strip())
out = {"commit": commit + ("+uncommitted kernel changes" if dirty else ""), "tag": tag,
       "how": "tools/batch_profile.sh: ncu --metrics gpu__time_duration.sum,smsp__issue_active...,dram__bytes_read/write.sum,... "
              "--clock-control none over every kernel launch of ONE warm wavefront batch of bench.py's own command; "
              "per kernel class: DRAM bytes per launch (mean), issue-slot use (time-weighted), lanes per warp instruction",
       "scenes": scenes}
with open(os.path.join(ROOT, "profiles", "ncu_metrics.json"), "w") as f:
    json.dump(out, f, indent=1)
print(json.dumps({s: {k: v for k, v in d.items() if k in ("k_nee", "k_extend")} for s, d in scenes.items()}, indent=1))
