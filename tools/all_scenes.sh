#!/bin/bash
# Run on the GPU box: one bench line per BASELINE config (bench.py --scene ...), each with the reference's CPU baseline on
# the box's host cores and its 1-thread rate -> gpurun_out/<tag>_bench_<scene>.json
TAG=${1:-r2}
OUT=gpurun_out; mkdir -p $OUT
for scene in book2_final cornell_box cornell_box_small_lights cornell_box_with_volume three_orbs; do
  timeout 420 python bench.py --scene $scene --steps 8 --warmup 3 --cpu-seconds 12 --cpu-one-thread > $OUT/${TAG}_bench_$scene.json 2> $OUT/${TAG}_bench_$scene.err
  python - <<PY
import json
try:
    j=json.loads(open("$OUT/${TAG}_bench_$scene.json").read().strip().splitlines()[-1])
    c=j.get("cpu_baseline") or {}
    print("$scene", round(j["value"],1), "Mpaths/s", round(j["mrays_per_s"]), "Mrays/s (traced", round(j["rays"]["mrays_traced_per_s"]), ") e2e", round(j["e2e"]["value"],1), "| cpu", c.get("value"), c.get("cores"), (c.get("one_thread") or {}).get("value"), "| ms", j["roofline"]["kernel_ms"])
except Exception as e:
    print("$scene failed", e, open("$OUT/${TAG}_bench_$scene.err").read()[-400:])
PY
done
