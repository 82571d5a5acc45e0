#!/bin/bash
# Run on the GPU box: compute-sanitizer memcheck, racecheck and initcheck over tools/sanitize_run.py (every kernel variant at
# CI size) -> gpurun_out/<tag>_sanitizer_<tool>.txt.  SURVEY.md 5: "memcheck/racecheck on every kernel in CI-sized configs".
TAG=${1:-r2x}
OUT=gpurun_out; mkdir -p $OUT
python tools/sanitize_run.py > $OUT/${TAG}_sanitize_plain.txt 2>&1 || { echo "workload fails without the sanitizer"; tail -5 $OUT/${TAG}_sanitize_plain.txt; exit 1; }
for TOOL in memcheck racecheck initcheck; do
  timeout 1500 compute-sanitizer --tool $TOOL --error-exitcode 9 --print-limit 20 python tools/sanitize_run.py > $OUT/${TAG}_sanitizer_$TOOL.txt 2>&1
  echo "$TOOL rc=$?"; grep -E "ERROR SUMMARY|RACECHECK SUMMARY|sanitize_run ok" $OUT/${TAG}_sanitizer_$TOOL.txt | tail -3
done
