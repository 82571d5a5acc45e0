#!/bin/bash
# Run on the GPU box: GPU tests, tools/capture_profiles.sh <tag>, then the text summaries of the ncu reports (what is
# committed under profiles/); the .ncu-rep files of the cornell_box extend/shade captures are dropped afterwards so
# that gpurun_out/ stays below the 64 MiB that is copied back.
TAG=${1:-r1x}
OUT=gpurun_out
mkdir -p $OUT
(timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -5) > $OUT/${TAG}_pytest_gpu.log
timeout 120 python __graft_entry__.py smoke > $OUT/${TAG}_smoke.log 2>&1
timeout 900 bash tools/capture_profiles.sh $TAG > $OUT/${TAG}_capture.log 2>&1
for k in nee_cornell_box extend_cornell_box shade_cornell_box nee_book2_final extend_book2_final; do
    python tools/ncu_summary.py full $OUT/${TAG}_$k.ncu-rep > $OUT/${TAG}_${k}_full.txt 2>> $OUT/${TAG}_summary.err
done
python tools/ncu_summary.py launches $OUT/${TAG}_launches_cornell_box.csv > $OUT/${TAG}_launches_cornell_box.txt
rm -f $OUT/${TAG}_extend_cornell_box.ncu-rep $OUT/${TAG}_shade_cornell_box.ncu-rep $OUT/${TAG}_extend_book2_final.ncu-rep
for s in three_orbs cornell_box_small_lights cornell_box_with_volume; do
    python bench.py --scene $s --no-cpu-baseline > $OUT/${TAG}_bench_$s.json 2>/dev/null
done
cat $OUT/${TAG}_pytest_gpu.log $OUT/${TAG}_smoke.log
du -sh $OUT
