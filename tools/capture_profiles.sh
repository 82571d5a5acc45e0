#!/bin/bash
# Run on the GPU box (gpurun -- 'bash tools/capture_profiles.sh r1k'): the bench line, the ncu launch list of the same
# command and `--set full` captures of the dominant kernels, all into gpurun_out/<tag>_*.
# tools/ncu_summary.py / tools/ncu_lines.py turn them into the text summaries committed under profiles/.
set -u
TAG=${1:-r1x}
OUT=gpurun_out
mkdir -p $OUT
BENCH="python bench.py --steps 2 --warmup 3 --no-cpu-baseline"
python bench.py > $OUT/${TAG}_bench_cornell_box.json 2> $OUT/${TAG}_bench.err || exit 1
python bench.py --scene book2_final --no-cpu-baseline > $OUT/${TAG}_bench_book2_final.json 2>> $OUT/${TAG}_bench.err
$BENCH > /dev/null 2>&1 || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 2000 --csv --log-file $OUT/${TAG}_launches_cornell_box.csv \
    $BENCH > $OUT/ncu_launches.log 2>&1
# the three k_nee launches of bounces 1-3 of a warm batch (70 launches per batch: 10 x (extend, 5 shade, nee) + ...)
ncu --set full --clock-control none --import-source on -k regex:k_nee -s 21 -c 3 -f -o $OUT/${TAG}_nee_cornell_box \
    $BENCH > $OUT/ncu_nee.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_extend -s 21 -c 3 -f -o $OUT/${TAG}_extend_cornell_box \
    $BENCH > $OUT/ncu_ext.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_shade -s 181 -c 1 -f -o $OUT/${TAG}_shade_cornell_box \
    $BENCH > $OUT/ncu_shade.log 2>&1
B2="python bench.py --scene book2_final --steps 1 --warmup 3 --no-cpu-baseline"
ncu --set full --clock-control none --import-source on -k regex:k_extend -s 11 -c 2 -f -o $OUT/${TAG}_extend_book2_final \
    $B2 > $OUT/ncu_ext2.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_nee -s 11 -c 2 -f -o $OUT/${TAG}_nee_book2_final \
    $B2 > $OUT/ncu_nee2.log 2>&1
ls -la $OUT
