#!/bin/bash
# Run on the GPU box: per-bounce ncu metrics (time, active lanes, issue-slot use, L1 / L2 hit rates, warp instructions)
# of every k_extend / k_nee launch of one warm wavefront batch -> gpurun_out/<tag>_per_bounce_<scene>.txt
TAG=${1:-r1x}
OUT=gpurun_out
mkdir -p $OUT
M=gpu__time_duration.sum,smsp__thread_inst_executed_per_inst_executed.ratio,smsp__issue_active.avg.pct_of_peak_sustained_active,l1tex__t_sector_hit_rate.pct,lts__t_sector_hit_rate.pct,smsp__inst_executed.sum
for SCENE in cornell_box book2_final; do
    # launches 0-19 are the first (cold) batch; 20-39 the second: 10 bounces x (k_extend, k_nee)
    ncu --metrics $M --clock-control none -k regex:'k_extend|k_nee' -s 20 -c 20 --csv --log-file $OUT/${TAG}_pb_$SCENE.csv \
        python bench.py --scene $SCENE --steps 1 --warmup 3 --no-cpu-baseline > $OUT/ncu_pb_$SCENE.log 2>&1
    python tools/per_bounce_table.py $SCENE $OUT/${TAG}_pb_$SCENE.csv > $OUT/${TAG}_per_bounce_$SCENE.txt
    cat $OUT/${TAG}_per_bounce_$SCENE.txt
done
