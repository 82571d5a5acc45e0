#!/bin/bash
# Run on the GPU box: ncu metrics (time, active lanes, issue-slot use, DRAM bytes, L2 hit rate, warp instructions) of EVERY
# kernel launch of one warm wavefront batch, in launch order -> gpurun_out/<tag>_batch_<scene>.{csv,txt}
# usage: bash tools/batch_profile.sh <tag> <scene> [launches per batch, default: counted from a dry run]
TAG=${1:-r2x}
SCENE=${2:-book2_final}
OUT=gpurun_out
mkdir -p $OUT
M=gpu__time_duration.sum,smsp__thread_inst_executed_per_inst_executed.ratio,smsp__issue_active.avg.pct_of_peak_sustained_active,lts__t_sector_hit_rate.pct,smsp__inst_executed.sum,dram__bytes_read.sum,dram__bytes_write.sum
CMD="python bench.py --scene $SCENE --steps 1 --warmup 3 --no-cpu-baseline --spp-per-step ${SPP:-8}"
$CMD > $OUT/${TAG}_batchprof_bench_$SCENE.json 2> $OUT/${TAG}_batchprof_bench_$SCENE.err || { echo "bench failed"; exit 1; }
N=${3:-$(python - <<PY
import json
j=json.loads(open("$OUT/${TAG}_batchprof_bench_$SCENE.json").read().strip().splitlines()[-1])
# launches of the timed region / batches in it
per_step = j["config"]["spp_per_step_this_rank"] // j["config"]["wavefront_batch_spp"]
print(j["gpu_launches"] // (j["steps"] * per_step))
PY
)}
echo "launches per batch: $N"
ncu --metrics $M --clock-control none -s $N -c $N --csv --log-file $OUT/${TAG}_batch_$SCENE.csv $CMD > $OUT/ncu_batch_$SCENE.log 2>&1
python tools/batch_table.py $SCENE $OUT/${TAG}_batch_$SCENE.csv $OUT/${TAG}_batch_$SCENE.json > $OUT/${TAG}_batch_$SCENE.txt
tail -40 $OUT/${TAG}_batch_$SCENE.txt
