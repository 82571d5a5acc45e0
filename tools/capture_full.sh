#!/bin/bash
# Run on the GPU box: `ncu --set full` captures (with source) of one warm bounce-1 launch of k_extend / k_nee / k_shade
# usage: bash tools/capture_full.sh <tag> <scene> "<kernels>"
TAG=${1:-r2x}; SCENE=${2:-book2_final}; KERNELS=${3:-"k_extend k_nee"}
OUT=gpurun_out; mkdir -p $OUT
CMD="python bench.py --scene $SCENE --steps 1 --warmup 3 --no-cpu-baseline --spp-per-step ${SPP:-8}"
$CMD > /dev/null 2>&1 || { echo "bench failed"; exit 1; }
for K in $KERNELS; do
  SKIP=11; [ "$K" = "k_shade" ] && SKIP=${SHADE_SKIP:-56}
  ncu --set full --clock-control none --import-source on -k regex:$K -s $SKIP -c 1 -f -o $OUT/${TAG}_${K}_$SCENE $CMD > $OUT/ncu_full_$K.log 2>&1
  ls -la $OUT/${TAG}_${K}_$SCENE.ncu-rep
done
