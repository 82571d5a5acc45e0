#!/bin/bash
# A/B of library builds on the GPU box: tools/ab.sh <tag> <steps> <scene> lib1.so lib2.so ...   (each lib is benched twice, interleaved)
TAG=$1; STEPS=$2; SCENE=$3; shift 3
OUT=gpurun_out; mkdir -p $OUT
for rep in 1 2; do
  for lib in "$@"; do
    name=$(basename $lib .so)
    B200PT_LIB=$PWD/$lib python bench.py --scene $SCENE --steps $STEPS --warmup 3 --no-cpu-baseline > $OUT/${TAG}_${SCENE}_${name}_$rep.json 2> $OUT/${TAG}_${SCENE}_${name}_$rep.err || echo "FAILED $lib"
    python - <<PY
import json
try:
    j=json.loads(open("$OUT/${TAG}_${SCENE}_${name}_$rep.json").read().strip().splitlines()[-1])
    k=j["roofline"]["kernel_ms"]
    print("$name rep $rep: %.1f Mpaths/s  e2e %.1f  ms/step %.2f  kernels %s" % (j["value"], j["e2e"]["value"], j["ms_per_step"], {a: round(b,1) for a,b in k.items()}))
except Exception as e:
    print("$name rep $rep: no line", e)
PY
  done
done
