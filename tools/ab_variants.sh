#!/bin/bash
# A/B of library variants on the GPU box: bench.py kernel-class times per variant (B200PT_LIB selects the build)
# usage: bash tools/ab_variants.sh "<variant suffixes, '' = default>" [scene] [extra env assignments]
mkdir -p gpurun_out
SCENE=${2:-book2_final}
for v in $1; do
  [ "$v" = "default" ] && lib=pathtrace_b200/libb200pt.so || lib=pathtrace_b200/libb200pt_$v.so
  env $3 B200PT_LIB=$PWD/$lib python bench.py --scene $SCENE --steps 4 --warmup 3 --no-cpu-baseline > gpurun_out/ab_$v.json 2> gpurun_out/ab_$v.err
  python - <<PY
import json
try:
    j=json.loads(open("gpurun_out/ab_$v.json").read().strip().splitlines()[-1])
    print("$v", round(j["value"],1), "Mpaths/s e2e", round(j["e2e"]["value"],1), j["roofline"]["kernel_ms"])
except Exception as e:
    print("$v failed", e, open("gpurun_out/ab_$v.err").read()[-500:])
PY
done
