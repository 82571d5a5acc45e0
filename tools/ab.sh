#!/bin/bash
# A/B of library variants on the GPU box: tools/ab.sh <scene> <variant suffixes...>  ("-" = the default build)
# Variants are built with `make -C pathtrace_b200/csrc VARIANT=_x EXTRA=-D...`; each is benched twice, interleaved.
SCENE=$1; shift
mkdir -p gpurun_out
for rep in 1 2; do
  for v in "$@"; do
    sfx=$([ "$v" = "-" ] && echo "" || echo "_$v")
    B200PT_LIB=pathtrace_b200/libb200pt$sfx.so python bench.py --scene $SCENE --steps 8 --warmup 3 --no-cpu-baseline \
        > gpurun_out/ab_${SCENE}_${v}_$rep.json 2> gpurun_out/ab_err.log
    python - "$SCENE" "$v" "$rep" <<'P'
import json, sys
sc, v, rep = sys.argv[1:]
d = json.loads(open(f"gpurun_out/ab_{sc}_{v}_{rep}.json").read().strip().splitlines()[-1])
r = d["roofline"]
print(sc, v, rep, round(d["value"], 1), "e2e", round(d["e2e"]["value"], 1), r["kernel_ms"], "boxes/ray", round(r["boxes_per_ray"], 2))
P
  done
done
