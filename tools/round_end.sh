#!/bin/bash
# What the driver does at round end, in one GPU-box call: pytest -m gpu -x, smoke(), bench.py (reference arm first, then ours)
# with its default flags.  usage: bash tools/round_end.sh [tag]
TAG=${1:-r2}
OUT=gpurun_out; mkdir -p $OUT
S0=$SECONDS
python -m pytest tests -x -q -m gpu > $OUT/${TAG}_pytest_gpu.txt 2>&1; echo "pytest rc=$? ($((SECONDS-S0)) s)"; tail -2 $OUT/${TAG}_pytest_gpu.txt
S0=$SECONDS
python -c "import __graft_entry__ as g; g.smoke()" > $OUT/${TAG}_smoke.txt 2>&1; echo "smoke rc=$? ($((SECONDS-S0)) s)"; tail -1 $OUT/${TAG}_smoke.txt
S0=$SECONDS
python bench.py --impl reference > $OUT/${TAG}_bench_reference.json 2> $OUT/${TAG}_bench_reference.err; echo "reference arm rc=$? ($((SECONDS-S0)) s)"
S0=$SECONDS
python bench.py > $OUT/${TAG}_bench_default.json 2> $OUT/${TAG}_bench_default.err; echo "bench rc=$? ($((SECONDS-S0)) s)"
python - <<PY
import json
for name in ("reference", "default"):
    try:
        j=json.loads(open(f"$OUT/${TAG}_bench_{name}.json").read().strip().splitlines()[-1])
        print(name, {k: j.get(k) for k in ("impl","metric","value","unit","n_gpus","steps","warmup","ms_per_step","scaling","dtype","gpu_launches")})
        print("   config", j.get("config")); print("   e2e", j.get("e2e")); print("   cpu", j.get("cpu_baseline"))
        if name == "default":
            r=j["roofline"]; print("   roofline", {k:r[k] for k in ("bound","kernel","achieved","peak","unit","frac","bound_valid","traffic","traffic_source")}, r["issue_bound"]["frac"]); print("   clocks", j["clocks"])
    except Exception as e:
        print(name, "FAILED", e, open(f"$OUT/${TAG}_bench_{name}.err").read()[-600:])
PY
