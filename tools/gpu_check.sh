#!/bin/bash
# One GPU-box call: the GPU parity tests, then short bench lines of the two scenes the round is judged on.
# usage (from the repo root, under gpurun): bash tools/gpu_check.sh [tag]
TAG=${1:-check}
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/${TAG}_gpus.txt 2>&1
python -m pytest tests -m gpu -q > gpurun_out/${TAG}_pytest_gpu.txt 2>&1
echo "pytest rc=$?" >> gpurun_out/${TAG}_pytest_gpu.txt
tail -5 gpurun_out/${TAG}_pytest_gpu.txt
for scene in book2_final cornell_box; do
  python bench.py --scene $scene --steps 8 --warmup 3 --no-cpu-baseline > gpurun_out/${TAG}_bench_${scene}.json 2> gpurun_out/${TAG}_bench_${scene}.err
  echo "bench $scene rc=$?"; python - <<PY
import json
try:
    j=json.loads(open("gpurun_out/${TAG}_bench_${scene}.json").read().strip().splitlines()[-1])
    print("$scene", round(j["value"],1), "Mpaths/s e2e", round(j["e2e"]["value"],1), j["roofline"].get("kernel_ms"), "frac", j["roofline"]["frac"], j["roofline"]["bound_valid"], "traced share", j["rays"]["shadow_traced_share"])
except Exception as e:
    print("bench parse failed", e)
PY
done
