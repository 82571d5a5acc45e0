#!/bin/bash
# A/B of the coherence sort on the GPU box: kernel-class times of book2_final with B200PT_SORT=0 / 1
mkdir -p gpurun_out
python -m pytest tests/test_gpu_render.py tests/test_gpu_trace.py -q -x > gpurun_out/ab_pytest.txt 2>&1; tail -3 gpurun_out/ab_pytest.txt
for v in 0 1; do
  B200PT_SORT=$v python bench.py --steps 4 --warmup 3 --no-cpu-baseline > gpurun_out/ab_sort$v.json 2> gpurun_out/ab_sort$v.err
  python - <<PY
import json
j=json.loads(open("gpurun_out/ab_sort$v.json").read().strip().splitlines()[-1])
print("SORT=$v", round(j["value"],1), "Mpaths/s e2e", round(j["e2e"]["value"],1), j["roofline"]["kernel_ms"])
PY
done
