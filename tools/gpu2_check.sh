#!/bin/bash
# On a 2-GPU box: the device/lifecycle tests across real devices, the NCCL test, and bench lines at N = 1 and 2
mkdir -p gpurun_out
nvidia-smi -L
python -m pytest tests/test_gpu_lifecycle_and_devices.py -q > gpurun_out/g2_pytest.txt 2>&1; tail -5 gpurun_out/g2_pytest.txt
python bench.py --steps 4 --warmup 3 --no-cpu-baseline > gpurun_out/g2_bench_n1.json 2> gpurun_out/g2_bench_n1.err
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 4 --warmup 3 > gpurun_out/g2_bench_n2.json 2> gpurun_out/g2_bench_n2.err
python - <<PY
import json
for n in (1,2):
    try:
        j=json.loads(open(f"gpurun_out/g2_bench_n{n}.json").read().strip().splitlines()[-1])
        print(n, round(j["value"],1), "e2e", round(j["e2e"]["value"],1), j["scaling"], j.get("film_check"))
    except Exception as e:
        print(n, "failed", e, open(f"gpurun_out/g2_bench_n{n}.err").read()[-600:])
PY
