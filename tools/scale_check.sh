#!/bin/bash
# Run on an N-GPU box (gpurun --gpus 8): strong-scaling bench lines at N = 1, 2, 4, 8 (torchrun, NCCL), the device tests on
# real GPUs, and the north-star frame (book2_final 3840x2160 x 1024 spp) through the drop-in CLI with "devices": N.
TAG=${1:-r2}
OUT=gpurun_out; mkdir -p $OUT
NG=$(nvidia-smi -L | grep -c '^GPU ')
echo "GPUs: $NG"
python -m pytest tests/test_gpu_lifecycle_and_devices.py -q > $OUT/${TAG}_pytest_${NG}gpu.txt 2>&1; tail -2 $OUT/${TAG}_pytest_${NG}gpu.txt
for N in 1 2 4 8; do
  [ $N -gt $NG ] && continue
  if [ $N -eq 1 ]; then
    python bench.py --steps 8 --warmup 3 --no-cpu-baseline > $OUT/${TAG}_scale_book2_final_$N.json 2> $OUT/${TAG}_scale_$N.err
  else
    python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2951$N bench.py --gpus $N --steps 8 --warmup 3 > $OUT/${TAG}_scale_book2_final_$N.json 2> $OUT/${TAG}_scale_$N.err
  fi
done
python - <<PY
import json
base=None
for n in (1,2,4,8):
    try:
        j=json.loads(open(f"$OUT/${TAG}_scale_book2_final_{n}.json").read().strip().splitlines()[-1])
    except Exception as e:
        print(n, "missing", e); continue
    base = base or j["value"]
    print(n, round(j["value"],1), "Mpaths/s  eff", round(j["value"]/(n*base),4), " e2e", round(j["e2e"]["value"],1), j["scaling"], (j.get("film_check") or {}).get("ok"))
PY
# the north-star frame through the C library alone
mkdir -p $OUT/frame && cd $OUT/frame
for D in 1 $NG; do
python - <<PY
import json, sys
sys.path.insert(0, "../..")
import pathtrace_b200 as pt
c = pt.default_config_dict(width=3840, height=2160, samples=1024, scene="../../scenes/book2_final.json", ppm_output_path=f"frame_$D.ppm", traced_paths_output_path="p.txt", traced_paths_2d_output_path="p2.txt", devices=$D)
json.dump(c, open("config.json", "w"))
PY
T0=$(date +%s.%N)
../../pathtrace_b200/b200pt_main > cli_$D.txt 2>&1
echo "wall $(python -c "import time,sys; print(round(time.time()-$T0,2))") s (process start to exit, incl. scene build and PPM writes)" >> cli_$D.txt
grep -E "GPU\(s\)|computed|device time|wall" cli_$D.txt
done
python - <<PY
import numpy as np
def ppm(p):
    d=open(p,"rb").read(); h=b"P6\n3840 2160\n255\n"; assert d.startswith(h); return np.frombuffer(d[len(h):],dtype=np.uint8).astype(np.int32)
a,b=ppm("frame_1.ppm"),ppm("frame_$NG.ppm"); d=np.abs(a-b)
print("frame 1 GPU vs $NG GPUs: max byte diff", d.max(), "differing", float((d>0).mean()))
PY
rm -f frame_*.ppm
