#!/bin/bash
# Run on the GPU box.  compute-sanitizer is closed on this pool, so this is our own memcheck: the workload of
# tools/sanitize_run.py (every kernel variant at CI size) on the -DPT_DEBUG_BOUNDS build, where every node, instance, stack,
# queue, slot and sort-bucket index is range-checked on the device and a violation traps (= a CUDA error = a failed run);
# and a race check by determinism: the same workload's films must be bit-identical run to run.
TAG=${1:-r2}
OUT=gpurun_out; mkdir -p $OUT
make -C pathtrace_b200/csrc VARIANT=_dbg EXTRA=-DPT_DEBUG_BOUNDS > /dev/null 2>&1
B200PT_LIB=$PWD/pathtrace_b200/libb200pt_dbg.so python tools/sanitize_run.py > $OUT/${TAG}_bounds_check.txt 2>&1; echo "bounds-checked build rc=$?" >> $OUT/${TAG}_bounds_check.txt
B200PT_LIB=$PWD/pathtrace_b200/libb200pt_dbg.so python tools/variants_smoke.py >> $OUT/${TAG}_bounds_check.txt 2>&1; echo "variants_smoke rc=$?" >> $OUT/${TAG}_bounds_check.txt
python tools/determinism.py >> $OUT/${TAG}_bounds_check.txt 2>&1; echo "determinism rc=$?" >> $OUT/${TAG}_bounds_check.txt
tail -25 $OUT/${TAG}_bounds_check.txt
