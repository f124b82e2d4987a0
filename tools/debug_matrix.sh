#!/bin/bash
# For the default library and every var_*.so: the parity tests that exercise pass 1 (tiny, shapes, edge points, slots, cull drop-in,
# candidate counts), then the pass-1 bench at C2 and C1.
cd "$(dirname "$0")/.."
out=${1:-gpurun_out/matrix}
: > $out.txt
for lib in pointcloudprocessor_b200/libpcp_b200.so pointcloudprocessor_b200/var_*.so; do
  [ -f "$lib" ] || continue
  echo "=== $lib" >> $out.txt
  PCP_LIB_PATH=$PWD/$lib timeout 300 python -m pytest tests/test_gpu_parity.py tests/test_gpu_golden.py -k "not nid and not hsv and not calib" -x -q 2>&1 | tail -3 >> $out.txt
done
tools/bench_variants.sh $out.jsonl --frames 40 --reps 3 >> $out.txt 2>&1
tools/bench_variants.sh ${out}_C1.jsonl --workload C1 --frames 20 --reps 30 >> $out.txt 2>&1
cat $out.txt
