#!/bin/bash
# For the default library and every var_*.so: the multi-stream parity tests, then (for the ones that pass) the pass-1 bench.
cd "$(dirname "$0")/.."
out=${1:-gpurun_out/matrix}
: > $out.txt
for lib in pointcloudprocessor_b200/libpcp_b200.so pointcloudprocessor_b200/var_*.so; do
  [ -f "$lib" ] || continue
  echo "=== $lib" >> $out.txt
  PCP_LIB_PATH=$PWD/$lib timeout 200 python -m pytest tests/test_gpu_parity.py -k "frames_in_flight or async_pipeline or c2_shape or tiny" -x -q 2>&1 | tail -3 >> $out.txt
done
tools/bench_variants.sh $out.jsonl --frames 40 --reps 3 >> $out.txt 2>&1
cat $out.txt
