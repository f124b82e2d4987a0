#!/bin/bash
# Runs tools/bench_pass1.py for the default library and every pointcloudprocessor_b200/var_*.so kernel variant (PCP_LIB_PATH),
# each under its own timeout (a kernel that hangs must not hold the box).  Output: one JSON line per library.
cd "$(dirname "$0")/.."
out=${1:-gpurun_out/variants.jsonl}; shift
: > "$out"
for lib in pointcloudprocessor_b200/libpcp_b200.so pointcloudprocessor_b200/var_*.so; do
  [ -f "$lib" ] || continue
  PCP_LIB_PATH=$PWD/$lib timeout 240 python tools/bench_pass1.py --tag "$(basename $lib .so)" "$@" >> "$out" 2>> "$out.err" || echo "{\"lib\": \"$lib\", \"failed\": $?}" >> "$out"
done
python - "$out" <<'PY'
import json, sys
for l in open(sys.argv[1]):
    try: d = json.loads(l)
    except Exception: continue
    if "failed" in d: print(d); continue
    print(f"{d['tag']:>14}  pass1 {d['serial']['pass1_us']:6.2f} us  frac {d['serial']['pass1_frac_of_hbm']:.3f}  resolve {d['serial']['resolve_us']:5.2f}  serial frame {d['serial']['frame_us']:6.2f}  pipelined {d['pipelined']['frame_us']:6.2f} us  sha {d['sha']} {d['sha3']}  rec {d['counters_last_frame']['records']}")
PY
