#!/bin/bash
# Runs the short bench for the default library and every pointcloudprocessor_b200/var_*.so kernel variant (PCP_LIB_PATH).
cd "$(dirname "$0")/.."
for lib in pointcloudprocessor_b200/libpcp_b200.so pointcloudprocessor_b200/var_*.so; do
  [ -f "$lib" ] || continue
  PCP_LIB_PATH=$PWD/$lib python bench.py --steps 10 --no-cpu "$@" 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
r=d['roofline']
print('$lib', 'value %.3e'%d['value'], 'e2e %.3e'%d['e2e']['value'], 'ms/frame', {k:round(v*1e3,1) for k,v in r['kernel_ms_per_frame'].items()}, 'pipe_frac %.3f'%r['pipeline']['frac'])
"
done
