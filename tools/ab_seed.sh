python -m pytest tests -m gpu -x -q 2>&1 | tail -3
S="import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); r=d['roofline']
print('value %.3e'%d['value'], 'e2e %.3e'%d['e2e']['value'], 'ms/step %.2f'%d['ms_per_step'], {k:round(v*1e3,1) for k,v in r['kernel_ms_per_frame'].items()}, 'frac %.3f pipe %.3f'%(r['frac'], r['pipeline']['frac']), d.get('parity'))"
echo DEFAULT; python bench.py --steps 10 2>/dev/null | python -c "$S"
for v in pointcloudprocessor_b200/var_*.so; do case $v in *stats*) continue;; esac
echo $v; PCP_LIB_PATH=$PWD/$v python bench.py --steps 10 --no-cpu 2>/dev/null | python -c "$S"; done

