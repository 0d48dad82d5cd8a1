#!/bin/bash
# Experiment driver (GPU box): lane count sweep, and optional variant libraries under gpurun_variants/.
# usage: tools/exp_variants.sh out_prefix [variant.so ...]
out=$1; shift
run() { # label
  python bench.py --no-cpu --no-matching --steps 10 --warmup 3 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
r=d['roofline']['stage_ms_per_pass']
print('$1', 'value %.0f' % d['value'], 'e2e %.0f' % d['e2e']['value'], 'ms %.2f' % d['ms_per_step'], {k: round(v,2) for k,v in r.items()})
" >> $out
}
for l in 1 2 3 4; do ORBX_LANES=$l run "default lanes=$l"; done
cp orb_slam_b200/liborbx.so /tmp/liborbx_default.so
for v in "$@"; do
  cp $v orb_slam_b200/liborbx.so
  for l in 2 3; do ORBX_LANES=$l run "$(basename $v) lanes=$l"; done
done
cp /tmp/liborbx_default.so orb_slam_b200/liborbx.so
cat $out
