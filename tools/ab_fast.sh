#!/bin/bash
# A/B driver (GPU box): extractor parity tests, then stage times of the default library under env settings and of
# variant libraries.  usage: tools/ab_fast.sh [variant.so ...]
python -m pytest tests/test_gpu_extract.py tests/test_gpu_fullsize.py -x -q -m gpu 2>&1 | tail -n 4
run() {
  python bench.py --no-cpu --no-matching --no-other-configs --steps 10 --warmup 3 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
r=d['roofline']['stage_ms_per_pass']
print('$1', 'value %.0f' % d['value'], 'e2e %.0f' % d['e2e']['value'], 'ms %.2f' % d['ms_per_step'], {k: round(v,2) for k,v in r.items()})
"
}
run default
ORBX_FAST=legacy ORBX_RESIZE=legacy run legacy
cp orb_slam_b200/liborbx.so /tmp/liborbx_default.so
for v in "$@"; do
  cp $v orb_slam_b200/liborbx.so
  run "$(basename $v)"
done
cp /tmp/liborbx_default.so orb_slam_b200/liborbx.so
