#!/bin/bash
# A/B on ONE box: `tools/exp_ab.sh out label1 "ENV=.. " lib1 label2 "ENV" lib2 ...` (lib "-" = the in-tree build); 3 rounds.
out=$1; shift
cp orb_slam_b200/liborbx.so /tmp/liborbx_default.so
args=("$@")
for round in 1 2 3; do
  i=0
  while [ $i -lt ${#args[@]} ]; do
    label=${args[$i]}; envs=${args[$((i+1))]}; lib=${args[$((i+2))]}; i=$((i+3))
    if [ "$lib" = "-" ]; then cp /tmp/liborbx_default.so orb_slam_b200/liborbx.so; else cp $lib orb_slam_b200/liborbx.so; fi
    env $envs python bench.py --no-cpu --no-matching --steps 10 --warmup 3 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('$label', 'value %.0f' % d['value'], 'e2e %.0f' % d['e2e']['value'], 'ms %.2f' % d['ms_per_step'])" >> $out
  done
done
cp /tmp/liborbx_default.so orb_slam_b200/liborbx.so
cat $out
