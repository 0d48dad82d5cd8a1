run() {
  python bench.py --no-cpu --no-matching --no-other-configs --steps 10 --warmup 3 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('$1', 'value %.0f' % d['value'], 'ms %.2f' % d['ms_per_step'], 'e2e %.0f' % d['e2e']['value'])
"
}
for l in 1 2 3 4; do ORBX_LANES=$l run "lanes=$l"; done
for b in 128 512; do python bench.py --no-cpu --no-matching --no-other-configs --steps 10 --warmup 3 --batch $b 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('batch $b', 'value %.0f' % d['value'], 'ms %.2f' % d['ms_per_step'], 'e2e %.0f' % d['e2e']['value'])
"; done
