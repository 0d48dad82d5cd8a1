run() {
  python bench.py --no-cpu --no-matching --no-other-configs --steps 10 --warmup 3 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
r=d['roofline']['stage_ms_per_pass']
print('$1', 'value %.0f' % d['value'], 'ms %.2f' % d['ms_per_step'], {k: round(v,2) for k,v in r.items()})
"
}
run default
ORBX_SELECT=smem run select_smem
ORBX_SELECT=global run select_global
