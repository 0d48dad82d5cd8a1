run() {
  python bench.py --no-cpu --no-matching --no-other-configs --steps 10 --warmup 3 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
r=d['roofline']['stage_ms_per_pass']
print('$1', 'value %.0f' % d['value'], 'ms %.2f' % d['ms_per_step'], {k: round(v,2) for k,v in r.items()})
"
}
run default
for c in 3 2 1; do ORBX_DESC_CTAS_PER_SM=$c run "desc_ctas=$c"; ORBX_DESC_CTAS_PER_SM=$c ORBX_LANES=3 run "desc_ctas=$c lanes=3"; done
