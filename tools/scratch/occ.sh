run() {
  python bench.py --no-cpu --no-matching --no-other-configs --steps 10 --warmup 3 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
r=d['roofline']['stage_ms_per_pass']
print('$1', 'value %.0f' % d['value'], 'ms %.2f' % d['ms_per_step'], {k: round(v,2) for k,v in r.items()})
"
}
for p in 0 2000 9000 20000; do ORBX_FAST_SMEM_PAD=$p run "pad=$p"; done
python tools/single_frame_latency.py 2>&1 | tail -3
