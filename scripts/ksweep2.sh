# usage: ksweep2.sh "ENV=VAL" ...  -- 20- and 2000-step headline (fused batches call) per setting, plus the look-ahead row
B="python bench.py --no-cpu-baseline --no-sharded --e2e-steps 1"
for e in "$@"; do
  env $e $B --steps 20 --warmup 5 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
la=(d.get('other_rows') or {}).get('list_form_lookahead_mix',{})
print('$e', 20, round(d['ms_per_step']*1e3,2), round(d['roofline']['frac'],4), 'unfused', round(d['roofline']['one_launch_per_batch']['kernel_ms_mean']*1e3,2), 'iso', round(d['roofline']['frac_isolated'],4), 'lookahead', round(la.get('ms_per_step',0)*1e3,2))"
  env $e $B --no-other-rows --steps 2000 --warmup 50 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('$e', 2000, round(d['ms_per_step']*1e3,2), round(d['roofline']['frac'],4), 'unfused', round(d['roofline']['one_launch_per_batch']['kernel_ms_mean']*1e3,2))"
done
