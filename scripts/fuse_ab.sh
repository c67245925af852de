# headline bench with the batches of a call fused into one launch (default) and with one launch per batch (HLSMAP_NO_FUSE=1)
B="python bench.py --no-cpu-baseline --no-sharded --e2e-steps 1"
for e in "X=1" "HLSMAP_NO_FUSE=1"; do
  for k in "20 5" "2000 50"; do
    set -- $k
    env $e $B --steps $1 --warmup $2 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
la=(d.get('other_rows') or {}).get('list_form_lookahead_mix',{})
print('$e', d['steps'], round(d['ms_per_step']*1e3,2), 'us', round(d['roofline']['frac'],4), 'iso', round(d['roofline']['frac_isolated'],4), 'launches', d['gpu_launches'], 'lookahead', la.get('ms_per_step'), la.get('roofline_frac'))"
  done
done
