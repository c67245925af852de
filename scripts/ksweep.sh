# usage: ksweep.sh "ENV=VAL" ...  -- region time of K fused / unfused batches for several K
B="python bench.py --no-cpu-baseline --no-sharded --e2e-steps 1 --no-other-rows"
for e in "$@"; do
for k in 8 20 32 64; do
  env $e $B --steps $k --warmup 5 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('$e', d['steps'], 'region_us', round(d['ms_per_step']*1e3*d['steps'],1), 'per_batch', round(d['ms_per_step']*1e3,2), 'launches', d['gpu_launches'])"
done; done
