# usage: ab_env.sh "ENV=VAL ..." ["ENV=VAL ..." ...]  -- headline bench (20 and 2000 steps) and one configs[4] chunk per setting
B="python bench.py --no-cpu-baseline --no-sharded --e2e-steps 1"
i=0
for setting in "$@"; do
  i=$((i+1))
  env $setting $B --steps 20 --warmup 5 > gpurun_out/env_${i}_20.json 2> gpurun_out/env_${i}.err
  env $setting $B --no-other-rows --steps 2000 --warmup 50 > gpurun_out/env_${i}_2000.json 2>> gpurun_out/env_${i}.err
  env $setting python scripts/run_config5_piece.py 0 7 > gpurun_out/env_${i}_c5.txt 2>&1
  python - "$setting" $i <<'PY'
import json, sys
s, i = sys.argv[1], sys.argv[2]
for k in ("20", "2000"):
    t = open("gpurun_out/env_%s_%s.json" % (i, k)).read().strip()
    if not t:
        print(s, k, "EMPTY"); continue
    d = json.loads(t.splitlines()[-1])
    o = d.get("other_rows") or {}
    la = o.get("list_form_lookahead_mix", {}).get("ms_per_step")
    print(s, "|", k, "steps:", round(d["ms_per_step"] * 1e3, 2), "us", round(d["roofline"]["frac"], 4), "isolated", round(d["roofline"]["frac_isolated"], 4), "lookahead us", la and round(la * 1e3, 2))
print(s, "|", open("gpurun_out/env_%s_c5.txt" % i).read().strip().splitlines()[-1][-120:])
PY
done
