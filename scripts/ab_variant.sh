# usage: ab_variant.sh <variant> <flags> [tests]  -- bench the default build and a variant build back to back
V="$1"; F="$2"
B="python bench.py --no-other-rows --no-cpu-baseline --no-sharded --e2e-steps 1"
run() {   # name
  $B --steps 20 --warmup 5 > gpurun_out/ab_$1_20.json 2> gpurun_out/ab_$1.err
  $B --steps 2000 --warmup 50 > gpurun_out/ab_$1_2000.json 2>> gpurun_out/ab_$1.err
  python scripts/run_config5_piece.py 0 7 > gpurun_out/ab_$1_c5.txt 2>&1
}
if [ -n "$3" ]; then
  HLSMAP_VARIANT=$V HLSMAP_VARIANT_FLAGS="$F" timeout 900 python -m pytest $3 -m gpu -x -q > gpurun_out/ab_${V}_tests.txt 2>&1
  tail -n 3 gpurun_out/ab_${V}_tests.txt
fi
run base
export HLSMAP_VARIANT=$V HLSMAP_VARIANT_FLAGS="$F"
run $V
python - <<PY
import json
for n in ("base", "$V"):
    for k in ("20", "2000"):
        t = open("gpurun_out/ab_%s_%s.json" % (n, k)).read().strip()
        if not t:
            print(n, k, "EMPTY"); continue
        d = json.loads(t.splitlines()[-1])
        print(n, k, round(d["ms_per_step"] * 1e3, 2), "us", round(d["roofline"]["frac"], 4), d["roofline"].get("frac_isolated"))
    print(n, open("gpurun_out/ab_%s_c5.txt" % n).read().strip().splitlines()[-1][:300])
PY
