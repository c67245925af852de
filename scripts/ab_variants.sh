# usage: ab_variants.sh name:flags [name:flags ...]  -- headline bench (20 and 2000 steps) + one configs[4] chunk for the default build and each variant library
B="python bench.py --no-cpu-baseline --no-sharded --e2e-steps 1 --no-other-rows"
for spec in "base:" "$@"; do
  n="${spec%%:*}"; f="${spec#*:}"
  if [ "$n" = base ]; then unset HLSMAP_VARIANT HLSMAP_VARIANT_FLAGS; else export HLSMAP_VARIANT="$n" HLSMAP_VARIANT_FLAGS="$f"; fi
  $B --steps 20 --warmup 5 > gpurun_out/v_${n}_20.json 2> gpurun_out/v_${n}.err
  $B --steps 2000 --warmup 50 > gpurun_out/v_${n}_2000.json 2>> gpurun_out/v_${n}.err
  python scripts/run_config5_piece.py 0 7 > gpurun_out/v_${n}_c5.txt 2>&1
  python - "$n" <<'PY'
import json, sys
n = sys.argv[1]
out = [n]
for k in ("20", "2000"):
    t = open("gpurun_out/v_%s_%s.json" % (n, k)).read().strip()
    if not t:
        out.append("EMPTY"); continue
    d = json.loads(t.splitlines()[-1])
    out += [k + ":", round(d["ms_per_step"] * 1e3, 2), "us", round(d["roofline"]["frac"], 4), "iso", round(d["roofline"]["frac_isolated"], 4)]
out.append(open("gpurun_out/v_%s_c5.txt" % n).read().strip().splitlines()[-1][-45:])
print(*out)
PY
done
