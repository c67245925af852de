# usage: ncu_variant.sh <variant or ""> <flags> <out name>  -- one full capture of the list-form kernel in steady state
v="$1"; export HLSMAP_VARIANT_FLAGS="$2"; out="$3"
if [ -n "$v" ]; then export HLSMAP_VARIANT="$v"; else unset HLSMAP_VARIANT HLSMAP_VARIANT_FLAGS; fi
B="python bench.py --no-other-rows --no-cpu-baseline --no-sharded --e2e-steps 1 --steps 20 --warmup 5"
$B > gpurun_out/${out}_plain.json 2> gpurun_out/${out}_plain.err || exit 1
ncu --set full --clock-control none --import-source on -k regex:resolve_search --launch-skip 12 --launch-count 2 \
    -o gpurun_out/${out} -f $B > gpurun_out/${out}_ncu.log 2>&1
