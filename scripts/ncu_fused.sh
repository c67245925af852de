# one full ncu capture of the fused list-form launches of `bench.py --steps 5 --warmup 3` (after the same command ran clean)
B="python bench.py --no-other-rows --no-cpu-baseline --no-sharded --e2e-steps 1 --steps 5 --warmup 3"
$B > gpurun_out/ncu_fused_plain.json 2> gpurun_out/ncu_fused_plain.err || exit 1
ncu --set full --clock-control none --import-source on -k regex:resolve_search_kernel --launch-skip 4 --launch-count 2 \
    -o gpurun_out/ncu_fused -f $B > gpurun_out/ncu_fused.log 2>&1
