CMD="python bench.py --steps 1 --warmup 0 --no-cpu-baseline --no-e2e --mib 64"
$CMD > gpurun_out/plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:'k_parse|k_ansv_tile|k_ansv_far|k_plcp_irreducible' -c 4 -f -o gpurun_out/prof_parse2 $CMD > gpurun_out/ncu5.log 2>&1
echo "ncu exit $?"
