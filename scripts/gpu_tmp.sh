./scripts/ubench/rank > gpurun_out/ubench_rank.txt 2>&1; cat gpurun_out/ubench_rank.txt
CMD="python bench.py --steps 1 --warmup 0 --no-cpu-baseline --no-e2e --mib 64"
$CMD > gpurun_out/plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:'k_parse|k_ansv|k_plcp_irreducible|k_plcp_apply' -c 4 -f -o gpurun_out/prof_parse $CMD > gpurun_out/ncu3.log 2>&1
echo "ncu exit $?"
