CMD="python bench.py --steps 1 --warmup 0 --no-cpu-baseline --no-e2e --mib 64"
$CMD > gpurun_out/plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:'k_scatter|k_rerank_apply' -s 5 -c 4 -f -o gpurun_out/prof_scatter2 $CMD > gpurun_out/ncu4.log 2>&1
echo "ncu exit $?"
