#!/bin/bash
# every metric of --set full for three launches of the scatter kernel (raw page, CSV), bench configuration
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 0 --no-cpu-baseline --no-extras --no-e2e"
$CMD > gpurun_out/plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/plain.log; exit 1; }
ncu --set full --clock-control none -k regex:k_scatter_tma -s 1 -c 3 -f -o /tmp/prof_sc $CMD > gpurun_out/ncu_sc.log 2>&1
echo "ncu exit $?"
ncu -i /tmp/prof_sc.ncu-rep --page raw --csv > gpurun_out/scatter_raw.csv 2>/dev/null
ncu -i /tmp/prof_sc.ncu-rep --page details > gpurun_out/scatter_details.txt 2>/dev/null
ls -la gpurun_out/scatter_raw.csv gpurun_out/scatter_details.txt
