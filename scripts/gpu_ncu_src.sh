#!/bin/bash
# ncu --set full + source of kernels matching regex $1 (first $2 launches) on a 64 MiB slice of the
# bench workload; the source page is exported to gpurun_out/src_$3.csv, raw metrics to raw_$3.csv
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 0 --no-cpu-baseline --no-extras --no-e2e --mib 64"
$CMD > gpurun_out/plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/plain.log; exit 1; }
ncu --set full --clock-control none --import-source on -k regex:"$1" -c ${2:-2} -f -o /tmp/rep_$3 $CMD > gpurun_out/ncu_$3.log 2>&1
echo "ncu exit $?"
ncu -i /tmp/rep_$3.ncu-rep --page source --csv > gpurun_out/src_$3.csv 2>/dev/null
ncu -i /tmp/rep_$3.ncu-rep --page raw --csv > gpurun_out/raw_$3.csv 2>/dev/null
ls -la gpurun_out | tail -5
