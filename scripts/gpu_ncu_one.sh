#!/bin/bash
# ncu --set full capture of one kernel (regex $1) on a 64 MiB slice of the bench workload; report name $2
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 0 --no-cpu-baseline --no-extras --no-e2e --mib 64"
$CMD > gpurun_out/plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/plain.log; exit 1; }
ncu --set full --clock-control none --import-source on -k regex:"$1" -c ${3:-2} -f -o gpurun_out/$2 $CMD > gpurun_out/ncu_$2.log 2>&1
echo "ncu exit $?"
