#!/bin/bash
# ncu launch list (per-launch device time) of one bench step; args: extra bench flags
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-extras --no-e2e $*"
$CMD > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 2000 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu.log 2>&1
echo "exit $?"; tail -2 gpurun_out/plain.log | cut -c1-300; wc -l gpurun_out/launches.csv
