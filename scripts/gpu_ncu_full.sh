#!/bin/bash
# ncu --set full captures of the top kernels on a 64 MiB slice of the bench workload.
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 0 --no-cpu-baseline --no-e2e --mib 64"
$CMD > gpurun_out/plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/plain.log; exit 1; }
ncu --set full --clock-control none --import-source on -k regex:'k_local|k_parse|k_rerank_apply|k_scatter' -c 8 -f -o gpurun_out/prof_top $CMD > gpurun_out/ncu1.log 2>&1
echo "ncu1 exit $?"
ncu --set full --clock-control none --import-source on -k regex:'k_ansv_tile|k_plcp_irreducible|k_init|k_plcp_apply' -c 4 -f -o gpurun_out/prof_rest $CMD > gpurun_out/ncu2.log 2>&1
echo "ncu2 exit $?"
ls -la gpurun_out/
