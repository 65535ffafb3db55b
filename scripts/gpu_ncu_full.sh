#!/bin/bash
# ncu --set full captures of the top kernels on a 64 MiB slice of the bench workload.
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 0 --no-cpu-baseline --no-e2e --mib 64"
$CMD > gpurun_out/plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/plain.log; exit 1; }
ncu --set full --clock-control none --import-source on -k regex:'k_radix_scatter' -s 8 -c 2 -f -o gpurun_out/prof_scatter $CMD > gpurun_out/ncu1.log 2>&1
echo "ncu1 exit $?"
ncu --set full --clock-control none --import-source on -k regex:'k_parse|k_ansv|k_build_keys|k_rerank_apply|k_init_keys|k_plcp_irreducible' -c 6 -f -o gpurun_out/prof_rest $CMD > gpurun_out/ncu2.log 2>&1
echo "ncu2 exit $?"
ls -la gpurun_out/
