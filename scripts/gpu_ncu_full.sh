#!/bin/bash
# ncu --set full captures at the bench configuration (256 MiB english, level 5).  The reports are
# exported to CSV on the box (raw metrics for every launch, source view for the first launches)
# and removed, so that gpurun_out/ stays under its size limit.
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 0 --no-cpu-baseline --no-e2e"
$CMD > gpurun_out/plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/plain.log; exit 1; }
ncu --set full --clock-control none -k regex:'k_scatter|k_local|k_parse' -c 30 -f -o /tmp/prof_main $CMD > gpurun_out/ncu1.log 2>&1
echo "ncu1 exit $?"
ncu -i /tmp/prof_main.ncu-rep --page raw --csv > gpurun_out/main_raw.csv 2>/dev/null
ncu --set full --clock-control none --import-source on -k regex:'k_rerank_apply|k_ansv_tile|k_plcp_irreducible|k_init|k_plcp_apply|k_emit_tokens|k_local|k_parse' -c 8 -f -o /tmp/prof_rest $CMD > gpurun_out/ncu2.log 2>&1
echo "ncu2 exit $?"
ncu -i /tmp/prof_rest.ncu-rep --page raw --csv > gpurun_out/rest_raw.csv 2>/dev/null
ncu -i /tmp/prof_rest.ncu-rep --page source --csv > gpurun_out/rest_src.csv 2>/dev/null
ls -la gpurun_out/
