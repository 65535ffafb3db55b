#!/bin/bash
# every metric of --set full for the launches of the other kernels of a step (raw page, CSV), bench configuration
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 0 --no-cpu-baseline --no-extras --no-e2e"
$CMD > gpurun_out/plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/plain.log; exit 1; }
ncu --set full --clock-control none -k regex:'k_local|k_rerank_apply|k_ansv_tile|k_plcp_irreducible|k_plcp_apply|k_parse|k_choose|k_init|k_emit_tokens' -c 40 -f -o /tmp/prof_ot $CMD > gpurun_out/ncu_ot.log 2>&1
echo "ncu exit $?"
ncu -i /tmp/prof_ot.ncu-rep --page raw --csv > gpurun_out/others_raw.csv 2>/dev/null
ls -la gpurun_out/others_raw.csv
