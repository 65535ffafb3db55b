#!/bin/bash
# ncu captures at the bench configuration (256 MiB english, level 5).  The --set full report is
# exported to CSV on the box (raw metrics for every launch of the main kernels; source view of a
# few) and removed, so that gpurun_out/ stays under its size limit.
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 0 --no-cpu-baseline --no-extras --no-e2e"
$CMD > gpurun_out/plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/plain.log; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none -c 2000 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu0.log 2>&1
echo "ncu0 exit $?"
ncu --set full --clock-control none -k regex:'k_scatter|k_local|k_parse|k_choose|k_rerank_apply|k_ansv_tile|k_plcp_irreducible|k_plcp_apply|k_init|k_emit_tokens' -c 48 -f -o /tmp/prof_main $CMD > gpurun_out/ncu1.log 2>&1
echo "ncu1 exit $?"
ncu -i /tmp/prof_main.ncu-rep --page raw --csv > gpurun_out/main_raw.csv 2>/dev/null
for K in k_scatter_tma k_local k_ansv_tile; do
  ncu --set full --clock-control none --import-source on -k regex:$K -c 1 -f -o /tmp/prof_src_$K $CMD > gpurun_out/ncu2_$K.log 2>&1
  echo "ncu2 $K exit $?"
  ncu -i /tmp/prof_src_$K.ncu-rep --page source --csv > gpurun_out/src_$K.csv 2>/dev/null
done
ls -la gpurun_out/
