#!/bin/bash
# per-phase cycles of k_scatter_tma (variants/prof.so built with SALZ_SCATTER_PROF=1), modes 1 and 3, 64 MiB of the bench text
mkdir -p gpurun_out
cp salz_b200/libsalz_cuda.so /tmp/orig.so; cp variants/prof.so salz_b200/libsalz_cuda.so
for m in ${MODES:-1 3}; do
  SALZ_SCATTER=$m timeout 300 python bench.py --steps 1 --warmup 0 --no-cpu-baseline --no-extras --no-e2e --mib 64 2>&1 | grep "scatter prof" | awk -v m=$m 'NR<=6 {print "SALZ_SCATTER=" m, $0}'
done | tee gpurun_out/scatter_prof.txt
cp /tmp/orig.so salz_b200/libsalz_cuda.so
