#!/bin/bash
# e2e of the bench step for different cuts of a one-batch host stream (SALZ_SPLIT_PCT / SALZ_SPLIT_LAST_PCT)
mkdir -p gpurun_out
export SALZ_BENCH_CACHE=/tmp/c
for cfg in "SALZ_SPLIT_PCT=25" "SALZ_SPLIT_PCT=20 SALZ_SPLIT_LAST_PCT=15" "SALZ_SPLIT_PCT=15 SALZ_SPLIT_LAST_PCT=10" "SALZ_SPLIT_PCT=25 SALZ_SPLIT_LAST_PCT=25" "SALZ_SPLIT_PCT=12" "SALZ_SPLIT_PCT=10 SALZ_SPLIT_LAST_PCT=6"; do
  echo "=== $cfg"
  env $cfg timeout 200 python bench.py --steps 5 --warmup 2 --no-cpu-baseline --no-extras 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); print('device %.2f ms  e2e %.2f ms  (%.0f MB/s) verified %s' % (d['ms_per_step'], d['e2e']['ms_per_step'], d['e2e']['value'], d['verified']))
"
done 2>&1 | tee gpurun_out/split.txt
