#!/bin/bash
# A/B of run-time switches on one build: each line of $1 (a file of "ENV=.. ENV=.." settings, "-" = defaults)
# gets a quick parity run and a bench line (stage times).  Extra bench flags: $2..
mkdir -p gpurun_out
export SALZ_BENCH_CACHE=/tmp/salz_bench_cache
CFG=$1; shift
while IFS= read -r line; do
  [ -z "$line" ] && continue
  [ "$line" = "-" ] && line=""
  echo "=== ${line:-defaults}"
  env $line timeout 120 python -m pytest tests/test_gpu.py -m gpu -x -q -k "medium or small or batched or alphabet" 2>&1 | tail -1
  env $line timeout 180 python bench.py --steps 3 --warmup 2 --no-cpu-baseline --no-e2e --no-extras $* 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); print('value %.1f MB/s' % d['value'], 'roofline', d['roofline']['kernel'], '%.3f' % d['roofline']['frac'], [(r['kernel'], round(r['frac'], 3)) for r in d['roofline_others']], {k: round(v, 2) for k, v in d['extra']['stage_ms_per_step'].items()})
    else: print(l.rstrip())
"
done < "$CFG" 2>&1 | tee gpurun_out/ab.txt
