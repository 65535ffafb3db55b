#!/bin/bash
mkdir -p gpurun_out
run() { python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-e2e "$@" 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); print('$*', 'value %.1f' % d['value'], {k: round(v, 2) for k, v in d['extra']['stage_ms_per_step'].items()})
    elif 'rror' in l: print(l.rstrip())
"; }
for v in variants/*.so; do
  cp $v salz_b200/libsalz_cuda.so
  echo "=== $v"
  run --mib 256
  SALZ_PARSE_SEGMENTS=4 run --mib 256
  SALZ_PARSE_SEGMENTS=8 run --mib 256
done
