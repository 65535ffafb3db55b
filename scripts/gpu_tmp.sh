#!/bin/bash
mkdir -p gpurun_out
run() { python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-e2e "$@" 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); print('$SALZ_PARSE_SEGMENTS $*', 'value %.1f' % d['value'], 'launches', d['gpu_launches'], {k: round(v, 2) for k, v in d['extra']['stage_ms_per_step'].items() if k in ('ms_parse','ms_total')})
    elif 'rror' in l: print(l.rstrip())
"; }
for v in variants/*.so; do
  cp $v salz_b200/libsalz_cuda.so
  echo "=== $v"
  timeout 600 python -m pytest tests/test_gpu.py -m gpu -x -q -k "parse_segments" 2>&1 | tail -1
  for k in 0 8 32; do export SALZ_PARSE_SEGMENTS=$k; run --mib 256; done
  unset SALZ_PARSE_SEGMENTS
  run --workload english256 --level 9
done
