#!/bin/bash
mkdir -p gpurun_out
run() { python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-e2e "$@" 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); print('$SALZ_PARSE_SEGMENTS $*', 'value %.1f' % d['value'], 'launches', d['gpu_launches'], {k: round(v, 2) for k, v in d['extra']['stage_ms_per_step'].items()})
    elif 'rror' in l: print(l.rstrip())
"; }
for k in 1 2 4 8 16; do export SALZ_PARSE_SEGMENTS=$k; run --mib 256; done
for k in 1 4 16 64; do export SALZ_PARSE_SEGMENTS=$k; run --mib 64; done
