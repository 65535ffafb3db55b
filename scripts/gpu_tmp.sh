#!/bin/bash
mkdir -p gpurun_out
run() { python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-e2e "$@" 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); print('$*', 'value %.1f' % d['value'], {k: round(v, 2) for k, v in d['extra']['stage_ms_per_step'].items()})
    elif 'rror' in l: print(l.rstrip())
"; }
run --mib 256
run --mib 64
SALZ_PARSE_SEGMENTS=1 run --mib 64
run --mib 16
SALZ_PARSE_SEGMENTS=1 run --mib 16
run --workload english256 --level 9
