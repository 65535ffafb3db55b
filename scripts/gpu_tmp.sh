#!/bin/bash
mkdir -p gpurun_out
for mib in 64 128 256; do
python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-e2e --mib $mib 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); print('mib $mib value %.1f' % d['value'], {k: round(v, 2) for k, v in d['extra']['stage_ms_per_step'].items()})
"
done 2>&1 | tee gpurun_out/tmp.txt
