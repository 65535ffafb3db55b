#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
run() { python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-e2e "$@" 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); print('$*', 'value %.1f' % d['value'], 'launches', d['gpu_launches'], {k: round(v, 2) for k, v in d['extra']['stage_ms_per_step'].items()})
    elif 'rror' in l: print(l.rstrip())
"; }
run --mib 256
run --workload dna1g --batch-blocks 512
run --workload single512 --batch-blocks 512
run --workload random512 --batch-blocks 512
run --workload english256 --level 9
run --workload single512 --level 9 --mib 128
