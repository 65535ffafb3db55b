#!/bin/bash
# bench lines for the small-block CLI levels: 64 KiB .. 512 KiB blocks, 256 MiB English-like text
mkdir -p gpurun_out
for args in "--level 1 --batch-blocks 2048" "--level 3 --batch-blocks 1024" "--level 4" "--level 7"; do
  echo "=== $args"
  timeout 600 python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-extras $args 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); print('value %.1f MB/s  e2e %.1f' % (d['value'], d['e2e']['value']), 'rounds', d['extra']['doubling_rounds_per_step'], 'ratio %.2f' % d['extra']['ratio'], 'blocks', d['config']['blocks_per_gpu'], {k: round(v, 2) for k, v in d['extra']['stage_ms_per_step'].items()})
    else: print(l.rstrip())
"
done 2>&1 | tee gpurun_out/levels.txt
