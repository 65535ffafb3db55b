#!/bin/bash
# stage times against batch size (fixed cost per batch)
mkdir -p gpurun_out
for m in 16 32 64 128; do
echo "=== mib $m"
timeout 600 python bench.py --steps 3 --warmup 2 --no-cpu-baseline --mib $m 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); print('value %.1f MB/s e2e %.1f ms %.2f launches %d' % (d['value'], d['e2e']['value'], d['e2e']['ms_per_step'], d['gpu_launches']), {k: round(v, 2) for k, v in d['extra']['stage_ms_per_step'].items()})
"
done 2>&1 | tee gpurun_out/batchsize.txt
