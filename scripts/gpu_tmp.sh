#!/bin/bash
for pct in 50 35 25 15; do echo "=== split $pct"; SALZ_SPLIT_PCT=$pct timeout 600 python bench.py --steps 4 --warmup 2 --no-cpu-baseline 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); print('value %.1f MB/s e2e %.1f ms %.2f' % (d['value'], d['e2e']['value'], d['e2e']['ms_per_step']))
"; done
