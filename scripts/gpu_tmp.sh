#!/bin/bash
mkdir -p gpurun_out
cp variants/warm16.so salz_b200/libsalz_cuda.so
timeout 900 python -m pytest tests/test_gpu.py -m gpu -x -q -k "parse_segments or golden or medium" 2>&1 | tail -5
bash scripts/gpu_variants.sh
for w in 0 8 16; do cp variants/warm$w.so salz_b200/libsalz_cuda.so; echo "=== warm$w dna1g / random512 / single512"; for args in "--workload dna1g --batch-blocks 512" "--workload random512 --batch-blocks 512" "--workload single512 --batch-blocks 512"; do
timeout 600 python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-e2e $args 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); print('value %.1f MB/s' % d['value'], {k: round(v, 2) for k, v in d['extra']['stage_ms_per_step'].items()})
"; done; done
