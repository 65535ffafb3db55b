#!/bin/bash
mkdir -p gpurun_out
cp variants/a_base.so salz_b200/libsalz_cuda.so
timeout 600 python -m pytest tests/test_gpu.py -m gpu -x -q -k "pipelined" 2>&1 | tail -3
timeout 600 python bench.py --steps 3 --warmup 2 --no-cpu-baseline 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); print('value %.1f MB/s' % d['value'], 'e2e', d['e2e'])
"
bash scripts/gpu_variants.sh
bash scripts/gpu_variants.sh --workload dna1g --batch-blocks 512
