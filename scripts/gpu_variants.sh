#!/bin/bash
# A/B of prebuilt library variants in variants/*.so : parity smoke + bench stage times
mkdir -p gpurun_out
for v in variants/*.so; do
  cp $v salz_b200/libsalz_cuda.so
  echo "=== $v"
  timeout 600 python -m pytest tests/test_gpu.py -m gpu -x -q -k "medium or small or batched" 2>&1 | tail -1
  timeout 600 python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-extras --no-e2e $* 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); print('value %.1f MB/s' % d['value'], 'roofline %.3f' % d['roofline']['frac'], {k: round(v, 2) for k, v in d['extra']['stage_ms_per_step'].items()})
    else: print(l.rstrip())
"
done 2>&1 | tee gpurun_out/variants.txt
