#!/bin/bash
# bench lines for the other BASELINE configs (parity of these shapes is in tests/test_gpu.py)
mkdir -p gpurun_out
for args in "--workload dna1g --batch-blocks 512" "--workload single512 --batch-blocks 512" "--workload random512 --batch-blocks 512" "--workload english256 --level 9" "--workload single512 --level 9 --mib 128"; do
  echo "=== $args"
  timeout 900 python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-extras $args 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); print('value %.1f MB/s  e2e %.1f' % (d['value'], d['e2e']['value']), 'roofline %.3f' % d['roofline']['frac'], 'rounds', d['extra']['doubling_rounds_per_step'], 'ratio %.2f' % d['extra']['ratio'], {k: round(v, 2) for k, v in d['extra']['stage_ms_per_step'].items()}); print(l.strip())
    else: print(l.rstrip())
"
done 2>&1 | tee gpurun_out/workloads.txt
