timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
for args in "--workload single512 --batch-blocks 512" "--workload english256"; do
timeout 900 python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-e2e $args 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); print('value %.1f MB/s' % d['value'], {k: round(v, 2) for k, v in d['extra']['stage_ms_per_step'].items()})
    else: print(l.rstrip())
"
done
