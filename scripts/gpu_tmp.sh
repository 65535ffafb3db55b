#!/bin/bash
mkdir -p gpurun_out
cp variants/b_fused.so salz_b200/libsalz_cuda.so
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
bash scripts/gpu_variants.sh
cp variants/b_fused.so salz_b200/libsalz_cuda.so
timeout 600 python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-e2e --workload english256 --level 9 | grep -o "stage_ms_per_step[^}]*}"
