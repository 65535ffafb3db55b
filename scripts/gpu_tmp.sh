#!/bin/bash
mkdir -p gpurun_out
cp variants/fused.so salz_b200/libsalz_cuda.so
timeout 900 python -m pytest tests/test_gpu.py -m gpu -x -q -k "parse_segments or golden or medium or known or small" 2>&1 | tail -5
bash scripts/gpu_variants.sh
