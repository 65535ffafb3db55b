#!/bin/bash
# compute-sanitizer on a small end-to-end case; $1 = memcheck | racecheck | synccheck | initcheck
# (one tool per box visit, B200_PROFILING.md).  The plain run must pass first.
TOOL=${1:-memcheck}
mkdir -p gpurun_out
timeout 300 python scripts/sanitize_case.py > gpurun_out/sanitize_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/sanitize_plain.log; exit 1; }
timeout 1500 compute-sanitizer --tool $TOOL --print-limit 30 python scripts/sanitize_case.py > gpurun_out/sanitize_$TOOL.log 2>&1
echo "sanitizer exit $?"
tail -15 gpurun_out/sanitize_$TOOL.log
