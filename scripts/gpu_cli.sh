#!/bin/bash
# wall time of programs/salz on the bench file (tmpfs), phases printed by SALZ_CLI_TIMING; $@ = batch sizes (MiB) to try
mkdir -p gpurun_out /dev/shm/salzcli
python - <<'PY'
import os, sys, numpy as np
sys.path.insert(0, '.')
p = '/tmp/salz_bench_cache/english_268439552_1_1048576.npy'
if os.path.exists(p):
    d = np.load(p)
else:
    from salz_b200 import synth
    d = synth.english_like((256 << 20) + 4096, seed=1)
d.tofile('/dev/shm/salzcli/a.bin')
PY
make -s -C lib
rm -f gpurun_out/cli.txt
for mib in "" "$@"; do
  for i in 1 2; do
    echo "=== SALZ_BATCH_MIB=${mib:-default}" | tee -a gpurun_out/cli.txt
    ( time env SALZ_CLI_TIMING=1 ${mib:+SALZ_BATCH_MIB=$mib} programs/salz -k -f -5 /dev/shm/salzcli/a.bin ) 2>&1 | tee -a gpurun_out/cli.txt
  done
done
rm -rf /dev/shm/salzcli
