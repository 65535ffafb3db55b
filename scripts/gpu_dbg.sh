#!/bin/bash
# SA of blocks of growing size against the oracle (debugging aid)
mkdir -p gpurun_out
cat > /tmp/dbg.py <<'PY'
import sys, numpy as np
sys.path.insert(0, '.')
import salz_b200
from salz_b200 import synth
from oracle.binding import Port
o = Port()
enc = salz_b200.SalzCuda(device=0, max_block_len=1 << 20, max_batch_blocks=4)
for name, gen in [("eng", lambda n: synth.english_like(n, seed=5)), ("dna", lambda n: synth.dna_like(n, seed=7, block=n)),
                  ("a", lambda n: np.full(n, 0x61, np.uint8)), ("rnd", lambda n: synth.uniform_random(n, seed=3))]:
    for n in [3000, 5000, 9000, 20000, 100000, 300000, (1 << 20) - 8]:
        t = gen(n)
        sa = enc.suffix_arrays([t])[0]
        ref = o.suffix_array(t)
        ok = np.array_equal(sa, ref)
        msg = ""
        if not ok:
            bad = np.flatnonzero(sa != ref)
            isperm = np.array_equal(np.sort(sa), np.arange(n))
            msg = " nbad=%d first=%d perm=%s got=%s want=%s" % (len(bad), bad[0], isperm, sa[bad[0]:bad[0]+4], ref[bad[0]:bad[0]+4])
        print(name, n, "OK" if ok else "BAD" + msg, "rounds", enc.stats()["doubling_rounds"], flush=True)
PY
timeout 900 python /tmp/dbg.py > gpurun_out/dbg.log 2>&1
echo "exit $?"
tail -40 gpurun_out/dbg.log
