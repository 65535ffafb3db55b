"""First-contact GPU diagnostic: every stage of every small case against the
oracle, with the first mismatch printed.  Run on the GPU box:
    timeout 600 python scripts/gpu_check.py
"""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle.binding import Port, Ref  # noqa: E402  (diagnostic script: checker only)
import salz_b200  # noqa: E402
from salz_b200 import synth  # noqa: E402


def first_diff(a, b):
    a = np.asarray(a); b = np.asarray(b)
    if a.shape != b.shape:
        return "shape %s vs %s" % (a.shape, b.shape)
    d = np.nonzero(a != b)
    if len(d[0]) == 0:
        return None
    i = tuple(x[0] for x in d)
    return "first diff at %s: got %s want %s (total %d)" % (i, a[i], b[i], len(d[0]))


def check_case(enc, orc, name, data):
    st = enc.stages(data)
    want = orc.stages(data)
    ok = True
    n = len(data) - 8
    for k in ("sa", "plcp", "lcp"):
        d = first_diff(st[k], want[k])
        if d:
            print("  [%s] %s: %s" % (name, k, d)); ok = False
            break
    else:
        d = first_diff(st["cand"][1:], want["cand"][1:])
        if d:
            print("  [%s] cand: %s" % (name, d)); ok = False
        else:
            d = first_diff(st["parse"][1:n], want["parse"][1:n])
            if d:
                print("  [%s] parse: %s" % (name, d)); ok = False
            elif st["out"] != want["out"]:
                a = np.frombuffer(st["out"], np.uint8); b = np.frombuffer(want["out"], np.uint8)
                m = min(len(a), len(b))
                print("  [%s] out: len %d vs %d, %s" % (name, len(a), len(b), first_diff(a[:m], b[:m]))); ok = False
    return ok


def main():
    port = Port()
    orc = Ref() if Ref.available() else port
    print("oracle:", type(orc).__name__)
    enc = salz_b200.SalzCuda(0, 1 << 20, 8)
    cases = synth.small_cases(3)
    cases += [("zeros", np.zeros(100, np.uint8)),
              ("zero_tail", np.concatenate([np.arange(50, dtype=np.uint8) + 1, np.zeros(20, np.uint8)])),
              ("abcdefghi", np.frombuffer(b"abcdefghi", np.uint8)),
              ("eng64k", synth.english_like(65536 + 100, seed=9)),
              ("dna200k", synth.dna_like(200000, seed=3, block=200000)),
              ("a100k", np.full(100000, 97, np.uint8)),
              ("rand50k", synth.uniform_random(50000, seed=1)),
              ("eng1M", synth.english_like(1 << 20, seed=5))]
    for mode in (0, 1):
        enc.set_sort_mode(mode)
        bad = 0
        t0 = time.time()
        for name, data in cases:
            try:
                if not check_case(enc, orc, name, data):
                    bad += 1
            except Exception as e:  # noqa: BLE001
                print("  [%s] EXC %r" % (name, e)); bad += 1
                break
        print("sort mode %d: %d/%d cases bad, %.1fs" % (mode, bad, len(cases), time.time() - t0), flush=True)
    # batched stream vs per-block oracle
    data = np.concatenate([synth.english_like(3 << 20, seed=11), synth.dna_like(2 << 20, seed=12),
                           synth.uniform_random(1 << 20, seed=13), np.full((1 << 20) + 777, 98, np.uint8)])
    bl = 1 << 20
    enc.reset_stats()
    t0 = time.time()
    stream, sizes = enc.encode_stream(data, bl)
    t1 = time.time()
    want = b""
    for off in range(0, len(data), bl):
        blk = orc.encode(data[off: off + bl])
        want += len(blk).to_bytes(4, "little") + blk
    print("stream: %d bytes in %.3fs, match=%s (want %d got %d)" % (len(data), t1 - t0, stream == want, len(want), len(stream)))
    print("stats:", enc.stats())


if __name__ == "__main__":
    main()
