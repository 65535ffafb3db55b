"""Generates tests/golden/golden.json from the compiled reference
(oracle/_ref/libsalz_ref.so, built in place from /root/reference).

Run in the authoring container only:  python tests/golden/make_golden.py
The vectors are small (inputs <= 3000 bytes) and are stored as hex, together
with the reference's intermediate tables where they are short enough.
"""
import hashlib
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle.binding import Ref  # noqa: E402
from salz_b200 import synth  # noqa: E402


def main():
    ref = Ref()
    vectors = []
    named = [
        ("kat_abcdefghi", b"abcdefghi"),
        ("kat_a64", b"a" * 64),
        ("kat_abc_xyz", b"abcabcabcabcabcabcabcabcXYZ01234"),
        ("zeros100", bytes(100)),
        ("zero_tail", bytes(range(1, 51)) + bytes(20)),
    ]
    for name, arr in synth.small_cases(seed=7)[:20]:
        named.append((name, arr.tobytes()))
    for name, data in named:
        st = ref.stages(data)
        v = {"name": name, "input": data.hex(), "output": st["out"].hex()}
        if len(data) <= 400:
            v["sa"] = st["sa"].tolist()
            v["lcp"] = st["lcp"].tolist()
            v["cand"] = st["cand"][1:].reshape(-1).tolist()
            v["parse"] = st["parse"][1:].reshape(-1).tolist()
        vectors.append(v)
    # larger known answers as digests only
    big = []
    a = np.full(1 << 24, 0x61, np.uint8)
    out = ref.encode(a)
    big.append({"name": "a_x_16MiB", "gen": "single_byte(1<<24)", "len": len(out),
                "sha256": hashlib.sha256(out).hexdigest(), "head": out[:16].hex()})
    r = synth.uniform_random(1 << 23, seed=4)
    out = ref.encode(r)
    big.append({"name": "random_8MiB", "gen": "uniform_random(1<<23, seed=4)", "len": len(out),
                "sha256": hashlib.sha256(out).hexdigest(), "head": out[:4].hex()})
    r = synth.uniform_random(1 << 24, seed=4)
    out = ref.encode(r)
    big.append({"name": "random_16MiB", "gen": "uniform_random(1<<24, seed=4)", "len": len(out),
                "sha256": hashlib.sha256(out).hexdigest(), "head": out[:4].hex()})
    e = synth.english_like((1 << 20), seed=5)
    out = ref.encode(e)
    big.append({"name": "english_1MiB", "gen": "english_like(1<<20, seed=5)", "len": len(out),
                "sha256": hashlib.sha256(out).hexdigest(), "head": out[:16].hex()})
    d = synth.dna_like(1 << 20, seed=7)
    out = ref.encode(d)
    big.append({"name": "dna_1MiB", "gen": "dna_like(1<<20, seed=7)", "len": len(out),
                "sha256": hashlib.sha256(out).hexdigest(), "head": out[:16].hex()})
    m = synth.markov_text(300000, seed=42)
    out = ref.encode(m)
    big.append({"name": "markov_300k", "gen": "markov_text(300000, seed=42)", "len": len(out),
                "sha256": hashlib.sha256(out).hexdigest(), "head": out[:16].hex()})
    # vnibble known answers (ref lib/salz.c:352,565 are exported symbols)
    import ctypes as C
    vn = []
    for v in [0, 7, 8, 71, 72, 583, 584, 4679, 4680, 37447, 37448, 65535, 299591, 299592]:
        res = C.c_uint64(0)
        k = ref.lib.encode_vnibble_le(v, C.byref(res))
        vn.append({"v": v, "size": int(ref.lib.vnibble_size(v)), "nibbles": k,
                   "bits": res.value & ((1 << (4 * k)) - 1)})
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden.json")
    with open(path, "w") as f:
        json.dump({"source": "compiled reference salz @1bf47a3 + libsais 2.7.1 (oracle/_ref)",
                   "vectors": vectors, "digests": big, "vnibble": vn}, f, separators=(",", ":"))
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
