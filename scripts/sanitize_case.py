"""Small end-to-end case for compute-sanitizer (memcheck / racecheck): every kernel of the
pipeline runs at least once (packed local path, global radix path, PLAIN and SALZ blocks,
multi-pass scatter launch), and the bytes are checked against the compiled reference."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import salz_b200  # noqa: E402
from salz_b200 import synth  # noqa: E402
from oracle.binding import Ref  # noqa: E402  (checker only)


def main():
    orc = Ref()
    bl = 1 << 15
    enc = salz_b200.SalzCuda(device=0, max_block_len=bl, max_batch_blocks=4)
    data = np.concatenate([synth.english_like(2 * bl, seed=1), synth.dna_like(bl, seed=2, block=bl),
                           synth.uniform_random(bl, seed=3), np.full(bl + 777, 0x61, np.uint8)])
    stream, sizes = enc.encode_stream(data, bl)
    want = b""
    for off in range(0, len(data), bl):
        blk = orc.encode(data[off: off + bl])
        want += len(blk).to_bytes(4, "little") + blk
    assert stream == want, "stream differs from the reference"
    for name, d in (("eng", synth.english_like(20000, seed=5)), ("a", np.full(9000, 7, np.uint8)),
                    ("per", np.resize(np.frombuffer(b"abcab", np.uint8), 12345))):
        got, ref = enc.stages(d), orc.stages(d)
        for k in ("sa", "lcp", "psvnsv"):
            assert np.array_equal(got[k], ref[k]), (name, k)
        assert got["out"] == ref["out"], name
    enc.close()
    print("sanitize_case ok: %d blocks, %d bytes" % (len(sizes), len(stream)))


if __name__ == "__main__":
    main()
