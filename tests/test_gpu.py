"""GPU tier (run on the B200 box: pytest -m gpu): the CUDA path, called through
the C ABI (include/salz_cuda.h via ctypes) and through the host library /
CLI, against the oracle.  Everything is integer/byte work: the bar is
bit-exact equality, stage by stage and on the final bytes."""
import ctypes as C
import hashlib
import json
import os
import subprocess
import sys

import numpy as np
import pytest

import salz_b200
from salz_b200 import synth

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "golden.json")))


def _first_diff(a, b):
    a = np.asarray(a); b = np.asarray(b)
    if a.shape != b.shape:
        return "shape %s vs %s" % (a.shape, b.shape)
    d = np.nonzero(a != b)
    if len(d[0]) == 0:
        return None
    i = tuple(x[0] for x in d)
    return "first diff at %s: got %s want %s (%d differ)" % (i, a[i], b[i], len(d[0]))


def _check_stages(enc, orc, name, data):
    got, want = enc.stages(data), orc.stages(data)
    n = len(data) - 8
    for k in ("sa", "plcp", "lcp"):
        assert _first_diff(got[k], want[k]) is None, (name, k, _first_diff(got[k], want[k]))
    # PSV / NSV as the reference's build_psvnsv_array leaves them (ref lib/salz.c:471-490)
    assert _first_diff(got["psvnsv"], want["psvnsv"]) is None, (name, "psvnsv", _first_diff(got["psvnsv"], want["psvnsv"]))
    assert _first_diff(got["cand"][1:], want["cand"][1:]) is None, (name, "cand")
    assert _first_diff(got["parse"][1:n], want["parse"][1:n]) is None, (name, "parse")
    assert got["out"] == want["out"], (name, "bytes")


def test_known_answers(enc):
    # SURVEY.md Appendix A
    assert enc.encode(b"abcdefghi").hex() == "09000000616263646566676869"
    assert enc.encode(b"a" * 64).hex() == "120000010000000000000c6061006161616161616161"
    assert enc.encode(b"abcabcabcabcabcabcabcabcXYZ01234").hex() == \
        "1400000100000000000028186162630258595a3031323334"


def test_golden_vectors(enc):
    for vec in GOLDEN["vectors"]:
        data = bytes.fromhex(vec["input"])
        st = enc.stages(data)
        assert st["out"].hex() == vec["output"], vec["name"]
        if "sa" in vec:
            assert st["sa"].tolist() == vec["sa"], vec["name"]
            assert st["lcp"].tolist() == vec["lcp"], vec["name"]
            assert st["cand"][1:].reshape(-1).tolist() == vec["cand"], vec["name"]
            assert st["parse"][1:].reshape(-1).tolist() == vec["parse"], vec["name"]


def test_golden_digests_1mib(enc):
    gens = {"english_1MiB": lambda: synth.english_like(1 << 20, seed=5),
            "dna_1MiB": lambda: synth.dna_like(1 << 20, seed=7),
            "markov_300k": lambda: synth.markov_text(300000, seed=42)}
    for d in GOLDEN["digests"]:
        if d["name"] in gens:
            out = enc.encode(gens[d["name"]]())
            assert len(out) == d["len"] and hashlib.sha256(out).hexdigest() == d["sha256"], d["name"]


@pytest.mark.parametrize("mode", [0, 1])
def test_small_cases_all_stages(enc, orc, mode):
    enc.set_sort_mode(mode)
    try:
        cases = synth.small_cases(3) + [
            ("zeros", np.zeros(100, np.uint8)),
            ("zero_tail", np.concatenate([np.arange(50, dtype=np.uint8) + 1, np.zeros(20, np.uint8)])),
            ("n9", np.frombuffer(b"abcdefghi", np.uint8)),
            ("n10", np.frombuffer(b"aaaaaaaaaa", np.uint8)),
            ("n11", np.frombuffer(b"abababababa", np.uint8)),
        ]
        for name, data in cases:
            _check_stages(enc, orc, name, data)
    finally:
        enc.set_sort_mode(1)


def test_medium_cases_all_stages(enc, orc):
    cases = [("eng64k", synth.english_like(65536 + 100, seed=9)),
             ("dna200k", synth.dna_like(200000, seed=3, block=200000)),
             ("a100k", np.full(100000, 97, np.uint8)),
             ("rand50k", synth.uniform_random(50000, seed=1)),
             ("ab_period", np.resize(np.frombuffer(b"ab", np.uint8), 70001)),
             ("markov100k", synth.markov_text(100000, seed=42)),
             ("eng1M", synth.english_like(1 << 20, seed=5)),
             ("dna1M", synth.dna_like(1 << 20, seed=7)),
             ("a1M", np.full(1 << 20, 0x61, np.uint8)),
             ("rand1M", synth.uniform_random(1 << 20, seed=4))]
    # factors that jump over many parse chunks between stretches of ordinary text (the chunk-entry
    # chain of emit.cu is walked speculatively per range of chunks and must survive being skipped)
    eng = synth.english_like(400000, seed=17)
    cases.append(("eng_longcopies", np.concatenate([eng[:200000], eng[:150000], eng[200000:260000],
                                                    eng[100000:190000], eng[260000:], eng[:70000]])))
    for name, data in cases:
        _check_stages(enc, orc, name, data)


def test_adversarial_suffix_structures(enc, orc):
    """Fibonacci and Thue-Morse words, a de Bruijn sequence, runs of every length, a long period with
    one defect: many doubling rounds, groups that cross every window, the bag reorder and the
    unchanged-window fast path, one-chain parses.  All stages against the reference; then the same
    inputs as ONE ragged batch (different alphabets and round-0 key widths per block), and in a
    context whose block capacity is above 2^21 (wide elements, pairs on the global path)."""
    cases = synth.adversarial_cases()
    for name, data in cases:
        _check_stages(enc, orc, name, data)
    big = synth.fibonacci_word(1 << 20)
    assert enc.encode(big) == orc.encode(big), "fibonacci 1 MiB"
    outs = enc.encode_blocks([c[1] for c in cases])
    for (name, data), got in zip(cases, outs):
        assert got == orc.encode(data), ("batched", name)
    e = salz_b200.SalzCuda(0, 1 << 22, 2)
    try:
        for name, data in cases[:4] + [("fibonacci3M", synth.fibonacci_word(3 << 20))]:
            assert e.encode(data) == orc.encode(data), ("wide", name)
    finally:
        e.close()


def test_order_check_fallback_path(tmp_path):
    """The default radix ranking takes the slot returned by the counting atomicAdd and VERIFIES that
    every pass was stable; if the check ever fails the suffix sort is repeated with the peer-mask
    ranking.  The check does not fail on this hardware, so the repeat path is forced here, in a
    process of its own (the switch is read once per process)."""
    script = tmp_path / "fb.py"
    script.write_text(
        "import sys, numpy as np\n"
        "sys.path.insert(0, %r)\n"
        "import salz_b200\n"
        "from salz_b200 import synth\n"
        "from oracle.binding import Ref\n"
        "orc = Ref()\n"
        "e = salz_b200.SalzCuda(0, 1 << 20, 4)\n"
        "for d in (synth.english_like(700000, seed=3), synth.dna_like(300000, seed=4, block=300000), np.full(50000, 97, np.uint8)):\n"
        "    assert e.encode(d) == orc.encode(d)\n"
        "outs = e.encode_blocks([synth.english_like(1 << 20, seed=s) for s in (5, 6, 7)])\n"
        "assert all(o == orc.encode(synth.english_like(1 << 20, seed=s)) for o, s in zip(outs, (5, 6, 7)))\n"
        "print('ok')\n" % ROOT)
    env = dict(os.environ, SALZ_FORCE_ORDER_FALLBACK="1")
    p = subprocess.run([sys.executable, str(script)], capture_output=True, text=True, env=env, timeout=600)
    assert p.returncode == 0 and "ok" in p.stdout, p.stderr[-2000:]
    assert "repeating with the peer-mask ranking" in p.stderr
    # and without the switch nothing is repeated
    p = subprocess.run([sys.executable, str(script)], capture_output=True, text=True, timeout=600)
    assert p.returncode == 0 and "ok" in p.stdout, p.stderr[-2000:]
    assert "repeating" not in p.stderr


@pytest.mark.parametrize("k", [1, 2, 8, 64, -4, -64])
def test_parse_segments_exact_for_every_cut(enc, orc, k):
    """The segmented parse (estimates to the right of a segment, passes until nothing
    changes) must give the reference's (offs, len, cost) and bytes for any number of
    segments, on text, deep repeats, one long factor and incompressible data.  k < 0 forces
    the shortened first pass (only the left end of every segment is walked)."""
    enc.set_parse_segments(k)
    try:
        cases = [("eng1M", synth.english_like(1 << 20, seed=11)),
                 ("dna1M", synth.dna_like(1 << 20, seed=12)),
                 ("a600k", np.full(600000, 0x61, np.uint8)),
                 ("period7", np.resize(np.frombuffer(b"abcdefg", np.uint8), 300001)),
                 ("rand300k", synth.uniform_random(300000, seed=13)),
                 ("markov300k", synth.markov_text(300000, seed=42))]
        for name, data in cases:
            _check_stages(enc, orc, "%s/k%d" % (name, k), data)
    finally:
        enc.set_parse_segments(0)


def test_alphabet_edge_cases(enc, orc):
    """Round 0 sorts dense codes of the block's alphabet (1..sigma, 0 = past the end):
    every sigma class, byte values 0 and 255, and suffixes that end inside the first key."""
    rng = np.random.default_rng(21)
    cases = [("sigma256", np.tile(np.arange(256, dtype=np.uint8), 40)),
             ("sigma255", rng.integers(1, 256, 30000).astype(np.uint8)),
             ("sigma2_0_255", rng.choice(np.array([0, 255], np.uint8), 40000)),
             ("sigma3", rng.choice(np.array([0, 7, 200], np.uint8), 50000)),
             ("sigma17", rng.integers(100, 117, 50000).astype(np.uint8)),
             ("sigma33", rng.integers(0, 33, 50000).astype(np.uint8)),
             ("zeros_tail", np.concatenate([rng.integers(0, 4, 20000).astype(np.uint8), np.zeros(64, np.uint8)])),
             ("ff_run", np.full(5000, 255, np.uint8)),
             ("short_keys", np.frombuffer(b"abracadabra" * 3, np.uint8))]
    for name, data in cases:
        _check_stages(enc, orc, name, data)
    sas = enc.suffix_arrays([c[1] for c in cases])
    for (name, data), sa in zip(cases, sas):
        assert np.array_equal(sa, orc.suffix_array(data)), name


def test_encode_blocks_error_semantics(enc, orc):
    """salz_encode_safe conventions (ref lib/salz.c:777-823)."""
    data = synth.english_like(5000, seed=3)
    full = orc.encode(data)
    blocks = [b"", b"x" * 7, b"x" * 8, data, b"x" * 9]
    out = enc.encode_blocks(blocks)
    assert out[0] is None and out[1] is None and out[2] is None
    assert out[3] == full and out[4] == orc.encode(b"x" * 9)
    for cap in (3, 4, len(full) - 1, len(full), len(full) + 1):
        assert enc.encode(data, cap=cap) == orc.encode(data, cap=cap), cap
    rnd = synth.uniform_random(5000, seed=3)
    for cap in (5004, 5003, 5004 + 8 * 79 - 1, 5004 + 8 * 79):
        assert enc.encode(rnd, cap=cap) == orc.encode(rnd, cap=cap), cap


def test_batched_stream_matches_per_block_oracle(enc, orc):
    """CLI block loop (ref programs/salzcli.c:143-179): framed records, ragged tail,
    more blocks than one batch holds, PLAIN and SALZ blocks mixed."""
    bl = 1 << 18
    data = np.concatenate([synth.english_like(9 * bl, seed=11), synth.dna_like(4 * bl, seed=12, block=bl),
                           synth.uniform_random(2 * bl, seed=13), np.full(3 * bl + 777, 98, np.uint8)])
    stream, sizes = enc.encode_stream(data, bl)
    want = b""
    wsizes = []
    for off in range(0, len(data), bl):
        blk = orc.encode(data[off: off + bl])
        wsizes.append(len(blk))
        want += len(blk).to_bytes(4, "little") + blk
    assert sizes.tolist() == wsizes
    assert stream == want


def test_pipelined_host_stream_paths(orc):
    """salz_cuda_encode_stream with host buffers overlaps copies and kernels (two text and
    two output buffers): a stream that fits one batch is cut in two (1:3), a longer one runs
    batch after batch; pinned and pageable memory; the device-resident entry on the same
    context afterwards.  All must give the per-block reference records."""
    import torch
    bl = 1 << 16
    e = salz_b200.SalzCuda(0, bl, 48)
    try:
        data = np.concatenate([synth.english_like(70 * bl, seed=31), synth.uniform_random(5 * bl, seed=32),
                               synth.dna_like(60 * bl + 4321, seed=33, block=bl)])   # 136 blocks: 3 batches of 48
        want = b""
        for off in range(0, len(data), bl):
            blk = orc.encode(data[off: off + bl])
            want += len(blk).to_bytes(4, "little") + blk
        stream, sizes = e.encode_stream(data, bl)
        assert stream == want
        short = data[: 40 * bl + 999]                                            # 41 blocks: one batch, cut in two (16 + 25)
        nrec = 0
        pos = 0
        for _ in range(41):
            ln = int.from_bytes(want[pos: pos + 4], "little")
            pos += 4 + ln
            nrec += 1
            if nrec == 40:
                break
        stream2, sizes2 = e.encode_stream(short, bl)
        assert stream2[:pos] == want[:pos] and len(sizes2) == 41
        # pinned source and destination through the pointer entry
        src = torch.from_numpy(data.copy()).pin_memory()
        cap = salz_b200.api.stream_len_max(len(data), bl)
        dst = torch.empty(cap, dtype=torch.uint8).pin_memory()
        n = e.encode_stream_ptr(src.data_ptr(), len(data), bl, dst.data_ptr(), cap)
        n = n[0] if isinstance(n, tuple) else n
        assert bytes(dst[:n].numpy()) == want
        # the other entries still work on the same context (buffers restored)
        assert e.encode(data[:bl]) == orc.encode(data[:bl])
    finally:
        e.close()
    # other batch capacities: one batch cut in two (34 + 102) / two full batches (ragged last block)
    for bmax in (160, 68):
        e = salz_b200.SalzCuda(0, bl, bmax)
        try:
            stream, sizes = e.encode_stream(data, bl)
            assert stream == want and len(sizes) == 136
        finally:
            e.close()


def test_incompressible_tail_block_is_plain_like_the_cli(enc, orc):
    """A short, nearly incompressible tail block whose SALZ form overshoots ITS OWN
    salz_encoded_len_max by a few bytes (a factor saves two direct bytes but can open one more
    8-byte bit word): the reference CLI hands every block the capacity of a full block (ref
    programs/salzcli.c:129), so it writes the tail as PLAIN; salz_encode_safe with the tail's own
    capacity fails.  The stream entry follows the CLI, the block entry the API."""
    bl = 1 << 16
    tail = synth.uniform_random(4096, seed=3128)   # the tail of bench.py's 1 GiB mixed stream
    full_cap = salz_b200.encoded_len_max(bl)
    assert orc.encode(tail) is None and len(orc.encode(tail, cap=full_cap)) == 4100
    data = np.concatenate([synth.english_like(2 * bl, seed=71), tail])
    stream, sizes = enc.encode_stream(data, bl)
    want = b""
    for off in range(0, len(data), bl):
        blk = orc.encode(data[off: off + bl], cap=full_cap)
        want += len(blk).to_bytes(4, "little") + blk
    assert sizes.tolist()[-1] == 4100 and stream == want
    assert enc.encode(tail) is None                              # per-block API: the tail's own capacity, like the reference
    assert enc.encode(tail, cap=full_cap) == orc.encode(tail, cap=full_cap)


def test_stream_rejects_tiny_tail(enc):
    bl = 1 << 15
    with pytest.raises(salz_b200.SalzCudaError):
        enc.encode_stream(np.zeros(bl + 8, np.uint8), bl)


def test_suffix_arrays_entry(enc, orc):
    texts = [synth.english_like(50000, seed=1), synth.dna_like(70001, seed=2, block=70001),
             np.full(30000, 7, np.uint8), synth.uniform_random(9999, seed=5), np.frombuffer(b"banana", np.uint8)]
    sas = enc.suffix_arrays(texts)
    for t, sa in zip(texts, sas):
        assert np.array_equal(sa, orc.suffix_array(t))


def test_level9_block(orc):
    """One 16 MiB block (CLI level 9): bytes identical, header quirk F4 on random data."""
    e = salz_b200.SalzCuda(0, 1 << 24, 1)
    try:
        data = synth.english_like(1 << 24, seed=31)
        assert e.encode(data) == orc.encode(data)
        a = np.full(1 << 24, 0x61, np.uint8)
        out = e.encode(a)
        d = [x for x in GOLDEN["digests"] if x["name"] == "a_x_16MiB"][0]
        assert len(out) == d["len"] == 262166 and hashlib.sha256(out).hexdigest() == d["sha256"]
        r = synth.uniform_random(1 << 24, seed=4)
        out = e.encode(r)
        d = [x for x in GOLDEN["digests"] if x["name"] == "random_16MiB"][0]
        assert out[:4].hex() == "00000000" == d["head"]
        assert len(out) == d["len"] and hashlib.sha256(out).hexdigest() == d["sha256"]
    finally:
        e.close()


def test_repeated_runs_are_deterministic(orc):
    """The sort leans on relaxed loads/stores for two decoupled look-backs, on ranks read and
    written in one launch (stamps) and on self-cleaning shared-memory masks; compute-sanitizer's
    racecheck is closed on this GPU pool (profiles/r02_sanitizer_closed.txt).  A race there shows
    up as run-to-run differences under different schedules: the same stream, many times, three
    batch shapes (different tile interleavings), both ranking modes - always the reference's bytes."""
    bl = 1 << 16
    data = np.concatenate([synth.english_like(37 * bl, seed=51), synth.dna_like(20 * bl, seed=52, block=bl),
                           np.full(6 * bl, 0x41, np.uint8), synth.uniform_random(3 * bl + 999, seed=53)])
    want = b""
    for off in range(0, len(data), bl):
        blk = orc.encode(data[off: off + bl])
        want += len(blk).to_bytes(4, "little") + blk
    for bmax in (67, 23, 5):
        e = salz_b200.SalzCuda(0, bl, bmax)
        try:
            for rep in range(6):
                e.set_sort_mode(rep & 1)
                assert e.encode_stream(data, bl)[0] == want, (bmax, rep)
        finally:
            e.close()


def test_wide_blocks_all_stages(orc):
    """Block capacity above 2^21 (CLI levels 7-9): group, second key and position no longer fit
    one u64, so the bag holds (group, p), k_local works with window-local group indices and the
    groups that cross windows travel as pairs.  Same stages, same bytes - on inputs small enough
    to compare every stage, and on one 4 MiB block of each kind."""
    e = salz_b200.SalzCuda(0, 1 << 22, 2)
    try:
        cases = synth.small_cases(5)[:15] + [
            ("eng300k", synth.english_like(300000, seed=3)),
            ("dna300k", synth.dna_like(300000, seed=4, block=300000)),
            ("a200k", np.full(200000, 0x61, np.uint8)),
            ("period5", np.resize(np.frombuffer(b"abcab", np.uint8), 150001)),
            ("rand100k", synth.uniform_random(100000, seed=6)),
        ]
        for name, data in cases:
            _check_stages(e, orc, "wide/" + name, data)
        for name, data in (("eng4M", synth.english_like(1 << 22, seed=7)), ("dna4M", synth.dna_like(1 << 22, seed=8, block=1 << 22)),
                           ("a4M", np.full(1 << 22, 0x62, np.uint8))):
            assert e.encode(data) == orc.encode(data), name
        texts = [c[1] for c in cases[15:]]
        for t, sa in zip(texts, e.suffix_arrays(texts)):
            assert np.array_equal(sa, orc.suffix_array(t))
    finally:
        e.close()


def _host_lib():
    subprocess.run(["make", "-s", "-C", os.path.join(ROOT, "lib")], check=True)
    lib = C.CDLL(os.path.join(ROOT, "lib", "libsalz.so"))
    lib.salz_encode_safe.argtypes = [C.c_char_p, C.c_size_t, C.c_char_p, C.POINTER(C.c_size_t)]
    lib.salz_decode_safe.argtypes = [C.c_char_p, C.c_size_t, C.c_char_p, C.POINTER(C.c_size_t)]
    return lib


def test_host_library_encode_roundtrip(orc):
    """lib/salz.h entry points: salz_encode_safe runs the CUDA path, salz_decode_safe restores."""
    lib = _host_lib()
    for data in (synth.english_like(300000, seed=1).tobytes(), b"abcdefghi", bytes(5000)):
        cap = salz_b200.encoded_len_max(len(data))
        dst = C.create_string_buffer(cap)
        n = C.c_size_t(cap)
        assert lib.salz_encode_safe(data, len(data), dst, C.byref(n)) == 0
        assert dst.raw[: n.value] == orc.encode(data)
        back = C.create_string_buffer(len(data) + 8)
        m = C.c_size_t(len(data))
        assert lib.salz_decode_safe(dst.raw[: n.value], n.value, back, C.byref(m)) == 0
        assert back.raw[: m.value] == data
    n = C.c_size_t(100)
    assert lib.salz_encode_safe(b"12345678", 8, C.create_string_buffer(100), C.byref(n)) == -1


def test_host_library_concurrent_callers(orc):
    """salz_encode_safe from several threads at once on different buffers (the reference is
    re-entrant, ref lib/salz.c:777-823 keeps everything in a per-call ctx; this library
    serialises the callers on its process-wide CUDA context): every caller gets its own
    block's bytes."""
    from concurrent.futures import ThreadPoolExecutor
    lib = _host_lib()
    inputs = [synth.english_like(200000 + 1000 * i, seed=10 + i).tobytes() for i in range(4)]
    inputs += [synth.dna_like(150000, seed=20, block=150000).tobytes(), bytes(70000), b"abcdefghi" * 3]
    want = [orc.encode(d) for d in inputs]

    def one(d):
        cap = salz_b200.encoded_len_max(len(d))
        dst = C.create_string_buffer(cap)
        n = C.c_size_t(cap)
        rc = lib.salz_encode_safe(d, len(d), dst, C.byref(n))   # ctypes drops the GIL for the call
        return rc, dst.raw[: n.value]

    for _ in range(3):
        with ThreadPoolExecutor(len(inputs)) as ex:
            got = list(ex.map(one, inputs))
        for (rc, out), w in zip(got, want):
            assert rc == 0 and out == w


def _stream_properties(orc, data, bl, sample):
    """Size-independent checks of a whole stream: every record decodes back to its block
    (host decoder), the sizes add up to the stream, sampled blocks equal the oracle's bytes."""
    lib = _host_lib()
    nblocks = (len(data) + bl - 1) // bl
    e = salz_b200.SalzCuda(0, bl, min(nblocks, max(1, (64 << 20) // bl + 1)))
    try:
        stream, sizes = e.encode_stream(data, bl)
    finally:
        e.close()
    assert len(sizes) == nblocks
    assert len(stream) == int(sizes.astype(np.int64).sum()) + 4 * len(sizes)
    pos = 0
    back = C.create_string_buffer(bl + 8)
    picks = {0, len(sizes) - 1} | {int(x) for x in sample if x < len(sizes)}
    for b, sz in enumerate(sizes.tolist()):
        assert int.from_bytes(stream[pos: pos + 4], "little") == sz
        rec = stream[pos + 4: pos + 4 + sz]
        blk = data[b * bl: (b + 1) * bl].tobytes()
        m = C.c_size_t(len(blk))
        assert lib.salz_decode_safe(rec, sz, back, C.byref(m)) == 0 and back.raw[: m.value] == blk, b
        if b in picks:
            assert rec == orc.encode(blk), b
        pos += 4 + sz
    assert pos == len(stream)


def test_full_size_roundtrip_properties(orc):
    """64 MiB + 4096 English-like, level 5 (the bench shape, BASELINE configs[1])."""
    _stream_properties(orc, synth.english_like((64 << 20) + 4096, seed=1), 1 << 20, (17, 40))


@pytest.mark.parametrize("kind,mib,level,sample", [
    ("dna", 1024, 5, (3, 31, 700)),  # configs[2] at full size: planted repeats up to 2^19, ~20 doubling rounds
    ("mixed", 256, 5, (7, 9, 20, 150)),  # configs[3] (one GPU's share): english / dna / random sections in one stream
    ("single", 512, 5, (5, 300)),    # configs[4]a at full size: every round on the global path, one parse chain per block
    ("random", 512, 5, (11, 400)),   # configs[4]b at full size: every block PLAIN
    ("mixed", 48, 9, (1,)),           # 16 MiB blocks: pairs format
    ("english", 32, 1, (100, 301)),   # 64 KiB blocks: hundreds of blocks per batch
])
def test_full_size_properties_other_shapes(orc, kind, mib, level, sample):
    bl = 1 << (15 + level)
    size = (mib << 20) + 4096
    data = {"dna": lambda: synth.dna_like(size, seed=2, block=bl),
            "mixed": lambda: synth.mixed_stream(size, seed=3),
            "single": lambda: synth.single_byte(size),
            "random": lambda: synth.uniform_random(size, seed=4),
            "english": lambda: synth.english_like(size, seed=1)}[kind]()
    _stream_properties(orc, data, bl, sample)


def test_cli_compress_matches_reference_cli(tmp_path):
    """`salz -k -5 file` writes the same bytes as the reference CLI and the reference decodes it."""
    ref_cli = os.path.join(ROOT, "oracle", "_ref", "salz_ref")
    if not os.path.exists(ref_cli):
        pytest.skip("prebuilt reference CLI absent")
    _host_lib()
    data = np.concatenate([synth.english_like(5 << 20, seed=2), synth.uniform_random((1 << 20) + 4096, seed=3)])
    a = tmp_path / "a.bin"; b = tmp_path / "b.bin"
    a.write_bytes(data.tobytes()); b.write_bytes(data.tobytes())
    env = dict(os.environ, SALZ_BATCH_MIB="4")
    subprocess.run([os.path.join(ROOT, "programs", "salz"), "-k", "-q", "-5", str(a)], check=True, env=env)
    subprocess.run([ref_cli, "-k", "-q", "-5", str(b)], check=True)
    assert (tmp_path / "a.bin.salz").read_bytes() == (tmp_path / "b.bin.salz").read_bytes()
    a.unlink()
    subprocess.run([ref_cli, "-d", "-q", str(a) + ".salz"], check=True)
    assert a.read_bytes() == data.tobytes()
    # exact multiple of the block size fails like the reference
    c = tmp_path / "c.bin"
    c.write_bytes(bytes(1 << 15))
    assert subprocess.run([os.path.join(ROOT, "programs", "salz"), "-k", "-q", "-0", str(c)]).returncode != 0


def test_reference_cli_linked_against_this_library(tmp_path):
    """Link-level drop-in: the unmodified reference CLI, linked to
    lib/libsalz.so, writes the same file as the reference CLI with its own encoder."""
    dropin = os.path.join(ROOT, "oracle", "_ref", "salz_ref_dropin")
    ref_cli = os.path.join(ROOT, "oracle", "_ref", "salz_ref")
    if not (os.path.exists(dropin) and os.path.exists(ref_cli)):
        pytest.skip("prebuilt reference CLIs absent")
    _host_lib()
    data = np.concatenate([synth.english_like(3 * (1 << 17) + 5000, seed=6), synth.uniform_random(1 << 17, seed=7),
                           np.full(70000, 65, np.uint8)])
    a = tmp_path / "a.bin"; b = tmp_path / "b.bin"
    a.write_bytes(data.tobytes()); b.write_bytes(data.tobytes())
    subprocess.run([dropin, "-k", "-q", "-2", str(a)], check=True)
    subprocess.run([ref_cli, "-k", "-q", "-2", str(b)], check=True)
    assert (tmp_path / "a.bin.salz").read_bytes() == (tmp_path / "b.bin.salz").read_bytes()
