"""Differential test of the host decoder (lib/salz.c, row f1) on MALFORMED streams.

"Accept exactly what the reference accepts": valid streams are mutated (bit flips, byte
changes, truncation, header damage) and decoded three ways

  * the reference decoder (oracle/_ref, ref lib/salz.c:829-1228) - in a forked child, because
    it has undefined behaviour on some malformed inputs: `src = dst - factor_offs`
    (ref lib/salz.c:1147) is not checked against the bytes produced so far, and its 8-byte
    copy loop (ref lib/salz.c:1150-1163) may run past dst_len; it crashes (SIGSEGV / heap
    corruption) on about one mutated stream in eight;
  * lib/libsalz.so's salz_decode_safe (the product);
  * `model_decode` below: the reference's arithmetic step by step in Python, which also tells
    whether the stream makes the reference read before the start of the output.

Wherever the reference's behaviour is defined, the product must agree with it byte for byte
(return code, length, bytes); where the reference reads before its output buffer the product
must reject; the product must never crash.
"""
import ctypes as C
import faulthandler
import os
import pickle

import numpy as np
import pytest

from salz_b200 import synth

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
M32 = 0xFFFFFFFF


def model_decode(stream, cap):
    """-> (result bytes or None, reads_before_start).  Follows ref lib/salz.c:829-1192 with its
    32-bit arithmetic; a factor that reaches before the output start makes the result undefined."""
    if len(stream) < 4:
        return None, False
    hdr = int.from_bytes(stream[:4], "little")
    typ, plen = hdr >> 24, hdr & 0xFFFFFF
    if typ > 1 or plen > len(stream) - 4:
        return None, False
    src = stream[4:4 + plen]
    if typ == 0:
        return (None if plen > cap else bytes(src)), False
    pos, bits, avail = 0, 0, 0
    out = bytearray()

    def queue():
        nonlocal pos, bits, avail
        if pos + 8 > plen:
            return False
        bits = int.from_bytes(src[pos:pos + 8], "little")
        pos += 8
        avail = 64
        return True

    def read_bits(cnt):
        nonlocal bits, avail
        if avail == 0 and not queue():
            return None
        if cnt <= avail:
            v = bits >> (64 - cnt)
            bits = (bits << cnt) & (2**64 - 1)
            avail -= cnt
            return v
        v = bits >> (64 - avail)
        cnt -= avail
        if not queue():
            return None
        v = (v << cnt) | (bits >> (64 - cnt))
        bits = (bits << cnt) & (2**64 - 1)
        avail -= cnt
        return v

    def read_unary():
        nonlocal bits, avail
        if avail == 0 and not queue():
            return None
        res = 0
        while bits == 0:
            res = (res + avail) & M32
            if not queue():
                return None
        z = 64 - bits.bit_length()
        bits = (bits << (z + 1)) & (2**64 - 1)
        avail -= z + 1
        return (res + z) & M32

    while pos != plen:
        tok = read_bits(1)
        if tok is None:
            return None, False
        if tok == 0:
            if pos >= plen or len(out) >= cap:
                return None, False
            out.append(src[pos])
            pos += 1
            continue
        nib = read_bits(4)
        if nib is None:
            return None, False
        var = nib & 7
        k = 1
        while nib < 8 and k < 11:
            nib = read_bits(4)
            if nib is None:
                return None, False
            var = (((var + 1) << 3) | (nib & 7)) & M32
            k += 1
        if pos >= plen:
            return None, False
        offs = ((((var << 8) & M32) | src[pos]) + 1) & M32
        pos += 1
        q = read_unary()
        if q is None:
            return None, False
        low = read_bits(3)
        if low is None:
            return None, False
        ln = ((((q << 3) & M32) | low) + 3) & M32
        if len(out) + ln > cap:
            return None, False
        if offs > len(out) or offs == 0:
            return None, True   # the reference reads dst[-k]: undefined
        for _ in range(ln):
            out.append(out[-offs])
    return bytes(out), False


def _ours(lib, stream, cap):
    buf = C.create_string_buffer(cap + 64)
    n = C.c_size_t(cap)
    rc = lib.salz_decode_safe(bytes(stream), len(stream), buf, C.byref(n))
    assert buf.raw[cap:] == b"\0" * 64, "the decoder wrote past its capacity"
    return None if rc != 0 else buf.raw[: n.value]


def _ref_many(ref, items):
    """reference decoder over [(stream, cap)] in forked children; a crash is recorded as 'crash'"""
    out, i = [], 0
    while i < len(items):
        r, w = os.pipe()
        pid = os.fork()
        if pid == 0:
            os.close(r)
            faulthandler.disable()                       # a crash here is an expected outcome, not a report
            os.dup2(os.open(os.devnull, os.O_WRONLY), 2)  # glibc's heap-corruption messages
            try:
                with os.fdopen(w, "wb") as f:
                    for m, cap in items[i:]:
                        pickle.dump(ref.decode(m, cap), f)
                        f.flush()
            finally:
                os._exit(0)
        os.close(w)
        got = []
        with os.fdopen(r, "rb") as f:
            while True:
                try:
                    got.append(pickle.load(f))
                except (EOFError, pickle.UnpicklingError):
                    break
        _, status = os.waitpid(pid, 0)
        out += got
        i += len(got)
        if i < len(items) and status != 0:
            out.append("crash")
            i += 1
    return out


def _mutations(enc, rng, count):
    for trial in range(count):
        m = bytearray(enc)
        kind = trial % 6
        if kind == 0:
            m[int(rng.integers(len(m)))] ^= 1 << int(rng.integers(8))
        elif kind == 1:
            m[int(rng.integers(len(m)))] = int(rng.integers(256))
        elif kind == 2:
            m = m[: int(rng.integers(1, len(m) + 1))]
        elif kind == 3:   # header: type / length bytes
            m[int(rng.integers(min(len(m), 4)))] = int(rng.integers(256))
        elif kind == 4:
            for _ in range(3):
                m[int(rng.integers(len(m)))] = int(rng.integers(256))
        else:             # payload cut, header length adjusted to match
            keep = int(rng.integers(0, len(m) - 3))
            m = bytearray((keep | (m[3] << 24)).to_bytes(4, "little")) + m[4:4 + keep]
        yield bytes(m)


def test_decoder_agrees_with_reference_on_malformed_streams(ref):
    lib = C.CDLL(os.path.join(ROOT, "lib", "libsalz.so"))
    lib.salz_decode_safe.argtypes = [C.c_char_p, C.c_size_t, C.c_char_p, C.POINTER(C.c_size_t)]
    rng = np.random.default_rng(20261105)
    cases = [(n, d) for n, d in synth.small_cases(seed=13) if len(d) <= 6000]
    cases += [("eng", synth.english_like(5000, seed=1)), ("a", np.full(3000, 97, np.uint8)),
              ("rnd", synth.uniform_random(1500, seed=2))]
    items, names = [], []
    for name, data in cases:
        enc = ref.encode(data)
        if enc is None:
            continue
        for m in _mutations(enc, rng, 90):
            for cap in (len(data), max(0, len(data) - 1), len(data) + 13):
                items.append((m, cap))
                names.append(name)
    refs = _ref_many(ref, items)
    assert len(refs) == len(items)
    stat = dict(crash=0, undefined=0, accept=0, reject=0)
    for (m, cap), a, name in zip(items, refs, names):
        b = _ours(lib, m, cap)
        want, undefined = model_decode(m, cap)
        if undefined:
            stat["undefined"] += 1
            assert b is None, (name, "a factor reaching before the output start must be rejected")
            continue
        assert b == want, (name, cap, None if b is None else len(b), None if want is None else len(want))
        if a == "crash":   # the reference's wide copy ran past dst_len; the model stands in for it
            stat["crash"] += 1
            continue
        assert a == b, (name, cap, None if a is None else len(a), None if b is None else len(b))
        stat["accept" if b is not None else "reject"] += 1
    print("decoder fuzz:", len(items), "streams", stat)
    # the run must have exercised all classes
    assert stat["accept"] > 100 and stat["reject"] > 100 and stat["undefined"] > 10
