"""ctypes bindings for the test-only oracle libraries.

`Port`  = oracle/libsalz_oracle.so  (our C restatement, salz_oracle.c)
`Ref`   = oracle/_ref/libsalz_ref.so (the upstream reference compiled in place,
          ref_shim.c; prebuilt in the authoring container, travels to the GPU box)
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
u8p = C.POINTER(C.c_uint8)
i32p = C.POINTER(C.c_int32)


def build(ref_root="/root/reference"):
    """Compile the restatement and, when the reference tree is present, _ref."""
    subprocess.run(["make", "-s", "-C", HERE, f"REF={ref_root}"], check=True)


def _ptr(a, t):
    return a.ctypes.data_as(t) if a is not None else None


def _as_u8(data):
    if isinstance(data, (bytes, bytearray, memoryview)):
        return np.frombuffer(bytes(data), dtype=np.uint8)
    return np.ascontiguousarray(data, dtype=np.uint8)


def encoded_len_max(n):
    """lib/salz.h:25-28"""
    return 4 + n + ((n + 63) // 64) * 64 // 8


class _Base:
    enc_name = dec_name = None

    def encode(self, data, cap=None):
        src = _as_u8(data)
        cap = encoded_len_max(len(src)) if cap is None else cap
        dst = np.zeros(max(cap, 1), dtype=np.uint8)
        n = C.c_size_t(cap)
        rc = getattr(self.lib, self.enc_name)(_ptr(src, u8p), len(src), _ptr(dst, u8p), C.byref(n))
        return None if rc != 0 else dst[: n.value].tobytes()

    def decode(self, data, cap):
        src = _as_u8(data)
        dst = np.zeros(max(cap, 1), dtype=np.uint8)
        n = C.c_size_t(cap)
        rc = getattr(self.lib, self.dec_name)(_ptr(src, u8p), len(src), _ptr(dst, u8p), C.byref(n))
        return None if rc != 0 else dst[: n.value].tobytes()


class Port(_Base):
    enc_name, dec_name = "oracle_encode", "oracle_decode"

    def __init__(self):
        path = os.path.join(HERE, "libsalz_oracle.so")
        if not os.path.exists(path):
            build()
        self.lib = L = C.CDLL(path)
        L.oracle_encode.argtypes = [u8p, C.c_size_t, u8p, C.POINTER(C.c_size_t)]
        L.oracle_decode.argtypes = [u8p, C.c_size_t, u8p, C.POINTER(C.c_size_t)]
        L.oracle_stages.argtypes = [u8p, C.c_size_t, i32p, i32p, i32p, i32p, i32p, u8p, C.POINTER(C.c_size_t)]
        L.oracle_suffix_array.argtypes = [u8p, C.c_int32, i32p]
        L.oracle_vnibble_size.argtypes = [C.c_uint32]
        L.oracle_vnibble_size.restype = C.c_size_t
        L.oracle_factor_cost.argtypes = [C.c_int32, C.c_int32]

    def suffix_array(self, data):
        src = _as_u8(data)
        sa = np.zeros(len(src), dtype=np.int32)
        assert self.lib.oracle_suffix_array(_ptr(src, u8p), len(src), _ptr(sa, i32p)) == 0
        return sa

    def stages(self, data):
        """dict(sa, plcp, lcp, cand[n,4], parse[n,3], out) for one block of N bytes."""
        src = _as_u8(data)
        N = len(src)
        n = N - 8
        sa = np.zeros(n, np.int32); plcp = np.zeros(n, np.int32); lcp = np.zeros(n, np.int32)
        cand = np.zeros((n, 4), np.int32); parse = np.zeros((n, 3), np.int32)
        cap = encoded_len_max(N)
        dst = np.zeros(cap, np.uint8); dl = C.c_size_t(cap)
        rc = self.lib.oracle_stages(_ptr(src, u8p), N, _ptr(sa, i32p), _ptr(plcp, i32p), _ptr(lcp, i32p),
                                    _ptr(cand, i32p), _ptr(parse, i32p), _ptr(dst, u8p), C.byref(dl))
        if rc != 0:
            return None
        return dict(sa=sa, plcp=plcp, lcp=lcp, cand=cand, parse=parse, out=dst[: dl.value].tobytes())


class Ref(_Base):
    enc_name, dec_name = "salz_encode_safe", "salz_decode_safe"

    @staticmethod
    def available():
        return os.path.exists(os.path.join(HERE, "_ref", "libsalz_ref.so"))

    def __init__(self):
        self.lib = L = C.CDLL(os.path.join(HERE, "_ref", "libsalz_ref.so"))
        L.salz_encode_safe.argtypes = [u8p, C.c_size_t, u8p, C.POINTER(C.c_size_t)]
        L.salz_decode_safe.argtypes = [u8p, C.c_size_t, u8p, C.POINTER(C.c_size_t)]
        L.ref_stages.argtypes = [u8p, C.c_size_t, i32p, i32p, i32p, i32p, u8p, C.POINTER(C.c_size_t)]
        L.ref_suffix_array.argtypes = [u8p, i32p, C.c_int32]
        L.ref_suffix_array_omp.argtypes = [u8p, i32p, C.c_int32, C.c_int32]
        L.ref_plcp_lcp.argtypes = [u8p, i32p, i32p, i32p, C.c_int32]
        L.vnibble_size.argtypes = [C.c_uint32]
        L.vnibble_size.restype = C.c_size_t
        L.encode_vnibble_le.argtypes = [C.c_uint32, C.POINTER(C.c_uint64)]
        L.encode_vnibble_le.restype = C.c_size_t

    def suffix_array(self, data, threads=1):
        src = _as_u8(data)
        sa = np.zeros(len(src), dtype=np.int32)
        if threads > 1:
            rc = self.lib.ref_suffix_array_omp(_ptr(src, u8p), _ptr(sa, i32p), len(src), threads)
        else:
            rc = self.lib.ref_suffix_array(_ptr(src, u8p), _ptr(sa, i32p), len(src))
        assert rc == 0
        return sa

    def plcp_lcp(self, data, sa):
        src = _as_u8(data)
        n = len(src)
        plcp = np.zeros(n, np.int32); lcp = np.zeros(n, np.int32)
        assert self.lib.ref_plcp_lcp(_ptr(src, u8p), _ptr(sa, i32p), _ptr(plcp, i32p), _ptr(lcp, i32p), n) == 0
        return plcp, lcp

    def stages(self, data):
        src = _as_u8(data)
        N = len(src)
        n = N - 8
        sa = np.zeros(n, np.int32); pn = np.zeros((n, 2), np.int32)
        cand = np.zeros((n, 4), np.int32); parse = np.zeros((n, 3), np.int32)
        cap = encoded_len_max(N)
        dst = np.zeros(cap, np.uint8); dl = C.c_size_t(cap)
        rc = self.lib.ref_stages(_ptr(src, u8p), N, _ptr(sa, i32p), _ptr(pn, i32p), _ptr(cand, i32p),
                                 _ptr(parse, i32p), _ptr(dst, u8p), C.byref(dl))
        if rc != 0:
            return None
        plcp, lcp = self.plcp_lcp(src[:n], sa)
        return dict(sa=sa, psvnsv=pn, plcp=plcp, lcp=lcp, cand=cand, parse=parse, out=dst[: dl.value].tobytes())
