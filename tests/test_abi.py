"""CPU tier: the C-ABI library loads and exports every symbol include/*.h
declares; the host library keeps the reference's symbols; the product fails
loudly (never falls back) without a GPU; the host decoder is exact."""
import ctypes as C
import os
import re
import subprocess

import numpy as np
import pytest

import salz_b200
from salz_b200 import synth
from salz_b200.build import build

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _no_gpu():
    return salz_b200.load_library().salz_cuda_device_count() == 0


def test_build_and_exports():
    build()
    lib = salz_b200.load_library()
    hdr = open(os.path.join(ROOT, "include", "salz_cuda.h")).read()
    declared = set(re.findall(r"\b(salz_cuda_[a-z_]+)\s*\(", hdr))
    assert declared, "no declarations parsed"
    for name in declared:
        assert hasattr(lib, name), name
    assert declared == set(salz_b200.api.EXPORTS)


def test_host_library_symbols():
    subprocess.run(["make", "-s", "-C", os.path.join(ROOT, "lib")], check=True)
    lib = C.CDLL(os.path.join(ROOT, "lib", "libsalz.so"))
    for name in ("salz_encode_safe", "salz_decode_safe", "encode_vnibble_le", "vnibble_size"):
        assert hasattr(lib, name), name


def test_no_cpu_fallback():
    if not _no_gpu():
        pytest.skip("GPU present")
    with pytest.raises(salz_b200.SalzCudaError):
        salz_b200.SalzCuda()
    lib = C.CDLL(os.path.join(ROOT, "lib", "libsalz.so"))
    src = (C.c_uint8 * 100)(*([65] * 100))
    dst = (C.c_uint8 * 200)()
    n = C.c_size_t(200)
    assert lib.salz_encode_safe(src, 100, dst, C.byref(n)) == -1


def test_multi_gpu_entry_points_fail_loudly_without_gpu():
    """salz_cuda_multi_create / salz_cuda_comm_create: no device -> NULL -> the mirror raises
    (no CPU stand-in for the sharded path either); the block ranges need no device."""
    from salz_b200 import api
    assert api.block_range(8193, 0, 8) == (0, 1025) and api.block_range(8193, 7, 8) == (7169, 8193)
    if not _no_gpu():
        pytest.skip("GPU present")
    with pytest.raises(salz_b200.SalzCudaError):
        salz_b200.SalzCudaMulti(ndev=0)
    with pytest.raises(salz_b200.SalzCudaError):
        salz_b200.ShardComm(0, bytes(128), 0, 1)


def test_host_decoder_and_vnibble(port, ref):
    lib = C.CDLL(os.path.join(ROOT, "lib", "libsalz.so"))
    lib.salz_decode_safe.argtypes = [C.c_char_p, C.c_size_t, C.c_char_p, C.POINTER(C.c_size_t)]
    lib.vnibble_size.restype = C.c_size_t
    lib.encode_vnibble_le.restype = C.c_size_t
    cases = synth.small_cases(seed=13) + [("eng", synth.english_like(300000, seed=1)),
                                          ("a", np.full(200000, 97, np.uint8)),
                                          ("rnd", synth.uniform_random(70000, seed=2))]
    for name, data in cases:
        enc = ref.encode(data)
        if enc is None:  # nearly incompressible block overflowing encoded_len_max
            continue
        buf = C.create_string_buffer(len(data) + 8)
        n = C.c_size_t(len(data))
        assert lib.salz_decode_safe(enc, len(enc), buf, C.byref(n)) == 0, name
        assert buf.raw[: n.value] == data.tobytes(), name
        # truncated and corrupted streams must fail cleanly or decode to something bounded
        n = C.c_size_t(len(data))
        assert lib.salz_decode_safe(enc, 3, buf, C.byref(n)) == -1
        n = C.c_size_t(len(data))
        assert lib.salz_decode_safe(enc, len(enc) - 1, buf, C.byref(n)) == -1
    for v in [0, 7, 8, 71, 72, 583, 584, 4679, 4680, 37447, 37448, 65535, 299591, 299592, 2396744, 1 << 31]:
        a, b = C.c_uint64(0), C.c_uint64(0)
        ka = lib.encode_vnibble_le(C.c_uint32(v), C.byref(a))
        kb = ref.lib.encode_vnibble_le(v, C.byref(b))
        mask = (1 << (4 * kb)) - 1 if kb < 16 else (1 << 64) - 1
        assert ka == kb and (a.value & mask) == (b.value & mask), v
        assert lib.vnibble_size(C.c_uint32(v)) == ref.lib.vnibble_size(v)


def test_cli_decompress_roundtrip(tmp_path, ref):
    """Reference CLI compresses, our CLI decompresses (format compatibility)."""
    data = synth.english_like(3 * (1 << 15) + 4096, seed=2)
    p = tmp_path / "f.txt"
    p.write_bytes(data.tobytes())
    subprocess.run([os.path.join(ROOT, "oracle", "_ref", "salz_ref"), "-k", "-q", "-0", str(p)], check=True)
    p.unlink()
    subprocess.run([os.path.join(ROOT, "programs", "salz"), "-d", "-q", str(p) + ".salz"], check=True)
    assert p.read_bytes() == data.tobytes()
