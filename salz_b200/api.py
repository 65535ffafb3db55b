"""ctypes binding of include/salz_cuda.h (libsalz_cuda.so, built in-tree)."""
import ctypes as C
import os
import struct

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libsalz_cuda.so")
SALZ_MAGIC = 0x53414C5A  # ref programs/salzcli.c:28

u8p = C.POINTER(C.c_uint8)
i32p = C.POINTER(C.c_int32)
u32p = C.POINTER(C.c_uint32)
szp = C.POINTER(C.c_size_t)

EXPORTS = [
    "salz_cuda_device_count", "salz_cuda_ctx_create", "salz_cuda_ctx_destroy", "salz_cuda_last_error",
    "salz_cuda_encode_blocks", "salz_cuda_encode_stream", "salz_cuda_encode_stream_device",
    "salz_cuda_debug_stages", "salz_cuda_suffix_arrays", "salz_cuda_get_stats", "salz_cuda_reset_stats",
    "salz_cuda_set_sort_mode", "salz_cuda_set_parse_segments", "salz_cuda_host_alloc", "salz_cuda_host_free", "salz_cuda_set_stream",
    # multi-GPU (csrc/multi_gpu.cu)
    "salz_cuda_block_range", "salz_cuda_comm_unique_id", "salz_cuda_comm_create", "salz_cuda_comm_destroy",
    "salz_cuda_comm_rank", "salz_cuda_comm_world", "salz_cuda_comm_gathers", "salz_cuda_comm_allgather_sizes",
    "salz_cuda_multi_create", "salz_cuda_multi_destroy", "salz_cuda_multi_devices", "salz_cuda_multi_ctx",
    "salz_cuda_multi_last_error", "salz_cuda_multi_encode_stream",
    # no-allocation entry point (csrc/lcp.cu)
    "salz_cuda_lcp_noalloc",
]
NOALLOC_MAX_N = 1 << 18
UNIQUE_ID_BYTES = 128


class SalzCudaError(RuntimeError):
    pass


class Stats(C.Structure):
    _fields_ = [(n, C.c_uint64) for n in (
        "blocks", "input_bytes", "output_bytes", "launches", "sort_launches", "sort_elems", "sort_bytes",
        "doubling_rounds", "alg_bytes_sa", "alg_bytes_rest")] + [(n, C.c_double) for n in (
        "ms_sa", "ms_sort", "ms_lcp", "ms_ansv", "ms_parse", "ms_emit", "ms_total", "ms_h2d", "ms_d2h")]

    def asdict(self):
        return {n: getattr(self, n) for n, _ in self._fields_}


_lib = None


def load_library():
    """Loads libsalz_cuda.so; raises (never falls back) when it is missing."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise SalzCudaError(
            f"{LIB_PATH} is missing: build it with `python -m salz_b200.build` "
            "(there is no CPU fallback for the compression path)")
    L = C.CDLL(LIB_PATH)
    L.salz_cuda_device_count.restype = C.c_int
    L.salz_cuda_ctx_create.restype = C.c_void_p
    L.salz_cuda_ctx_create.argtypes = [C.c_int, C.c_size_t, C.c_size_t]
    L.salz_cuda_ctx_destroy.argtypes = [C.c_void_p]
    L.salz_cuda_last_error.restype = C.c_char_p
    L.salz_cuda_last_error.argtypes = [C.c_void_p]
    L.salz_cuda_encode_blocks.argtypes = [C.c_void_p, C.POINTER(C.c_void_p), szp, C.POINTER(C.c_void_p), szp,
                                          C.POINTER(C.c_int), C.c_size_t]
    L.salz_cuda_encode_stream.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_size_t, C.c_void_p, C.c_size_t,
                                          szp, u32p, szp]
    L.salz_cuda_encode_stream_device.argtypes = L.salz_cuda_encode_stream.argtypes
    L.salz_cuda_debug_stages.argtypes = [C.c_void_p, u8p, C.c_size_t, i32p, i32p, i32p, i32p, i32p, i32p, u8p, szp]
    L.salz_cuda_suffix_arrays.argtypes = [C.c_void_p, u8p, u32p, C.c_size_t, C.c_size_t, i32p]
    L.salz_cuda_get_stats.argtypes = [C.c_void_p, C.POINTER(Stats)]
    L.salz_cuda_reset_stats.argtypes = [C.c_void_p]
    L.salz_cuda_set_sort_mode.argtypes = [C.c_void_p, C.c_int]
    L.salz_cuda_set_parse_segments.argtypes = [C.c_void_p, C.c_int]
    L.salz_cuda_set_stream.argtypes = [C.c_void_p, C.c_void_p]
    L.salz_cuda_host_alloc.restype = C.c_void_p
    L.salz_cuda_host_alloc.argtypes = [C.c_size_t]
    L.salz_cuda_host_free.argtypes = [C.c_void_p]
    u64p = C.POINTER(C.c_uint64)
    L.salz_cuda_lcp_noalloc.argtypes = [C.c_int, C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    L.salz_cuda_block_range.restype = None
    L.salz_cuda_block_range.argtypes = [C.c_size_t, C.c_int, C.c_int, szp, szp]
    L.salz_cuda_comm_unique_id.argtypes = [u8p]
    L.salz_cuda_comm_create.restype = C.c_void_p
    L.salz_cuda_comm_create.argtypes = [C.c_int, u8p, C.c_int, C.c_int]
    L.salz_cuda_comm_destroy.argtypes = [C.c_void_p]
    L.salz_cuda_comm_rank.argtypes = [C.c_void_p]
    L.salz_cuda_comm_world.argtypes = [C.c_void_p]
    L.salz_cuda_comm_gathers.restype = C.c_uint64
    L.salz_cuda_comm_gathers.argtypes = [C.c_void_p]
    L.salz_cuda_comm_allgather_sizes.argtypes = [C.c_void_p, u32p, C.c_size_t, C.c_size_t, u32p, u64p]
    L.salz_cuda_multi_create.restype = C.c_void_p
    L.salz_cuda_multi_create.argtypes = [C.POINTER(C.c_int), C.c_int, C.c_size_t, C.c_size_t]
    L.salz_cuda_multi_destroy.argtypes = [C.c_void_p]
    L.salz_cuda_multi_devices.argtypes = [C.c_void_p]
    L.salz_cuda_multi_ctx.restype = C.c_void_p
    L.salz_cuda_multi_ctx.argtypes = [C.c_void_p, C.c_int]
    L.salz_cuda_multi_last_error.restype = C.c_char_p
    L.salz_cuda_multi_last_error.argtypes = [C.c_void_p]
    L.salz_cuda_multi_encode_stream.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_size_t, C.c_void_p, C.c_size_t,
                                                szp, u32p, szp]
    _lib = L
    return L


def block_range(nblocks, rank, world):
    """[lo, hi) of the blocks rank `rank` of `world` encodes (salz_cuda_block_range)."""
    lo, hi = C.c_size_t(0), C.c_size_t(0)
    load_library().salz_cuda_block_range(nblocks, rank, world, C.byref(lo), C.byref(hi))
    return lo.value, hi.value


def _preload_nccl():
    """The C library binds NCCL by soname at run time (dlopen "libnccl.so.2").  In a Python
    process that also imports torch, load the copy bundled with PyTorch first so that both
    end up on ONE NCCL (the process must not hold two versions under the same soname)."""
    global _nccl_preloaded
    if _nccl_preloaded:
        return
    _nccl_preloaded = True
    try:
        import importlib.util
        spec = importlib.util.find_spec("nvidia")
        for base in (spec.submodule_search_locations if spec else []):
            path = os.path.join(base, "nccl", "lib", "libnccl.so.2")
            if os.path.exists(path):
                C.CDLL(path, mode=C.RTLD_GLOBAL)
                return
    except (OSError, ImportError, AttributeError):
        pass


_nccl_preloaded = False


def unique_id():
    """128-byte NCCL id made by rank 0 and shipped to the other ranks by the launcher."""
    _preload_nccl()
    buf = (C.c_uint8 * UNIQUE_ID_BYTES)()
    if load_library().salz_cuda_comm_unique_id(buf) != 0:
        raise SalzCudaError("salz_cuda_comm_unique_id failed (NCCL missing?)")
    return bytes(buf)


class ShardComm:
    """One rank of the size all-gather (NCCL, bound by the C library): include/salz_cuda.h."""

    def __init__(self, device, uid, rank, world):
        self.lib = load_library()
        _preload_nccl()
        buf = (C.c_uint8 * UNIQUE_ID_BYTES).from_buffer_copy(uid)
        self.h = self.lib.salz_cuda_comm_create(device, buf, rank, world)
        if not self.h:
            raise SalzCudaError("salz_cuda_comm_create failed")
        self.rank, self.world = rank, world

    def allgather_sizes(self, local_sizes, nblocks_total):
        """-> (sizes of all blocks, record offsets [nblocks_total + 1]); collective."""
        mine = np.ascontiguousarray(local_sizes, dtype=np.uint32)
        sizes = np.zeros(nblocks_total, np.uint32)
        offs = np.zeros(nblocks_total + 1, np.uint64)
        rc = self.lib.salz_cuda_comm_allgather_sizes(self.h, mine.ctypes.data_as(u32p), len(mine), nblocks_total,
                                                     sizes.ctypes.data_as(u32p),
                                                     offs.ctypes.data_as(C.POINTER(C.c_uint64)))
        if rc != 0:
            raise SalzCudaError("salz_cuda_comm_allgather_sizes failed")
        return sizes, offs

    def gathers(self):
        return int(self.lib.salz_cuda_comm_gathers(self.h))

    def close(self):
        if getattr(self, "h", None):
            self.lib.salz_cuda_comm_destroy(self.h)
            self.h = None


class SalzCudaMulti:
    """One process, several GPUs: salz_cuda_multi_* (a thread, a context and an NCCL rank per device)."""

    def __init__(self, ndev=0, max_block_len=1 << 20, max_batch_blocks=64, devices=None):
        self.lib = load_library()
        _preload_nccl()
        arr = (C.c_int * len(devices))(*devices) if devices else None
        self.h = self.lib.salz_cuda_multi_create(arr, len(devices) if devices else ndev, max_block_len, max_batch_blocks)
        if not self.h:
            raise SalzCudaError("salz_cuda_multi_create failed")
        self.ndev = self.lib.salz_cuda_multi_devices(self.h)

    def encode_stream(self, data, block_len):
        src = _u8(data)
        cap = stream_len_max(len(src), block_len)
        dst = np.empty(cap, np.uint8)
        nb = (len(src) + block_len - 1) // block_len
        sizes = np.zeros(nb, np.uint32)
        dl, nbo = C.c_size_t(0), C.c_size_t(0)
        rc = self.lib.salz_cuda_multi_encode_stream(self.h, src.ctypes.data, len(src), block_len, dst.ctypes.data, cap,
                                                    C.byref(dl), sizes.ctypes.data_as(u32p), C.byref(nbo))
        if rc != 0:
            raise SalzCudaError(self.lib.salz_cuda_multi_last_error(self.h).decode())
        return dst[: dl.value].tobytes(), sizes

    def encode_stream_ptr(self, src_ptr, src_len, block_len, dst_ptr, dst_cap, sizes=None):
        dl, nbo = C.c_size_t(0), C.c_size_t(0)
        sp = sizes.ctypes.data_as(u32p) if sizes is not None else None
        rc = self.lib.salz_cuda_multi_encode_stream(self.h, src_ptr, src_len, block_len, dst_ptr, dst_cap,
                                                    C.byref(dl), sp, C.byref(nbo))
        if rc != 0:
            raise SalzCudaError(self.lib.salz_cuda_multi_last_error(self.h).decode())
        return dl.value

    def close(self):
        if getattr(self, "h", None):
            self.lib.salz_cuda_multi_destroy(self.h)
            self.h = None


def encoded_len_max(n):
    """Worst-case encoded block length (ref lib/salz.h:25-28)."""
    return 4 + n + ((n + 63) // 64) * 64 // 8


def stream_len_max(src_len, block_len):
    nb = (src_len + block_len - 1) // block_len
    return src_len + nb * (8 + block_len // 8 + 8) + 64


def frame_file(block_len, stream_bytes):
    """8-byte file header (ref programs/salzcli.c:115-121) + framed block stream."""
    return struct.pack("<II", SALZ_MAGIC, block_len) + stream_bytes


def _u8(data):
    if isinstance(data, (bytes, bytearray, memoryview)):
        return np.frombuffer(bytes(data), dtype=np.uint8)
    return np.ascontiguousarray(data, dtype=np.uint8)


class SalzCuda:
    """One encoder context on one GPU (mirrors salz_encode_safe / the CLI block loop)."""

    def __init__(self, device=0, max_block_len=1 << 20, max_batch_blocks=64):
        self.lib = load_library()
        if self.lib.salz_cuda_device_count() <= 0:
            raise SalzCudaError("no CUDA device: the salz_b200 compression path has no CPU fallback")
        self.ctx = self.lib.salz_cuda_ctx_create(device, max_block_len, max_batch_blocks)
        if not self.ctx:
            raise SalzCudaError("salz_cuda_ctx_create failed")
        self.max_block_len = max_block_len
        self.max_batch_blocks = max_batch_blocks

    def close(self):
        if getattr(self, "ctx", None):
            self.lib.salz_cuda_ctx_destroy(self.ctx)
            self.ctx = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _err(self):
        return self.lib.salz_cuda_last_error(self.ctx).decode()

    def set_sort_mode(self, mode):
        self.lib.salz_cuda_set_sort_mode(self.ctx, mode)

    def set_parse_segments(self, k):
        """Segments per block of the parse: 0 = automatic, k >= 1 = at most k, k <= -1 = |k| with the shortened first pass."""
        self.lib.salz_cuda_set_parse_segments(self.ctx, k)

    def set_stream(self, cuda_stream):
        """Run on the caller's stream (an int cudaStream_t, e.g. torch.cuda.Stream().cuda_stream)."""
        self.lib.salz_cuda_set_stream(self.ctx, cuda_stream)

    def stats(self):
        s = Stats()
        self.lib.salz_cuda_get_stats(self.ctx, C.byref(s))
        return s.asdict()

    def reset_stats(self):
        self.lib.salz_cuda_reset_stats(self.ctx)

    # -- salz_encode_safe semantics, batched ---------------------------------
    def encode_blocks(self, blocks, caps=None):
        """blocks: list of bytes-like.  Returns list of bytes (or None where rc == -1)."""
        n = len(blocks)
        srcs = [_u8(b) for b in blocks]
        caps = [encoded_len_max(len(s)) for s in srcs] if caps is None else list(caps)
        dsts = [np.zeros(max(c, 1), np.uint8) for c in caps]
        sp = (C.c_void_p * n)(*[s.ctypes.data if len(s) else None for s in srcs])
        dp = (C.c_void_p * n)(*[d.ctypes.data for d in dsts])
        sl = (C.c_size_t * n)(*[len(s) for s in srcs])
        dl = (C.c_size_t * n)(*caps)
        rcs = (C.c_int * n)()
        if self.lib.salz_cuda_encode_blocks(self.ctx, sp, sl, dp, dl, rcs, n) != 0:
            raise SalzCudaError(self._err())
        return [dsts[i][: dl[i]].tobytes() if rcs[i] == 0 else None for i in range(n)]

    def encode(self, data, cap=None):
        return self.encode_blocks([data], None if cap is None else [cap])[0]

    # -- CLI block loop semantics ---------------------------------------------
    def encode_stream(self, data, block_len):
        """Host buffer -> framed block stream (u32 len + block, per block) and sizes."""
        src = _u8(data)
        cap = stream_len_max(len(src), block_len)
        dst = np.empty(cap, np.uint8)
        nb = (len(src) + block_len - 1) // block_len
        sizes = np.zeros(nb, np.uint32)
        dl = C.c_size_t(0)
        nbo = C.c_size_t(0)
        rc = self.lib.salz_cuda_encode_stream(self.ctx, src.ctypes.data, len(src), block_len, dst.ctypes.data, cap,
                                              C.byref(dl), sizes.ctypes.data_as(u32p), C.byref(nbo))
        if rc != 0:
            raise SalzCudaError(self._err())
        return dst[: dl.value].tobytes(), sizes

    def encode_stream_ptr(self, src_ptr, src_len, block_len, dst_ptr, dst_cap, sizes=None, device=False):
        """Raw-pointer variant (pinned host or device memory); returns bytes produced."""
        dl = C.c_size_t(0)
        nbo = C.c_size_t(0)
        fn = self.lib.salz_cuda_encode_stream_device if device else self.lib.salz_cuda_encode_stream
        sp = sizes.ctypes.data_as(u32p) if sizes is not None else None
        rc = fn(self.ctx, src_ptr, src_len, block_len, dst_ptr, dst_cap, C.byref(dl), sp, C.byref(nbo))
        if rc != 0:
            raise SalzCudaError(self._err())
        return dl.value

    def compress_file_bytes(self, data, level=5):
        """What `salz -<level>` writes for `data` (ref programs/salzcli.c:102-185)."""
        block_len = 1 << (15 + level)
        stream, _ = self.encode_stream(data, block_len)
        return frame_file(block_len, stream)

    # -- test hooks ---------------------------------------------------------------
    def stages(self, data):
        src = _u8(data)
        N = len(src)
        n = N - 8
        sa = np.zeros(n, np.int32); plcp = np.zeros(n, np.int32); lcp = np.zeros(n, np.int32)
        cand = np.zeros((n, 4), np.int32); parse = np.zeros((n, 3), np.int32); pn = np.zeros((n, 2), np.int32)
        cap = encoded_len_max(N)
        dst = np.zeros(cap, np.uint8); dl = C.c_size_t(cap)
        rc = self.lib.salz_cuda_debug_stages(
            self.ctx, src.ctypes.data_as(u8p), N, sa.ctypes.data_as(i32p), plcp.ctypes.data_as(i32p),
            lcp.ctypes.data_as(i32p), pn.ctypes.data_as(i32p), cand.ctypes.data_as(i32p),
            parse.ctypes.data_as(i32p), dst.ctypes.data_as(u8p), C.byref(dl))
        if rc != 0:
            raise SalzCudaError(self._err())
        return dict(sa=sa, plcp=plcp, lcp=lcp, psvnsv=pn, cand=cand, parse=parse, out=dst[: dl.value].tobytes())

    def suffix_arrays(self, texts):
        """texts: list of bytes-like (each <= max_block_len - 8).  Returns list of int32 arrays."""
        arrs = [_u8(t) for t in texts]
        stride = max(len(a) for a in arrs)
        stride = (stride + 15) // 16 * 16
        buf = np.zeros(len(arrs) * stride, np.uint8)
        for i, a in enumerate(arrs):
            buf[i * stride: i * stride + len(a)] = a
        lens = np.array([len(a) for a in arrs], np.uint32)
        out = np.zeros(len(arrs) * stride, np.int32)
        rc = self.lib.salz_cuda_suffix_arrays(self.ctx, buf.ctypes.data_as(u8p), lens.ctypes.data_as(u32p), stride,
                                              len(arrs), out.ctypes.data_as(i32p))
        if rc != 0:
            raise SalzCudaError(self._err())
        return [out[i * stride: i * stride + len(a)].copy() for i, a in enumerate(arrs)]
