"""salz_b200 - B200-native implementation of the salz compression hot path.

Host-side mirror of the reference's codec interface (lib/salz.h) on top of the
C ABI in include/salz_cuda.h.  The compute path is hand-written sm_100a CUDA
(salz_b200/csrc); there is no CPU fallback: without the built extension or
without a CUDA device every call raises.
"""
from .api import (  # noqa: F401
    SalzCuda,
    SalzCudaMulti,
    ShardComm,
    SalzCudaError,
    encoded_len_max,
    load_library,
    SALZ_MAGIC,
    frame_file,
)

__all__ = ["SalzCuda", "SalzCudaMulti", "ShardComm", "SalzCudaError", "encoded_len_max", "load_library", "SALZ_MAGIC", "frame_file"]
