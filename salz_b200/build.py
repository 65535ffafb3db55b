"""Builds salz_b200/libsalz_cuda.so (hand-written sm_100a kernels + C ABI) in-tree."""
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libsalz_cuda.so")
SOURCES = ["sa.cu", "lcp.cu", "ansv.cu", "parse.cu", "emit.cu", "api.cu", "multi_gpu.cu"]
NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
    "-Xcompiler", "-fPIC", "-shared", "-cudart", "static", "-ldl",
]


def needs_build():
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    deps = [os.path.join(CSRC, f) for f in os.listdir(CSRC)] + [
        os.path.join(HERE, "..", "include", "salz_cuda.h")]
    return any(os.path.getmtime(d) > t for d in deps)


def build(force=False, verbose=False):
    if not force and not needs_build():
        return LIB
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    defs = ["-D%s=%s" % (k, os.environ[k]) for k in ("SALZ_TT", "SALZ_ITEMS", "SALZ_LBW", "SALZ_LT", "SALZ_LITEMS", "SALZ_LCTAS", "SALZ_PC", "SALZ_HELPERS", "SALZ_PARSE_CTAS", "SALZ_NCH", "SALZ_PER", "SALZ_PLCP_CTAS", "SALZ_SCTAS", "SALZ_PLCP_ITEMS", "SALZ_WARM", "SALZ_TCTAS", "SALZ_TSLOTS", "SALZ_L2_HINTS", "SALZ_SCATTER_PROF", "SALZ_LOCAL_FAST", "SALZ_AL") if k in os.environ]
    cmd = [nvcc] + NVCC_FLAGS + defs + (["-Xptxas", "-v"] if verbose else []) + \
          ["-o", LIB] + [os.path.join(CSRC, s) for s in SOURCES]
    subprocess.run(cmd, check=True)
    return LIB


if __name__ == "__main__":
    import sys
    print(build(force=True, verbose="-v" in sys.argv))
