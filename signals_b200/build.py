"""Build libsignals_b200.so (sm_100a only) in-tree with nvcc.

    python -m signals_b200.build [--force] [--verbose]

The shared object sits next to this file so it travels to the GPU box with the repo snapshot.
"""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
SO = os.path.join(HERE, "libsignals_b200.so")
SOURCES = ["hmm_kernels.cu", "split_kernels.cu", "table_kernels.cu", "capi.cu", "noalloc.cu"]
HEADERS = ["sgb_math.h", "sgb_tables.h", "hmm_core.cuh", "kernels.h", "sgb_async.cuh", os.path.join("..", "..", "include", "signals_b200.h")]

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-std=c++17", "-lineinfo",
    "--fmad=false",                  # never contract a*b+c; every fused op in the math is an explicit fma
    "-Xcompiler", "-fPIC,-O2,-ffp-contract=off,-fvisibility=hidden",
    "--shared", "-cudart", "static",
]


def needs_build():
    if not os.path.exists(SO):
        return True
    t = os.path.getmtime(SO)
    files = [os.path.join(CSRC, f) for f in SOURCES + HEADERS] + [os.path.abspath(__file__)]
    return any(os.path.exists(f) and os.path.getmtime(f) > t for f in files)


def build(force=False, verbose=False, variant=None, defines=()):
    """variant/defines: experimental builds (libsignals_b200_<variant>.so, selected with SGB_LIB=<path>)"""
    so = SO if not variant else os.path.join(HERE, "libsignals_b200_%s.so" % variant)
    if not variant and not force and not needs_build():
        return SO
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    cmd = [nvcc] + NVCC_FLAGS + ["-D" + d for d in defines] + (["-Xptxas", "-v"] if verbose else []) + \
          [os.path.join(CSRC, s) for s in SOURCES] + ["-o", so]
    if verbose:
        print(" ".join(cmd))
    subprocess.check_call(cmd)
    return so


if __name__ == "__main__":
    variant = None
    defines = []
    for i, a in enumerate(sys.argv):
        if a == "--variant":
            variant = sys.argv[i + 1]
        if a == "--define":
            defines.append(sys.argv[i + 1])
    print(build(force="--force" in sys.argv, verbose="--verbose" in sys.argv, variant=variant, defines=defines))
