"""Build libtpalloc.so (the C ABI + sm_100a kernels) in-tree with nvcc.

No torch, no JIT cache: the .so sits next to this file so that it travels to the GPU box
with the repo snapshot.  `python build.py` or build() from __graft_entry__.
"""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libtpalloc.so")
SOURCES = ["csrc/tp_abi.cu", "csrc/tp_kernels.cu", "host/besteffort_policy.cpp"]
HEADERS = ["csrc/tp_common.cuh", "csrc/tp_kernels.cuh", "csrc/tp_split.cuh", "host/besteffort_policy.h"]
NVCC_FLAGS = ["-O3", "-std=c++17", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo",
              "-Xcompiler", "-fPIC,-O3,-Wall", "-shared", "-cudart", "static"]


def stale():
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    deps = [os.path.join(HERE, f) for f in SOURCES + HEADERS] + [os.path.join(ROOT, "include", "tp_allocator.h")]
    return any(os.path.getmtime(d) > t for d in deps if os.path.exists(d))


def build(force=False, verbose=False):
    if not (force or stale()):
        return LIB
    cmd = ["nvcc"] + NVCC_FLAGS + ["-I", os.path.join(ROOT, "include"), "-I", CSRC, "-o", LIB] \
        + [os.path.join(HERE, s) for s in SOURCES]
    if verbose:
        cmd += ["-Xptxas", "-v"]
    subprocess.run(cmd, check=True, cwd=CSRC)
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
