"""Build libtpalloc.so (the C ABI + sm_100a kernels) in-tree with nvcc.

No torch, no JIT cache: the .so sits next to this file so that it travels to the GPU box
with the repo snapshot.  Translation units are compiled in parallel (the SPLIT kernel is one
unit per lane-vector width), then linked.  `python build.py [--force] [-v]`, or build() from
__graft_entry__.
"""
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
CSRC = os.path.join(HERE, "csrc")
# A/B builds (tuning aid): TP_BUILD_TAG=x TP_BUILD_DEFS="-DTP_TRI_GT=12" python build.py -> libtpalloc_x.so, to be
# loaded with TP_ALLOC_LIB=<path>; the default build has no tag.
TAG = os.environ.get("TP_BUILD_TAG", "")
OBJ = os.path.join("/tmp", "tpalloc_build_" + TAG) if TAG else os.path.join(HERE, "build")  # (tagged objects stay out of the tree)
LIB = os.path.join(HERE, "libtpalloc" + ("_" + TAG if TAG else "") + ".so")
SPLIT_UNITS = [f"csrc/tp_split_m{m}_d{d}.cu" for m in (128, 96, 80, 64) for d in (4, 3, 2, 1, 0)]  # slowest first
TRI_UNITS = [f"csrc/tp_tri_m{m}_d{d}.cu" for m in (80, 64) for d in (3, 2, 1)]
SOURCES = TRI_UNITS + SPLIT_UNITS + ["csrc/tp_abi.cu", "csrc/tp_kernels.cu", "host/besteffort_policy.cpp"]
HEADERS = ["csrc/tp_common.cuh", "csrc/tp_kernels.cuh", "csrc/tp_device.cuh", "csrc/tp_split.cuh", "csrc/tp_tri.cuh",
           "host/besteffort_policy.h"]
ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
CFLAGS = ["-O3", "-std=c++17", "-lineinfo", "-Xcompiler", "-fPIC,-O3,-Wall"] + ARCH + os.environ.get("TP_BUILD_DEFS", "").split()
INCLUDES = ["-I", os.path.join(ROOT, "include"), "-I", CSRC]


def _deps():
    return [os.path.join(HERE, f) for f in HEADERS] + [os.path.join(ROOT, "include", "tp_allocator.h")]


def _obj(src):
    return os.path.join(OBJ, os.path.basename(src).rsplit(".", 1)[0] + ".o")


def _stale(target, sources):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(s) > t for s in sources if os.path.exists(s))


def stale():
    return _stale(LIB, [os.path.join(HERE, s) for s in SOURCES] + _deps())


def _compile(src, force, verbose):
    path, obj = os.path.join(HERE, src), _obj(src)
    if not (force or _stale(obj, [path] + _deps())):
        return ""
    cmd = ["nvcc"] + CFLAGS + INCLUDES + ["-c", path, "-o", obj] + (["-Xptxas", "-v"] if verbose else [])
    r = subprocess.run(cmd, capture_output=True, text=True, cwd=CSRC)
    if r.returncode != 0:
        raise RuntimeError(f"nvcc failed on {src}:\n{r.stdout}\n{r.stderr}")
    return r.stdout + r.stderr


def build(force=False, verbose=False):
    if not (force or stale()):
        return LIB
    os.makedirs(OBJ, exist_ok=True)
    with ThreadPoolExecutor(max_workers=min(len(SOURCES), os.cpu_count() or 4)) as ex:
        logs = list(ex.map(lambda s: _compile(s, force, verbose), SOURCES))
    if verbose:
        print("\n".join(logs))
    cmd = (["nvcc", "-shared", "-cudart", "static"] + ARCH + os.environ.get("TP_BUILD_LDFLAGS", "").split() +
           ["-o", LIB] + [_obj(s) for s in SOURCES])
    subprocess.run(cmd, check=True, cwd=CSRC)
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
