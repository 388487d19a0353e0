"""ctypes binding of the CPU oracle (oracle/_build/liboracle.so).

TEST INFRASTRUCTURE, NOT PRODUCT CODE: only tests/, __graft_entry__.smoke()
and bench.py's cpu_baseline / --impl reference legs import this module.  The
product (rocm-k8s-device-plugin_b200/) never does.

`RefPolicy` mirrors the reference's allocator.Policy (allocator.go:27-30):
Init(devs, topoDir) / Allocate(available, required, size).
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from dataclasses import dataclass

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "_build", "liboracle.so")
ID_LEN = 64


class _Device(C.Structure):
    _fields_ = [("id", C.c_char * ID_LEN), ("node_id", C.c_int), ("numa_node", C.c_int),
                ("dev_id", C.c_char * ID_LEN)]


class _Result(C.Structure):
    _fields_ = [("score", C.c_int), ("n_candidates", C.c_int), ("searched", C.c_int)]


class _ExhResult(C.Structure):
    _fields_ = [("score", C.c_int), ("m", C.c_int), ("scored", C.c_ulonglong),
                ("total", C.c_ulonglong), ("slow", C.c_ulonglong), ("n_free", C.c_int),
                ("k_free", C.c_int), ("n_groups", C.c_int), ("threads", C.c_int),
                ("key_hi", C.c_ulonglong), ("key_lo", C.c_ulonglong)]


@dataclass
class Device:
    """allocator.Device (device.go:56-65), the four fields the policy reads."""
    Id: str
    NodeId: int
    NumaNode: int
    DevId: str


class OracleError(Exception):
    def __init__(self, code: int, msg: str):
        super().__init__(msg)
        self.code = code


def build(force: bool = False) -> str:
    """Compile the oracle with oracle/Makefile (gcc only)."""
    srcs = [os.path.join(_HERE, f) for f in ("ref_policy.c", "exhaustive.c", "ref_policy.h", "ref_internal.h")]
    stale = force or not os.path.exists(_LIB_PATH) or any(
        os.path.getmtime(s) > os.path.getmtime(_LIB_PATH) for s in srcs)
    if stale:
        subprocess.run(["make", "-C", _HERE, "-s"], check=True)
    return _LIB_PATH


_lib = None


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(_LIB_PATH)
        L.ref_policy_new.restype = C.c_void_p
        L.ref_policy_free.argtypes = [C.c_void_p]
        L.ref_policy_init.argtypes = [C.c_void_p, C.POINTER(_Device), C.c_int, C.c_char_p]
        L.ref_policy_init_pairs.argtypes = [C.c_void_p, C.POINTER(_Device), C.c_int, C.POINTER(C.c_int),
                                            C.POINTER(C.c_int), C.POINTER(C.c_int), C.c_int]
        L.ref_policy_allocate.argtypes = [C.c_void_p, C.POINTER(C.c_char_p), C.c_int, C.POINTER(C.c_char_p),
                                          C.c_int, C.c_int, C.c_char_p, C.POINTER(_Result)]
        L.ref_exhaustive_allocate.argtypes = [C.c_void_p, C.POINTER(C.c_char_p), C.c_int,
                                              C.POINTER(C.c_char_p), C.c_int, C.c_int, C.c_char_p,
                                              C.POINTER(_ExhResult), C.c_int, C.c_ulonglong, C.c_ulonglong]
        L.ref_exhaustive_global_optimum.argtypes = [C.c_void_p, C.POINTER(C.c_char_p), C.c_int,
                                                    C.POINTER(C.c_char_p), C.c_int, C.c_int, C.c_char_p,
                                                    C.POINTER(_ExhResult), C.c_int]
        for f in ("ref_policy_weight_rows", "ref_policy_group_count", "ref_policy_pair_count",
                  "ref_policy_max_node"):
            getattr(L, f).argtypes = [C.c_void_p]
        L.ref_policy_weight.argtypes = [C.c_void_p, C.c_int, C.c_int]
        L.ref_policy_has_pair.argtypes = [C.c_void_p, C.c_int, C.c_int]
        L.ref_strerror.restype = C.c_char_p
        L.ref_strerror.argtypes = [C.c_int]
        _lib = L
    return _lib


def _strs(xs):
    arr = (C.c_char_p * max(len(xs), 1))()
    for i, s in enumerate(xs):
        arr[i] = s.encode()
    return arr


def _devs(devs):
    arr = (_Device * max(len(devs), 1))()
    for i, d in enumerate(devs):
        arr[i].id = d.Id.encode()
        arr[i].node_id = d.NodeId
        arr[i].numa_node = d.NumaNode
        arr[i].dev_id = d.DevId.encode()
    return arr


class RefPolicy:
    """The reference BestEffortPolicy, restated (besteffort_policy.go:45-151)."""

    def __init__(self):
        self._p = lib().ref_policy_new()
        self.last = None

    def __del__(self):
        if getattr(self, "_p", None):
            lib().ref_policy_free(self._p)
            self._p = None

    def _check(self, rc):
        if rc < 0:
            raise OracleError(rc, lib().ref_strerror(rc).decode())
        return rc

    def Init(self, devs, topo_dir: str):
        self.devices = list(devs)
        self._check(lib().ref_policy_init(self._p, _devs(devs), len(devs), topo_dir.encode()))

    def InitPairs(self, devs, pairs):
        """Init from (from, to, weight) triples instead of a sysfs tree."""
        self.devices = list(devs)
        n = len(pairs)
        f = (C.c_int * max(n, 1))(*[p[0] for p in pairs])
        t = (C.c_int * max(n, 1))(*[p[1] for p in pairs])
        w = (C.c_int * max(n, 1))(*[p[2] for p in pairs])
        self._check(lib().ref_policy_init_pairs(self._p, _devs(devs), len(devs), f, t, w, n))

    def Allocate(self, available, required, size):
        required = required or []
        out = C.create_string_buffer(ID_LEN * max(size, 1))
        res = _Result()
        rc = self._check(lib().ref_policy_allocate(self._p, _strs(available), len(available), _strs(required),
                                                   len(required), size, out, C.byref(res)))
        self.last = {"score": res.score, "n_candidates": res.n_candidates, "searched": bool(res.searched)}
        return [out.raw[i * ID_LEN:(i + 1) * ID_LEN].split(b"\0", 1)[0].decode() for i in range(rc)]

    def AllocateExhaustive(self, available, required, size, threads=0, rank_lo=0, rank_hi=0):
        """Same request answered by scoring every C(n,k') subset (exhaustive.c)."""
        required = required or []
        out = C.create_string_buffer(ID_LEN * max(size, 1))
        res = _ExhResult()
        rc = self._check(lib().ref_exhaustive_allocate(self._p, _strs(available), len(available),
                                                       _strs(required), len(required), size, out, C.byref(res),
                                                       threads, rank_lo, rank_hi))
        self.last = {f: getattr(res, f) for f, _ in _ExhResult._fields_}
        return [out.raw[i * ID_LEN:(i + 1) * ID_LEN].split(b"\0", 1)[0].decode() for i in range(rc)]

    def GlobalOptimum(self, available, required, size, threads=0):
        """Minimum of Score over ALL subsets (no Family test); ties: first in canonical order."""
        required = required or []
        out = C.create_string_buffer(ID_LEN * max(size, 1))
        res = _ExhResult()
        rc = self._check(lib().ref_exhaustive_global_optimum(self._p, _strs(available), len(available),
                                                             _strs(required), len(required), size, out,
                                                             C.byref(res), threads))
        self.last = {f: getattr(res, f) for f, _ in _ExhResult._fields_}
        return [out.raw[i * ID_LEN:(i + 1) * ID_LEN].split(b"\0", 1)[0].decode() for i in range(rc)]

    # introspection for the known-answer tests
    def weight_rows(self):
        return lib().ref_policy_weight_rows(self._p)

    def group_count(self):
        return lib().ref_policy_group_count(self._p)

    def pairs(self):
        """All (from, to, weight) entries of p2pWeights, sorted."""
        L = lib()
        m = L.ref_policy_max_node(self._p)
        return [(f, t, L.ref_policy_weight(self._p, f, t)) for f in range(m + 1) for t in range(m + 1)
                if L.ref_policy_has_pair(self._p, f, t)]


def get_test_devices(dev_count, partitions_per_dev, numa_count, start_node, end_node):
    """testInfo.getTestDevices (device_test.go:43-67)."""
    res = []
    node = start_node
    for i in range(dev_count):
        numa = dev_count // numa_count
        for j in range(partitions_per_dev):
            ident = f"amdgpu_xcp_{i * 8 + j}" if j else f"test{i + 1}"
            if node > end_node:
                break
            res.append(Device(ident, node, i // numa, str(i)))
            node += 1
    return res
