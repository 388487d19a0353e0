"""B200 topology-aware preferred-allocation policy: ctypes view of libtpalloc.so.

The product is the shared library (csrc/: C ABI + sm_100a kernels, host/: C++ mirror of the
reference's allocator.Policy).  This module only marshals arguments so that tests and
bench.py read like the reference's Go tests (besteffort_policy_test.go):

    pol = BestEffortPolicy()
    pol.Init(devices, topo_dir)            # allocator.Policy.Init      (allocator.go:28)
    ids = pol.Allocate(avail, req, size)   # allocator.Policy.Allocate  (allocator.go:29)

There is no Python or CPU implementation of the search behind it: loading fails loudly when
the library is missing, and every searched request runs on the GPU.
"""
from __future__ import annotations

import ctypes as C
import os
from dataclasses import dataclass

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("TP_ALLOC_LIB") or os.path.join(_HERE, "libtpalloc.so")  # (override: A/B runs of two builds)
ID_LEN = 64

TP_MAX_DEVICES = 128
TP_MAX_SUBSET = 16
STATUS = {
    "TP_OK": 0, "TP_ERR_INVALID_SIZE": -1, "TP_ERR_INVALID_AVAILABLE": -2, "TP_ERR_INVALID_REQUIRED": -3,
    "TP_ERR_INVALID_REQ_AVAILABLE": -4, "TP_ERR_INVALID_INIT": -5, "TP_ERR_NO_CANDIDATE": -6,
    "TP_ERR_INIT_FAILED": -7, "TP_ERR_BAD_INDEX": -20, "TP_ERR_BAD_ARGUMENT": -21, "TP_ERR_TOO_LARGE": -22,
    "TP_ERR_STATIC_SCRATCH": -23, "TP_ERR_CUDA": -30,
}
KERNEL_AUTO, KERNEL_LEX, KERNEL_SPLIT, KERNEL_FAMILY, KERNEL_TRI = 0, 1, 2, 3, 4


def kernel_name(kernel: int) -> str:
    return {1: "lex", 2: "split", 3: "family", 4: "tri"}.get(kernel, str(kernel))


class TpKey(C.Structure):
    _fields_ = [("hi", C.c_uint64), ("lo", C.c_uint64)]

    def tuple(self):
        return (self.hi, self.lo)


class TpResult(C.Structure):
    _fields_ = [("score", C.c_int32), ("groups_touched", C.c_int32), ("searched", C.c_int32),
                ("kernel", C.c_int32), ("subsets_scored", C.c_uint64), ("subsets_total", C.c_uint64),
                ("kernel_ms", C.c_float), ("gpu_launches", C.c_int32), ("key", TpKey)]

    def dict(self):
        d = {f: getattr(self, f) for f, _ in self._fields_ if f != "key"}
        d["key"] = (self.key.hi, self.key.lo)
        return d


class TpOptions(C.Structure):
    _fields_ = [("cuda_device", C.c_int32), ("kernel", C.c_int32), ("exhaustive_limit", C.c_uint64),
                ("max_inflight", C.c_int32), ("n_extra_devices", C.c_int32), ("extra_devices", C.c_int32 * 16),
                ("multi_gpu_min_subsets", C.c_uint64)]


class TpTopology(C.Structure):
    _fields_ = [("n_devices", C.c_int32), ("n_groups", C.c_int32), ("node_id", C.POINTER(C.c_int32)),
                ("group", C.POINTER(C.c_int32)), ("parent_rank", C.POINTER(C.c_int32)),
                ("weight", C.POINTER(C.c_uint8))]


class TpRequest(C.Structure):
    _fields_ = [("available", C.POINTER(C.c_int32)), ("n_available", C.c_int32),
                ("required", C.POINTER(C.c_int32)), ("n_required", C.c_int32), ("size", C.c_int32)]


class _HDevice(C.Structure):
    _fields_ = [("id", C.c_char * ID_LEN), ("node_id", C.c_int32), ("numa_node", C.c_int32),
                ("dev_id", C.c_char * ID_LEN)]


@dataclass
class Device:
    """allocator.Device (device.go:56-65)."""
    Id: str
    NodeId: int
    NumaNode: int
    DevId: str


class AllocatorError(Exception):
    """A Go `error` from the policy; str() is the reference's message."""

    def __init__(self, status: int, msg: str):
        super().__init__(msg)
        self.status = status


_lib = None
I32P = C.POINTER(C.c_int32)


def lib() -> C.CDLL:
    """Load libtpalloc.so; raises if it has not been built (no fallback of any kind)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(f"{LIB_PATH} is missing: run `python rocm-k8s-device-plugin_b200/build.py` "
                           "(nvcc, sm_100a). There is no CPU fallback.")
    L = C.CDLL(LIB_PATH)
    L.tp_strerror.restype = C.c_char_p
    L.tp_strerror.argtypes = [C.c_int]
    L.tp_last_cuda_error.restype = C.c_char_p
    L.tp_last_cuda_error.argtypes = [C.c_void_p]
    L.tp_create.argtypes = [C.POINTER(TpTopology), C.POINTER(TpOptions), C.POINTER(C.c_void_p)]
    L.tp_destroy.argtypes = [C.c_void_p]
    L.tp_destroy.restype = None
    req_args = [C.c_void_p, I32P, C.c_int32, I32P, C.c_int32, C.c_int32]
    L.tp_allocate.argtypes = req_args + [I32P, C.POINTER(TpResult)]
    L.tp_global_optimum.argtypes = req_args + [I32P, C.POINTER(TpResult)]
    L.tp_allocate_batch.argtypes = [C.c_void_p, C.c_int32, C.POINTER(TpRequest), I32P, C.c_int32, I32P,
                                    C.POINTER(TpResult)]
    L.tp_search_shard.argtypes = req_args + [C.c_int32, C.c_int32, C.POINTER(TpKey), C.POINTER(TpResult)]
    L.tp_decode.argtypes = req_args + [C.POINTER(TpKey), I32P]
    L.tp_shard_subsets.argtypes = req_args + [C.c_int32, C.c_int32, C.POINTER(C.c_uint64), I32P]
    L.tp_search_resident.argtypes = req_args + [C.c_int32, C.c_int32, C.c_int32, C.POINTER(C.c_float),
                                                C.POINTER(TpKey), C.POINTER(TpResult)]
    L.tp_build_weights.argtypes = [C.c_int32, I32P, I32P, I32P, C.POINTER(C.c_uint8), C.c_int32]
    L.tp_prepare.argtypes = req_args + [C.c_int32, C.c_int32, C.POINTER(C.c_void_p)]
    L.tp_launch.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    L.tp_prepared_subsets.restype = C.c_uint64
    L.tp_prepared_subsets.argtypes = [C.c_void_p]
    L.tp_prepared_free.argtypes = [C.c_void_p, C.c_void_p]
    L.tp_prepared_free.restype = None
    L.tp_measure_int_peak.argtypes = [C.c_int32, C.c_int32, C.POINTER(C.c_double)]
    L.tp_exchange_create.argtypes = [C.c_void_p, C.c_int32, C.c_int32, C.POINTER(C.c_void_p), C.c_char_p]
    L.tp_exchange_connect.argtypes = [C.c_void_p, C.c_char_p]
    L.tp_launch_exchange.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    L.tp_exchange_destroy.argtypes = [C.c_void_p]
    L.tp_exchange_destroy.restype = None
    L.tp_set_tuning.argtypes = [C.c_void_p, C.c_int32, C.c_uint64, C.c_uint64]
    L.tp_set_tuning.restype = None
    L.tp_set_batch_tuning.argtypes = [C.c_void_p, C.c_int32, C.c_int32, C.c_int32]
    L.tp_set_batch_tuning.restype = None
    L.tp_allocate_static.argtypes = [C.POINTER(TpTopology), C.c_void_p, I32P, C.c_int32, I32P, C.c_int32,
                                     C.c_int32, C.c_void_p, C.c_void_p, I32P, C.POINTER(TpResult)]
    L.tph_new.restype = C.c_void_p
    L.tph_new.argtypes = [C.POINTER(TpOptions)]
    L.tph_free.argtypes = [C.c_void_p]
    L.tph_free.restype = None
    L.tph_init.argtypes = [C.c_void_p, C.POINTER(_HDevice), C.c_int, C.c_char_p]
    L.tph_init_pairs.argtypes = [C.c_void_p, C.POINTER(_HDevice), C.c_int, I32P, I32P, I32P, C.c_int]
    strs = C.POINTER(C.c_char_p)
    L.tph_allocate.argtypes = [C.c_void_p, strs, C.c_int, strs, C.c_int, C.c_int, C.c_char_p, C.POINTER(TpResult)]
    L.tph_search_shard.argtypes = [C.c_void_p, strs, C.c_int, strs, C.c_int, C.c_int, C.c_int, C.c_int,
                                   C.POINTER(TpKey)]
    L.tph_decode.argtypes = [C.c_void_p, strs, C.c_int, strs, C.c_int, C.c_int, C.POINTER(TpKey), C.c_char_p]
    for f in ("tph_weight_rows", "tph_group_count", "tph_pair_count"):
        getattr(L, f).argtypes = [C.c_void_p]
    L.tph_pairs.argtypes = [C.c_void_p, I32P, I32P, I32P, C.c_int]
    L.tph_index_of.argtypes = [C.c_void_p, C.c_char_p]
    L.tph_handle.restype = C.c_void_p
    L.tph_handle.argtypes = [C.c_void_p]
    L.tph_last_error.restype = C.c_char_p
    L.tph_last_error.argtypes = [C.c_void_p]
    _lib = L
    return L


def strerror(status: int) -> str:
    return lib().tp_strerror(status).decode()


def _i32(xs):
    return (C.c_int32 * max(len(xs), 1))(*xs)


def _strs(xs):
    arr = (C.c_char_p * max(len(xs), 1))()
    for i, s in enumerate(xs):
        arr[i] = s.encode()
    return arr


def _hdevs(devs):
    arr = (_HDevice * max(len(devs), 1))()
    for i, d in enumerate(devs):
        arr[i].id = d.Id.encode()
        arr[i].node_id = d.NodeId
        arr[i].numa_node = d.NumaNode
        arr[i].dev_id = d.DevId.encode()
    return arr


def _ids(buf, n):
    return [buf.raw[i * ID_LEN:(i + 1) * ID_LEN].split(b"\0", 1)[0].decode() for i in range(n)]


class BestEffortPolicy:
    """allocator.BestEffortPolicy (besteffort_policy.go:45-151) backed by the GPU search."""

    def __init__(self, cuda_device: int = -1, kernel: int = KERNEL_AUTO, exhaustive_limit: int = 0,
                 max_inflight: int = 0, extra_devices=(), multi_gpu_min_subsets: int = 0):
        """extra_devices: ordinals of further GPUs of this process; searches of at least
        multi_gpu_min_subsets subsets (default 2^31) are then cut over all of them inside Allocate."""
        self._opts = TpOptions(cuda_device, kernel, exhaustive_limit, max_inflight, len(extra_devices),
                               (C.c_int32 * 16)(*extra_devices), multi_gpu_min_subsets)
        self._p = lib().tph_new(C.byref(self._opts))
        self.last = None

    def close(self):
        if getattr(self, "_p", None):
            lib().tph_free(self._p)
            self._p = None

    __del__ = close

    def _raise(self, rc):
        raise AllocatorError(rc, lib().tph_last_error(self._p).decode())

    def Init(self, devs, topo_dir: str = ""):
        self.devices = list(devs)
        rc = lib().tph_init(self._p, _hdevs(devs), len(devs), topo_dir.encode())
        if rc < 0:
            self._raise(rc)

    def InitWithWeights(self, devs, pairs):
        """Init from p2pWeights given as (from_node, to_node, weight) triples."""
        self.devices = list(devs)
        rc = lib().tph_init_pairs(self._p, _hdevs(devs), len(devs), _i32([p[0] for p in pairs]),
                                  _i32([p[1] for p in pairs]), _i32([p[2] for p in pairs]), len(pairs))
        if rc < 0:
            self._raise(rc)

    def Allocate(self, available, required, size):
        required = required or []
        out = C.create_string_buffer(ID_LEN * max(size, 1))
        res = TpResult()
        rc = lib().tph_allocate(self._p, _strs(available), len(available), _strs(required), len(required), size,
                                out, C.byref(res))
        self.last = res.dict()
        if rc < 0:
            self._raise(rc)
        return _ids(out, rc)

    def set_tuning(self, walk_q=0, item_cap=0, item_cap1=0):
        """tp_set_tuning (development hook): override the work plan of requests prepared from now on."""
        lib().tp_set_tuning(self.handle, walk_q, item_cap, item_cap1)

    def set_batch_tuning(self, share_shift=0, items_per_warp=0, light_first=False):
        """tp_set_batch_tuning (development hook): the work plan of tp_allocate_batch from now on."""
        lib().tp_set_batch_tuning(self.handle, share_shift, items_per_warp, 1 if light_first else 0)

    def GlobalOptimum(self, available, required, size):
        """tp_global_optimum: the best set over ALL subsets (advisory; not what the reference returns)."""
        required = required or []
        av, rq = self.indices(available), self.indices(required)
        out = _i32([0] * max(size, 1))
        res = TpResult()
        rc = lib().tp_global_optimum(self.handle, _i32(av), len(av), _i32(rq), len(rq), size, out, C.byref(res))
        self.last = res.dict()
        if rc < 0:
            raise AllocatorError(rc, strerror(rc))
        return [self.devices[i].Id for i in out[:rc]]

    def SearchShard(self, available, required, size, shard_index, shard_count):
        """(key | None if short-circuited): this shard's packed winner."""
        required = required or []
        key = TpKey()
        rc = lib().tph_search_shard(self._p, _strs(available), len(available), _strs(required), len(required),
                                    size, shard_index, shard_count, C.byref(key))
        if rc < 0:
            self._raise(rc)
        return None if rc == 1 else (key.hi, key.lo)

    def Decode(self, available, required, size, key):
        required = required or []
        k = TpKey(key[0], key[1])
        out = C.create_string_buffer(ID_LEN * max(size, 1))
        rc = lib().tph_decode(self._p, _strs(available), len(available), _strs(required), len(required), size,
                              C.byref(k), out)
        if rc < 0:
            self._raise(rc)
        return _ids(out, rc)

    # introspection
    def weight_rows(self):
        return lib().tph_weight_rows(self._p)

    def group_count(self):
        return lib().tph_group_count(self._p)

    def pairs(self):
        n = lib().tph_pair_count(self._p)
        f, t, w = _i32([0] * n), _i32([0] * n), _i32([0] * n)
        lib().tph_pairs(self._p, f, t, w, n)
        return sorted(zip(f[:n], t[:n], w[:n]))

    def index_of(self, ident):
        return lib().tph_index_of(self._p, ident.encode())

    def indices(self, ids):
        return [self.index_of(i) for i in ids]

    @property
    def handle(self):
        """The raw tp_policy* for index-level calls (tp_allocate, tp_search_resident, ...)."""
        return lib().tph_handle(self._p)


def search_resident(pol: BestEffortPolicy, available, required, size, iters=1, shard_index=0, shard_count=1):
    """tp_search_resident: request uploaded once, kernel launched `iters` times between CUDA events.

    Returns (total_ms, key, result dict)."""
    av, rq = pol.indices(available), pol.indices(required or [])
    ms = C.c_float()
    key = TpKey()
    res = TpResult()
    rc = lib().tp_search_resident(pol.handle, _i32(av), len(av), _i32(rq), len(rq), size, shard_index, shard_count,
                                  iters, C.byref(ms), C.byref(key), C.byref(res))
    if rc < 0:
        msg = strerror(rc)
        if rc == STATUS["TP_ERR_CUDA"]:
            msg += ": " + lib().tp_last_cuda_error(pol.handle).decode()
        raise AllocatorError(rc, msg)
    return ms.value, (key.hi, key.lo), res.dict()


def shard_subsets(pol: BestEffortPolicy, available, required, size, shard_index, shard_count):
    """tp_shard_subsets (host arithmetic only): (subsets this shard scores, kernel); None for a short-circuit."""
    av, rq = pol.indices(available), pol.indices(required or [])
    n, k = C.c_uint64(), C.c_int32()
    rc = lib().tp_shard_subsets(pol.handle, _i32(av), len(av), _i32(rq), len(rq), size, shard_index, shard_count,
                                C.byref(n), C.cast(C.byref(k), I32P))
    if rc < 0:
        raise AllocatorError(rc, strerror(rc))
    return None if rc == 1 else (n.value, k.value)


def allocate_indices(pol: BestEffortPolicy, av, rq, size):
    """tp_allocate on raw indices; returns (indices, result dict)."""
    out = _i32([0] * max(size, 1))
    res = TpResult()
    rc = lib().tp_allocate(pol.handle, _i32(av), len(av), _i32(rq), len(rq), size, out, C.byref(res))
    if rc < 0:
        raise AllocatorError(rc, strerror(rc))
    return list(out[:rc]), res.dict()


class PreparedSearch:
    """tp_prepare / tp_launch: stream-ordered search writing its key to caller-owned device memory."""

    def __init__(self, pol: BestEffortPolicy, available, required, size, shard_index=0, shard_count=1):
        self._pol = pol
        av, rq = pol.indices(available), pol.indices(required or [])
        q = C.c_void_p()
        rc = lib().tp_prepare(pol.handle, _i32(av), len(av), _i32(rq), len(rq), size, shard_index, shard_count,
                              C.byref(q))
        if rc != 0:
            raise AllocatorError(rc, strerror(rc) if rc < 0 else "request is answered by a short-circuit")
        self._q = q
        self.subsets = lib().tp_prepared_subsets(q)

    def launch(self, cuda_stream: int, d_out_ptr: int) -> int:
        """Enqueue on cuda_stream; d_out_ptr -> 3 x uint64 in device memory. Returns kernels launched."""
        rc = lib().tp_launch(self._pol.handle, self._q, C.c_void_p(cuda_stream), C.c_void_p(d_out_ptr))
        if rc < 0:
            raise AllocatorError(rc, strerror(rc) + ": " + lib().tp_last_cuda_error(self._pol.handle).decode())
        return rc

    def launch_exchange(self, exchange, cuda_stream: int, d_out_ptr: int) -> int:
        """tp_launch_exchange: as launch(), but d_out receives the minimum over ALL ranks' shards (and the
        sum of their counts): the exchange is fused into the kernel's last block."""
        rc = lib().tp_launch_exchange(self._pol.handle, self._q, exchange._x, C.c_void_p(cuda_stream),
                                      C.c_void_p(d_out_ptr))
        if rc < 0:
            raise AllocatorError(rc, strerror(rc) + ": " + lib().tp_last_cuda_error(self._pol.handle).decode())
        return rc

    def close(self):
        if getattr(self, "_q", None):
            lib().tp_prepared_free(self._pol.handle, self._q)
            self._q = None

    __del__ = close


class Exchange:
    """tp_exchange_*: the fused winner exchange of a multi-process run.  `all_gather(handle_bytes)` must
    return every rank's 64-byte handle in rank order (e.g. through torch.distributed.all_gather_object)."""

    def __init__(self, pol: BestEffortPolicy, rank: int, world: int, all_gather):
        self._x = C.c_void_p()
        handle = C.create_string_buffer(64)
        rc = lib().tp_exchange_create(pol.handle, rank, world, C.byref(self._x), handle)
        if rc < 0:
            raise AllocatorError(rc, strerror(rc) + ": " + lib().tp_last_cuda_error(pol.handle).decode())
        handles = all_gather(handle.raw)
        assert len(handles) == world and all(len(h) == 64 for h in handles)
        rc = lib().tp_exchange_connect(self._x, b"".join(handles))
        if rc < 0:
            raise AllocatorError(rc, "tp_exchange_connect: " + strerror(rc))

    def close(self):
        if getattr(self, "_x", None):
            lib().tp_exchange_destroy(self._x)
            self._x = None

    __del__ = close


def measure_int_peak(cuda_device=0, variant=0) -> float:
    """10^9 integer lane-operations per second of the dependency-free microbenchmark."""
    g = C.c_double()
    rc = lib().tp_measure_int_peak(cuda_device, variant, C.byref(g))
    if rc < 0:
        raise AllocatorError(rc, strerror(rc))
    return g.value


class Batch:
    """Marshalled arguments of one tp_allocate_batch call (built once, callable many times)."""

    def __init__(self, pol: BestEffortPolicy, requests):
        self.pol, self.n = pol, len(requests)
        self.arr = (TpRequest * max(self.n, 1))()
        self._keep = []
        self.stride = max([r[2] for r in requests] + [1])
        for i, (av, rq, size) in enumerate(requests):
            a, r = _i32(pol.indices(av)), _i32(pol.indices(rq or []))
            self._keep.append((a, r))
            self.arr[i] = TpRequest(a, len(av), r, len(rq or []), size)
        self.out = (C.c_int32 * (max(self.n, 1) * self.stride))()
        self.status = (C.c_int32 * max(self.n, 1))()
        self.infos = (TpResult * max(self.n, 1))()

    def run(self):
        rc = lib().tp_allocate_batch(self.pol.handle, self.n, self.arr, self.out, self.stride, self.status,
                                     self.infos)
        if rc < 0:
            raise AllocatorError(rc, strerror(rc))

    def results(self):
        res = []
        for i in range(self.n):
            st = self.status[i]
            if st < 0:
                res.append(AllocatorError(st, strerror(st)))
            else:
                res.append([self.pol.devices[j].Id for j in self.out[i * self.stride:i * self.stride + st]])
        return res, [self.infos[i].dict() for i in range(self.n)]


def allocate_batch(pol: BestEffortPolicy, requests):
    """tp_allocate_batch: requests = [(available ids, required ids, size)], the ContainerRequests of
    one GetPreferredAllocation RPC (amdgpu.go:307-317).  Returns ([ids] or AllocatorError per request, infos)."""
    b = Batch(pol, requests)
    b.run()
    return b.results()


def build_weights(group, numa_node, link_type, cuda_device=0):
    """tp_build_weights: calculatePairWeight (device.go:135-157) for all pairs on the GPU.
    link_type is an n*n row-major list of KFD link types, -1 where no link entry exists."""
    n = len(group)
    out = (C.c_uint8 * (n * n))()
    rc = lib().tp_build_weights(n, _i32(group), _i32(numa_node), _i32(link_type), out, cuda_device)
    if rc < 0:
        raise AllocatorError(rc, strerror(rc))
    return list(out)
