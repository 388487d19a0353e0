"""GPU tests of the callers either side of the search: the batch entry (the ContainerRequests
loop, amdgpu.go:307-317), concurrent callers (one goroutine per RPC in the reference) and the
pair-weight construction (calculatePairWeight, device.go:135-157)."""
import threading

import pytest

import topo
from __graft_entry__ import package
from oracle.binding import OracleError, RefPolicy

pytestmark = pytest.mark.gpu
tp = package()


def make(name="cpx"):
    devs, pairs = topo.fixture_topology(name)
    ref = RefPolicy()
    ref.InitPairs(devs, pairs)
    pol = tp.BestEffortPolicy(cuda_device=0)
    pol.InitWithWeights([tp.Device(d.Id, d.NodeId, d.NumaNode, d.DevId) for d in devs], pairs)
    return ref, pol, [d.Id for d in devs]


def pod_storm(ids, count, seed=0x5EED0001):
    """BASELINE config C5: availability p in {.25,.5,.75}, k in {1,2,4,8}, 20% one required."""
    gen = topo.SplitMix64(seed)
    reqs = []
    while len(reqs) < count:
        p = (0.25, 0.5, 0.75)[gen.below(3)]
        av = [i for i in ids if gen.chance(p)]
        k = (1, 2, 4, 8)[gen.below(4)]
        if len(av) < k:
            continue
        rq = [av[gen.below(len(av))]] if gen.chance(0.2) else []
        reqs.append((av, rq, k))
    return reqs


def test_pod_storm_batch_matches_reference():
    ref, pol, ids = make("cpx")
    reqs = pod_storm(ids, 2000)
    got, infos = tp.allocate_batch(pol, reqs)
    searched = 0
    for (av, rq, k), g, info in zip(reqs, got, infos):
        want = ref.Allocate(av, rq, k)
        assert g == want, (av, rq, k)
        if ref.last["searched"]:
            searched += 1
            assert info["score"] == ref.last["score"] and info["gpu_launches"] >= 1
    assert searched > 1500
    # errors are per request, in the reference's words, and do not poison the batch
    bad = [(ids[:3], [], 5), (ids, [], 0), (ids[:10], [], 2)]
    got, _ = tp.allocate_batch(pol, bad)
    assert str(got[0]) == "available devices count less than allocation size"
    assert str(got[1]) == "allocation size can not be negative"
    assert got[2] == ref.Allocate(ids[:10], [], 2)
    pol.close()


def test_two_batches_at_once_and_share_cuts():
    """Two tp_allocate_batch calls on one policy at the same time (separate buffers and streams per call), and
    the work plan of the batch at every share cut: few heavy requests (each cut into up to 16 shares), forced
    cuts, the light wave first."""
    ref, pol, ids = make("cpx")
    reqs = pod_storm(ids, 600, seed=0xBA7C4)
    want = [ref.Allocate(*r) for r in reqs]
    results = [None, None]

    def run(slot):
        results[slot] = tp.allocate_batch(pol, reqs)[0]

    threads = [threading.Thread(target=run, args=(t,)) for t in range(2)]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    assert results[0] == want and results[1] == want
    heavy = [r for r in reqs if r[2] == 8 and len(r[0]) >= 40][:12] + reqs[:20]
    want_heavy = [ref.Allocate(*r) for r in heavy]
    for shift, ipw, light_first in ((0, 0, False), (20, 2, False), (22, 16, True), (30, 1, False)):
        pol.set_batch_tuning(share_shift=shift, items_per_warp=ipw, light_first=light_first)
        got, infos = tp.allocate_batch(pol, heavy)
        assert got == want_heavy, (shift, ipw, light_first)
        for r, info in zip(heavy, infos):
            if info["searched"]:
                assert info["subsets_scored"] == info["subsets_total"], (shift, r)
    pol.close()


def test_concurrent_callers():
    """8 threads on one policy (grpc-go serves each RPC on its own goroutine, dpm/plugin.go:108-111)."""
    ref, pol, ids = make("mi308")
    reqs = pod_storm(ids, 400, seed=99)
    want = [ref.Allocate(*r) for r in reqs]
    got = [None] * len(reqs)

    # each thread needs its own view of `last`, so go through the index-level ABI
    def worker_idx(t):
        for i in range(t, len(reqs), 8):
            av, rq, k = reqs[i]
            out, _ = tp.allocate_indices(pol, pol.indices(av), pol.indices(rq), k)
            got[i] = [pol.devices[j].Id for j in out]

    threads = [threading.Thread(target=worker_idx, args=(t,)) for t in range(8)]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    assert got == want
    pol.close()


def test_concurrent_callers_with_different_shared_memory_needs():
    """Threads whose requests need the same kernels with different dynamic shared memory (8 of 20 .. 8 of 60 through
    TRI, 5 of n through SPLIT): the opt-in of a kernel's shared memory must not depend on who launched last."""
    ref, pol, ids = make("cpx")
    gen = topo.SplitMix64(0xD1FF)
    reqs = []
    for i in range(96):
        n = 20 + gen.below(41)
        av = sorted(ids, key=lambda _: gen.below(1 << 30))[:n]
        av = [x for x in ids if x in set(av)]
        reqs.append((av, [], 8 if i % 3 else 5))
    want = [ref.Allocate(*r) for r in reqs]
    got = [None] * len(reqs)
    errors = []

    def worker(t):
        try:
            for i in range(t, len(reqs), 6):
                av, rq, k = reqs[i]
                out, _ = tp.allocate_indices(pol, pol.indices(av), pol.indices(rq), k)
                got[i] = [pol.devices[j].Id for j in out]
        except Exception as e:  # noqa: BLE001 -- reported below
            errors.append(repr(e))

    threads = [threading.Thread(target=worker, args=(t,)) for t in range(6)]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    assert not errors, errors
    assert got == want
    pol.close()


def test_build_weights_on_gpu():
    devs, pairs = topo.hives_topology()
    n = len(devs)
    groups = {}
    group = [groups.setdefault(d.DevId, len(groups)) for d in devs]
    numa = [d.NumaNode for d in devs]
    link = [-1] * (n * n)
    for i, a in enumerate(devs):
        for j, b in enumerate(devs):
            if i != j:
                link[i * n + j] = 11 if int(a.DevId) // 2 == int(b.DevId) // 2 else 2
    link[0 * n + 5] = link[5 * n + 0] = 7    # some other link type -> otherLinkWeight
    link[1 * n + 9] = link[9 * n + 1] = -1   # no entry in p2pWeights -> 0
    w = tp.build_weights(group, numa, link)
    for i, a in enumerate(devs):
        for j, b in enumerate(devs):
            ty = link[i * n + j]
            want = 0 if (i == j or ty < 0) else topo.pair_weight(a, b, ty)
            assert w[i * n + j] == want, (i, j, ty)
