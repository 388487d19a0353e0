"""Host arithmetic of the sharded search (tp_shard_subsets, host-only handle): whatever the shard
count and the slicing of the work plan, the shards' counts must add up to C(free, k') -- the plan's
slices (pieces of the first walked position, single values cut by the second one) partition the
subset space.  The kernel decodes items with the same shared helper (tp_split_slice), so this pins
the item arithmetic without a GPU; the GPU tests then check what the items score."""
from math import comb

import pytest

import topo
from __graft_entry__ import package

tp = package()


def host_policy(devs, pairs):
    pol = tp.BestEffortPolicy(cuda_device=-2)  # TP_DEVICE_NONE
    pol.InitWithWeights([tp.Device(d.Id, d.NodeId, d.NumaNode, d.DevId) for d in devs], pairs)
    return pol, [d.Id for d in devs]


@pytest.mark.parametrize("shards", [1, 2, 3, 8])
def test_c4_shards_add_up(shards):
    pol, ids = host_policy(*topo.c4_topology())
    total = 0
    for s in range(shards):
        n, kernel = tp.shard_subsets(pol, ids, [], 8, s, shards)
        assert kernel == tp.KERNEL_TRI  # k' = 8 on 80 devices: the triple-table kernel, same work plan
        total += n
    assert total == comb(80, 8) == 28987537150


@pytest.mark.parametrize("caps", [None, ("200", "150"), ("64", "64"), ("5000", "300")])
def test_slices_partition_the_space(caps, monkeypatch):
    pol, ids = host_policy(*topo.fixture_topology("cpx"))
    if caps:
        pol.set_tuning(item_cap=int(caps[0]), item_cap1=int(caps[1]))
    gen = topo.SplitMix64(0x9A17)
    for size in (4, 5, 6, 7, 8, 9, 10, 12):  # walk depths 0 .. 4
        for _ in range(3):
            n_av = size + 3 + gen.below(24 if size <= 8 else 12)
            av = ids[:n_av]
            rq = av[:gen.below(2)]
            shards = 1 + gen.below(6)
            got = [tp.shard_subsets(pol, av, rq, size, s, shards) for s in range(shards)]
            assert sum(n for n, _ in got) == comb(n_av - len(rq), size - len(rq)), (size, n_av, len(rq), shards)


def test_short_circuit_and_errors():
    pol, ids = host_policy(*topo.fixture_topology("cpx"))
    assert tp.shard_subsets(pol, ids[:3], [], 3, 0, 1) is None
    with pytest.raises(tp.AllocatorError, match="available devices count less than allocation size"):
        tp.shard_subsets(pol, ids[:2], [], 3, 0, 1)
