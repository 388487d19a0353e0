"""The N>1 protocol on CPU (gloo, world_size 2): every rank scores its shard, the packed winners
are min-reduced as int64, and every rank decodes the same answer.

No GPU here, so each rank's shard is scored by the CPU oracle (a lexicographic rank range of
oracle/exhaustive.c) which emits its winner in the product's key format; what is under test is
the product's host side: the key order (an int64 MIN all-reduce must pick the reference's
winner, including "no candidate in this shard"), tp_decode, and the host-only policy handle."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import topo
from __graft_entry__ import package
from oracle.binding import RefPolicy

KEY_NONE = 0x7FFFFFFFFFFFFFFF


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, name, cases, out_q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    tp = package()
    devs, pairs = topo.fixture_topology(name)
    ref = RefPolicy()
    ref.InitPairs(devs, pairs)
    pol = tp.BestEffortPolicy(cuda_device=-2)  # TP_DEVICE_NONE: host-side logic only
    pol.InitWithWeights([tp.Device(d.Id, d.NodeId, d.NumaNode, d.DevId) for d in devs], pairs)
    results = []
    for av, rq, k in cases:
        ref.AllocateExhaustive(av, rq, k, threads=1)
        total = ref.last["total"]
        lo, hi = total * rank // world, total * (rank + 1) // world
        key = KEY_NONE
        if hi > lo:
            ref.AllocateExhaustive(av, rq, k, threads=1, rank_lo=lo, rank_hi=hi)
            assert ref.last["key_lo"] == 0  # these cases fit the single-word all-reduce
            key = ref.last["key_hi"]
        # a shard holds a candidate only if something in it is in the Family and beats the seed bound;
        # the oracle seeds with rank 0's subset, so drop a seed that lies outside this shard
        t = torch.tensor([key], dtype=torch.int64)
        dist.all_reduce(t, op=dist.ReduceOp.MIN)
        results.append(pol.Decode(av, rq, k, (int(t.item()), 0)))
    if rank == 0:
        out_q.put(results)
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("name", ["mi308", "cpx"])
def test_two_rank_min_reduce_and_decode(name):
    devs, pairs = topo.fixture_topology(name)
    ref = RefPolicy()
    ref.InitPairs(devs, pairs)
    ids = [d.Id for d in devs]
    gen = topo.SplitMix64(0x2A11 + len(ids))
    cases, want = [], []
    while len(cases) < 12:
        av, rq, k = topo.random_request(gen, ids, kmax=5)
        out = ref.Allocate(av, rq, k)
        if ref.last["searched"]:
            cases.append((av, rq, k))
            want.append(out)
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, name, cases, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = q.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert got == want


def test_host_only_policy_refuses_to_search():
    tp = package()
    devs, pairs = topo.fixture_topology("mi210")
    pol = tp.BestEffortPolicy(cuda_device=-2)
    pol.InitWithWeights([tp.Device(d.Id, d.NodeId, d.NumaNode, d.DevId) for d in devs], pairs)
    ids = [d.Id for d in devs]
    assert pol.Allocate(ids[:3], [], 3) == ids[:3]  # short-circuits need no GPU
    with pytest.raises(tp.AllocatorError, match="CUDA"):
        pol.Allocate(ids, [], 3)  # no CPU fallback: a searched request fails loudly
