"""Body of tests/test_gpu_exchange.py: run under torch.distributed.run, one rank per GPU.  Every rank searches its
shard with the exchange fused into the kernel (tp_launch_exchange) and must end up with the reference's answer."""
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, HERE)

import torch
import torch.distributed as dist

import topo
from __graft_entry__ import package
from oracle.binding import RefPolicy

tp = package()
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("gloo")


def gather(raw):
    got = [None] * world
    dist.all_gather_object(got, raw)
    return got


def run(name, devs, pairs, requests):
    ref = RefPolicy()
    ref.InitPairs(devs, pairs)
    pol = tp.BestEffortPolicy(cuda_device=local)
    pol.InitWithWeights([tp.Device(d.Id, d.NodeId, d.NumaNode, d.DevId) for d in devs], pairs)
    ex = tp.Exchange(pol, rank, world, gather)
    out = torch.zeros(3, dtype=torch.int64, device=f"cuda:{local}")
    stream = torch.cuda.current_stream()
    for av, rq, k, exhaustive in requests:
        want = ref.AllocateExhaustive(av, rq, k) if exhaustive else ref.Allocate(av, rq, k)
        if not exhaustive and not ref.last["searched"]:
            continue
        q = tp.PreparedSearch(pol, av, rq, k, shard_index=rank, shard_count=world)
        for _ in range(2):  # twice: both slot parities
            q.launch_exchange(ex, stream.cuda_stream, out.data_ptr())
            hi, lo, scored = out.cpu().tolist()
            got = pol.Decode(av, rq, k, (hi, lo & 0xFFFFFFFFFFFFFFFF))
            assert got == want, (name, rank, av, rq, k, got, want)
            total = tp.shard_subsets(pol, av, rq, k, 0, 1)
            if total and total[1] != tp.KERNEL_FAMILY:
                assert scored == total[0], (name, rank, scored, total)
        q.close()
    ex.close()
    pol.close()


devs, pairs = topo.fixture_topology("cpx")
ids = [d.Id for d in devs]
gen = topo.SplitMix64(0xE8C4)
reqs = [(ids, [], 8, False), (ids[:9], [], 2, False), (ids[:30], [ids[3]], 7, False)]
for _ in range(25):
    av, rq, k = topo.random_request(gen, ids, kmax=8, p_choices=(0.3, 0.6))
    reqs.append((av, rq, k, False))
run("cpx", devs, pairs, reqs)
devs, pairs = topo.wide_topology()
ids = [d.Id for d in devs]
singles = [ids[g * 8 + (g % 3)] for g in range(16)]
run("wide", devs, pairs, [(singles, [], 10, True), (singles[1:], [singles[4]], 10, True), (ids, [], 4, False)])
dist.barrier()
if rank == 0:
    print("exchange ok on", world, "ranks")
dist.destroy_process_group()
