"""Differential fuzz of the LARGE searches: random topologies of 48 .. 84 partitions (groups of 4 / 8, NUMA, link
types, missing links), requests of 7 .. 9 devices out of most of them (up to ~4e10 subsets), every answer against the
oracle's BFS; every other request also as 3 or 8 shards whose keys are min-reduced and decoded.
usage: fuzz_big.py <seconds> [seed]"""
import math, os, sys, time
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, R); sys.path.insert(0, os.path.join(R, "tests"))
import topo
from __graft_entry__ import package
from oracle.binding import RefPolicy, Device
tp = package()
budget = float(sys.argv[1]) if len(sys.argv) > 1 else 60
gen = topo.SplitMix64(int(sys.argv[2], 0) if len(sys.argv) > 2 else 0xB16)
t_end = time.time() + budget
n_topo = n_req = n_sharded = 0
subsets = 0
while time.time() < t_end:
    per = (4, 8)[gen.below(2)]
    groups = (48 // per) + gen.below((84 - 48) // per + 1)
    devs, node = [], 2 + gen.below(5)
    numa_count = 1 + gen.below(4)
    for g in range(groups):
        for j in range(per):
            ident = f"gpu{g:02d}" if j == 0 else f"amdgpu_xcp_{g * 8 + j}"
            devs.append(Device(ident, node, gen.below(numa_count), f"dev{g}"))
            node += 1 + gen.below(2)
    mode = gen.below(3)
    pairs = []
    for i, a in enumerate(devs):
        for b in devs[i + 1:]:
            if mode == 2 and gen.chance(0.1):
                continue
            ty = 11 if mode == 0 else (11, 2, 7)[gen.below(3)]
            pairs.append((a.NodeId, b.NodeId, topo.pair_weight(a, b, ty)))
    ref = RefPolicy(); ref.InitPairs(devs, pairs)
    pol = tp.BestEffortPolicy(cuda_device=0)
    pol.InitWithWeights([tp.Device(d.Id, d.NodeId, d.NumaNode, d.DevId) for d in devs], pairs)
    ids = [d.Id for d in devs]
    n_topo += 1
    for _ in range(6):
        k = 7 + gen.below(3)
        av = [i for i in ids if gen.chance((0.85, 0.95, 1.0)[gen.below(3)])]
        rq = [av[gen.below(len(av))]] if gen.chance(0.25) else []
        total = math.comb(len(av) - len(rq), k - len(rq))
        if total > 4e10 or len(av) <= k:
            continue
        want = ref.Allocate(av, rq, k)
        got = pol.Allocate(av, rq, k)
        if got != want or pol.last["score"] != ref.last["score"] or pol.last["subsets_scored"] != total:
            print("MISMATCH", dict(n=len(devs), per=per, mode=mode, av=len(av), rq=rq, k=k, got=got, want=want, gpu=pol.last, ref=ref.last))
            sys.exit(1)
        n_req += 1
        subsets += total
        if n_req % 2 == 0:
            P = (3, 8)[gen.below(2)]
            keys = [pol.SearchShard(av, rq, k, s, P) for s in range(P)]
            if pol.Decode(av, rq, k, min(keys)) != want:
                print("SHARD MISMATCH", dict(n=len(devs), av=len(av), rq=rq, k=k, P=P, keys=keys, want=want))
                sys.exit(1)
            n_sharded += 1
    pol.close()
print(f"fuzz_big ok: {n_topo} topologies, {n_req} requests ({n_sharded} also sharded), {subsets:.3e} subsets scored")
