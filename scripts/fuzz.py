"""Differential fuzz on the GPU: random topologies (group sizes, NUMA, link types, missing links)
and random requests; every answer must equal the oracle's BFS.  usage: fuzz.py <seconds> [seed]"""
import math, os, sys, time
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, R); sys.path.insert(0, os.path.join(R, "tests"))
import topo
from __graft_entry__ import package
from oracle.binding import RefPolicy, Device
tp = package()
budget = float(sys.argv[1]) if len(sys.argv) > 1 else 60
gen = topo.SplitMix64(int(sys.argv[2], 0) if len(sys.argv) > 2 else 0xF022)
t_end = time.time() + budget
n_topo = n_req = n_batch = 0
kernels = {}
while time.time() < t_end:
    groups = 2 + gen.below(11)
    devs, node = [], 2 + gen.below(5)
    numa_count = 1 + gen.below(4)
    for g in range(groups):
        size = 1 + gen.below(8)
        for j in range(size):
            if len(devs) >= 100:
                break
            ident = f"gpu{g:02d}" if j == 0 else f"amdgpu_xcp_{g * 8 + j}"
            devs.append(Device(ident, node, gen.below(numa_count), f"dev{g}"))
            node += 1 + gen.below(2)
    if len(devs) < 4:
        continue
    mode = gen.below(3)
    pairs = []
    for i, a in enumerate(devs):
        for b in devs[i + 1:]:
            if mode == 2 and gen.chance(0.1):
                continue  # missing link -> weight 0
            ty = 11 if mode == 0 else (11, 2, 7)[gen.below(3)]
            pairs.append((a.NodeId, b.NodeId, topo.pair_weight(a, b, ty)))
    ref = RefPolicy(); ref.InitPairs(devs, pairs)
    pol = tp.BestEffortPolicy(cuda_device=0)
    pol.InitWithWeights([tp.Device(d.Id, d.NodeId, d.NumaNode, d.DevId) for d in devs], pairs)
    ids = [d.Id for d in devs]
    n_topo += 1
    if n_topo % 3 == 0:  # forced small work items and cut values: the work plan's second level on every request
        pol.set_tuning(item_cap=64 + gen.below(400), item_cap1=64 + gen.below(300))
    spl = None
    if n_topo % 4 == 0:  # the round-1 kernel beside the default choice: same key
        spl = tp.BestEffortPolicy(cuda_device=0, kernel=tp.KERNEL_SPLIT)
        spl.InitWithWeights([tp.Device(d.Id, d.NodeId, d.NumaNode, d.DevId) for d in devs], pairs)
    for _ in range(40):
        av, rq, k = topo.random_request(gen, ids, kmax=min(9, len(ids) - 1), p_choices=(0.3, 0.6, 0.9), max_required=3)
        # keep the oracle's BFS (k-permutations on unpartitioned nodes) and the GPU search bounded
        if math.comb(len(av), k) > 3e9 or (groups > 8 and k > 6):
            continue
        want = ref.Allocate(av, rq, k)
        got = pol.Allocate(av, rq, k)
        if got != want or (ref.last["searched"] and pol.last["score"] != ref.last["score"]):
            print("MISMATCH", dict(groups=groups, n=len(devs), mode=mode, av=av, rq=rq, k=k, got=got, want=want,
                                   gpu=pol.last, ref=ref.last))
            sys.exit(1)
        n_req += 1
        kernels[pol.last["kernel"]] = kernels.get(pol.last["kernel"], 0) + 1
        if spl is not None and ref.last["searched"]:
            if spl.Allocate(av, rq, k) != want or spl.last["key"] != pol.last["key"]:
                print("SPLIT/TRI MISMATCH", dict(av=av, rq=rq, k=k, split=spl.last, default=pol.last))
                sys.exit(1)
    if n_topo % 5 == 0:  # the same kind of requests through ONE tp_allocate_batch call (block-per-request kernels, shares)
        reqs = []
        while len(reqs) < 24:
            av, rq, k = topo.random_request(gen, ids, kmax=min(9, len(ids) - 1), p_choices=(0.3, 0.6, 0.9), max_required=3)
            if math.comb(len(av), k) > 3e8 or (groups > 8 and k > 6):
                continue
            reqs.append((av, rq, k))
        if n_topo % 10 == 0:
            pol.set_batch_tuning(share_shift=20 + gen.below(6), items_per_warp=1 + gen.below(8))
        got, _ = tp.allocate_batch(pol, reqs)
        for r, g in zip(reqs, got):
            want = ref.Allocate(*r)
            if g != want:
                print("BATCH MISMATCH", dict(groups=groups, n=len(devs), req=r, got=g, want=want))
                sys.exit(1)
        n_batch += len(reqs)
    pol.close()
    if spl is not None:
        spl.close()
print(f"fuzz ok: {n_topo} topologies, {n_req} requests, kernels {kernels}, {n_batch} requests in batches")
