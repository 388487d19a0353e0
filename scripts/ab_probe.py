"""A/B of two builds on one box (tuning aid): TP_ALLOC_LIB=<lib> python scripts/ab_probe.py
Prints per-shard kernel times of the C4 search for 8, 4, 2 and 1 shards, and the 8-of-63 search."""
import os, sys
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, R); sys.path.insert(0, os.path.join(R, "tests"))
import topo
from __graft_entry__ import package
tp = package()
def pol_of(devs, pairs):
    pol = tp.BestEffortPolicy(cuda_device=0)
    pol.InitWithWeights([tp.Device(d.Id, d.NodeId, d.NumaNode, d.DevId) for d in devs], pairs)
    return pol, [d.Id for d in devs]
pol, ids = pol_of(*topo.c4_topology())
for cnt in (8, 4, 2, 1):
    out, tot = [], 0
    for s in range(cnt):
        tp.search_resident(pol, ids, [], 8, iters=1, shard_index=s, shard_count=cnt)
        ms, key, info = tp.search_resident(pol, ids, [], 8, iters=5, shard_index=s, shard_count=cnt)
        out.append(ms / 5); tot += info["subsets_scored"]
    print("c4 %d shards: max %.3f  " % (cnt, max(out)) + " ".join("%.3f" % x for x in out), "ok" if tot == 28987537150 else "COUNT MISMATCH %d" % tot)
pol2, ids2 = pol_of(*topo.fixture_topology("cpx"))
for n in (63, 47, 40):
    tp.search_resident(pol2, ids2[:n], [], 8, iters=1)
    ms, key, info = tp.search_resident(pol2, ids2[:n], [], 8, iters=5)
    print("cpx 8 of %d: %.3f ms" % (n, ms / 5))
