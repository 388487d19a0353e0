"""Time one shard of the C4 search on one GPU: python scripts/shard_time.py <shard_count>"""
import os, sys
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, R); sys.path.insert(0, os.path.join(R, "tests"))
import topo
from __graft_entry__ import package
tp = package()
devs, pairs = topo.c4_topology()
pol = tp.BestEffortPolicy(cuda_device=0)
pol.InitWithWeights([tp.Device(d.Id, d.NodeId, d.NumaNode, d.DevId) for d in devs], pairs)
ids = [d.Id for d in devs]
cnt = int(sys.argv[1])
tot = 0
for s in range(cnt):
    tp.search_resident(pol, ids, [], 8, iters=1, shard_index=s, shard_count=cnt)
    ms, key, info = tp.search_resident(pol, ids, [], 8, iters=5, shard_index=s, shard_count=cnt)
    tot += info["subsets_scored"]
    print("shard %d/%d: %.3f ms, %d subsets, key %x" % (s, cnt, ms / 5, info["subsets_scored"], key[0]))
print("sum", tot)
