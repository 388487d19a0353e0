"""One shard of the C4 search, a few launches: python scripts/shard_one.py <shard> <count>"""
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
s, cnt = int(sys.argv[1]), int(sys.argv[2])
ms, key, info = tp.search_resident(pol, ids, [], 8, iters=3, shard_index=s, shard_count=cnt)
print("shard %d/%d: %.3f ms" % (s, cnt, ms / 3))
