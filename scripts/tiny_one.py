"""One small SPLIT search, launched a few times (ncu target for the fixed cost): python scripts/tiny_one.py [n]"""
import os, sys
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, R); sys.path.insert(0, os.path.join(R, "tests"))
import topo
from __graft_entry__ import package
tp = package()
devs, pairs = topo.fixture_topology("cpx")
pol = tp.BestEffortPolicy(cuda_device=0)
pol.InitWithWeights([tp.Device(d.Id, d.NodeId, d.NumaNode, d.DevId) for d in devs], pairs)
ids = [d.Id for d in devs][:int(sys.argv[1]) if len(sys.argv) > 1 else 24]
tp.search_resident(pol, ids, [], 8, iters=1)
ms, key, info = tp.search_resident(pol, ids, [], 8, iters=5)
print("kernel", info["kernel"], "ms/search %.4f" % (ms / 5), info["subsets_total"])
