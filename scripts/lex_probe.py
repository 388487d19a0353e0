import os, sys
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, R); sys.path.insert(0, os.path.join(R, "tests"))
import topo
from __graft_entry__ import package
tp = package()
devs, pairs = topo.fixture_topology("cpx")
pol = tp.BestEffortPolicy(cuda_device=0)
pol.InitWithWeights([tp.Device(d.Id, d.NodeId, d.NumaNode, d.DevId) for d in devs], pairs)
prim = [f"test{i}" for i in range(1, 9)]
for k in (2, 3, 4, 5, 6):
    ms, key, info = tp.search_resident(pol, prim, [], k, iters=200)
    print("primaries k=%d: %.2f us/launch kernel %d scored %d" % (k, ms / 200 * 1e3, info["kernel"], info["subsets_scored"]))
ids = [d.Id for d in devs]
for k in (2, 3):
    ms, key, info = tp.search_resident(pol, ids, [], k, iters=200)
    print("all63 k=%d: %.2f us/launch kernel %d" % (k, ms / 200 * 1e3, info["kernel"]))
