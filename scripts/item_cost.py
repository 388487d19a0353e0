"""Cost of a work item: C3 at the default plan against the same search cut into many more items."""
import os, sys
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, R); sys.path.insert(0, os.path.join(R, "tests"))
import topo
from __graft_entry__ import package
tp = package()
devs, pairs = topo.cpx64_topology()
pol = tp.BestEffortPolicy(cuda_device=0)
pol.InitWithWeights([tp.Device(d.Id, d.NodeId, d.NumaNode, d.DevId) for d in devs], pairs)
ids = [d.Id for d in devs]
for cap, cap1 in ((0, 0), (32768, 16384), (32768, 8192), (32768, 4096), (32768, 2048)):
    pol.set_tuning(item_cap=cap, item_cap1=cap1)
    tp.search_resident(pol, ids, [], 8, iters=2)
    ms, key, info = tp.search_resident(pol, ids, [], 8, iters=10)
    print("cap0 %d cap1 %d: %.4f ms" % (cap, cap1, ms / 10), flush=True)
