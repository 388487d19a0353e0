"""Run one search several times (device-resident request) and print its rate: the ncu target.

usage: python scripts/run_one.py <c4|cpx|cpx64|mi308|...> <size> [iters] [kernel]
"""
import os, sys
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, R); sys.path.insert(0, os.path.join(R, "tests"))
import topo
from __graft_entry__ import package
tp = package()
name, size = sys.argv[1], int(sys.argv[2])
iters = int(sys.argv[3]) if len(sys.argv) > 3 else 3
kernel = int(sys.argv[4]) if len(sys.argv) > 4 else 0
SY = {"c4": topo.c4_topology, "hives": topo.hives_topology, "missing": topo.missing_links_topology, "cpx64": topo.cpx64_topology}
devs, pairs = SY[name]() if name in SY else topo.fixture_topology(name)
pol = tp.BestEffortPolicy(cuda_device=0, kernel=kernel)
pol.InitWithWeights([tp.Device(d.Id, d.NodeId, d.NumaNode, d.DevId) for d in devs], pairs)
ids = [d.Id for d in devs]
tp.search_resident(pol, ids, [], size, iters=1)
ms, key, info = tp.search_resident(pol, ids, [], size, iters=iters)
print(name, size, "kernel", info["kernel"], "ms/search %.4f" % (ms / iters), "subsets/s %.4g" % (info["subsets_total"] / (ms / iters * 1e-3)),
      "score", info["score"], pol.Decode(ids, [], size, key))
