"""A few GetPreferredAllocation calls on the latency path (C2: 4 of the 8 CPX primaries, 70 subsets): ncu target
for tp_exhaustive_lex_small.   python scripts/latency_one.py [calls]"""
import os, sys, time
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, R); sys.path.insert(0, os.path.join(R, "tests"))
import topo
from __graft_entry__ import package
tp = package()
devs, pairs = topo.fixture_topology("cpx")
pol = tp.BestEffortPolicy(cuda_device=0)
pol.InitWithWeights([tp.Device(d.Id, d.NodeId, d.NumaNode, d.DevId) for d in devs], pairs)
av = pol.indices([f"test{i}" for i in range(1, 9)])
calls = int(sys.argv[1]) if len(sys.argv) > 1 else 20
lat = []
for _ in range(calls):
    t = time.perf_counter()
    got, info = tp.allocate_indices(pol, av, [], 4)
    lat.append((time.perf_counter() - t) * 1e6)
print("latency path: kernel", info["kernel"], "subsets", info["subsets_total"], "median us %.1f" % sorted(lat)[len(lat) // 2], got)
