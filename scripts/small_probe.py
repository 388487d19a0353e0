"""Kernel time of small and mid-size single searches (fixed-cost probe): TP_ALLOC_LIB=<lib> python scripts/small_probe.py"""
import os, sys
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, R); sys.path.insert(0, os.path.join(R, "tests"))
import topo
from __graft_entry__ import package
tp = package()
devs, pairs = topo.fixture_topology("cpx")
pol = tp.BestEffortPolicy(cuda_device=0)
pol.InitWithWeights([tp.Device(d.Id, d.NodeId, d.NumaNode, d.DevId) for d in devs], pairs)
ids = [d.Id for d in devs]
gen = topo.SplitMix64(5)
out = []
for n in (20, 24, 31, 40, 47, 63):
    pool = sorted(ids, key=lambda _: gen.next())[:n]
    for pl in (pool, ids[:n]):
        tp.search_resident(pol, pl, [], 8, iters=1)
        ms, key, info = tp.search_resident(pol, pl, [], 8, iters=10)
        out.append("%d:%.1f" % (n, ms / 10 * 1e3))
print("us per search (random pool, first-n):", " ".join(out))
