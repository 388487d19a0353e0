"""Single-request time by pool, LEX against SPLIT (tuning aid).  python scripts/slow_probe.py"""
import os, sys
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, R); sys.path.insert(0, os.path.join(R, "tests"))
import topo
from __graft_entry__ import package
tp = package()
devs, pairs = topo.fixture_topology("cpx")
ids = [d.Id for d in devs]
pols = {}
for name, kern in (("auto", 0), ("lex", tp.KERNEL_LEX)):
    pols[name] = tp.BestEffortPolicy(cuda_device=0, kernel=kern)
    pols[name].InitWithWeights([tp.Device(d.Id, d.NodeId, d.NumaNode, d.DevId) for d in devs], pairs)
gen = topo.SplitMix64(5)
for k in (8,):
    for n in (16, 24, 31, 40, 47, 55, 63):
        pool = sorted(ids, key=lambda _: gen.next())[:n]
        ordered = ids[:n]
        for label, pl in (("random", pool), ("first-n", ordered)):
            out = []
            for name, pol in pols.items():
                tp.search_resident(pol, pl, [], k, iters=1)
                ms, key, info = tp.search_resident(pol, pl, [], k, iters=5)
                out.append("%s(k%d) %.3f ms" % (name, info["kernel"], ms / 5))
            print("k=%d n=%d %-8s subsets %.3g score %d m %d: %s" % (k, n, label, info["subsets_total"], info["score"],
                                                                 info["groups_touched"], "  ".join(out)))
