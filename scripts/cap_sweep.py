"""Kernel time of the C3 / C4 searches against the item-size knob (tp_set_tuning): python scripts/cap_sweep.py"""
import os, sys
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, R); sys.path.insert(0, os.path.join(R, "tests"))
import topo
from __graft_entry__ import package
tp = package()
for name, mk in (("cpx64", topo.cpx64_topology), ("c4", topo.c4_topology)):
    devs, pairs = mk()
    pol = tp.BestEffortPolicy(cuda_device=0)
    pol.InitWithWeights([tp.Device(d.Id, d.NodeId, d.NumaNode, d.DevId) for d in devs], pairs)
    ids = [d.Id for d in devs]
    for cap in (0, 32768, 65536, 131072, 262144, 524288):
        pol.set_tuning(item_cap=cap)
        tp.search_resident(pol, ids, [], 8, iters=2)
        ms, key, info = tp.search_resident(pol, ids, [], 8, iters=10)
        print("%s cap %d: %.4f ms" % (name, cap, ms / 10), flush=True)
    pol.close()
