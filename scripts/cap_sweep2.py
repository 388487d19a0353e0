"""C3 / C4 / one 1/8 shard of C4 against both item-size knobs (tp_set_tuning): python scripts/cap_sweep2.py"""
import os, sys
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, R); sys.path.insert(0, os.path.join(R, "tests"))
import topo
from __graft_entry__ import package
tp = package()
for name, mk, shards in (("cpx64", topo.cpx64_topology, 1), ("c4", topo.c4_topology, 1), ("c4/8", topo.c4_topology, 8)):
    devs, pairs = mk()
    pol = tp.BestEffortPolicy(cuda_device=0)
    pol.InitWithWeights([tp.Device(d.Id, d.NodeId, d.NumaNode, d.DevId) for d in devs], pairs)
    ids = [d.Id for d in devs]
    for cap in (0, 8192, 16384, 32768, 65536, 131072):
        line = "%s cap0 %6d:" % (name, cap)
        for cap1 in (0, 4096, 8192, 16384, 32768):
            if cap == 0 and cap1: continue
            pol.set_tuning(item_cap=cap, item_cap1=cap1)
            worst = 0
            for s in ((0,) if shards == 1 else (0, 3, 7)):
                tp.search_resident(pol, ids, [], 8, iters=2, shard_index=s, shard_count=shards)
                ms, key, info = tp.search_resident(pol, ids, [], 8, iters=8, shard_index=s, shard_count=shards)
                worst = max(worst, ms / 8)
            line += "  cap1 %5d: %.4f" % (cap1, worst)
        print(line, flush=True)
    pol.close()
