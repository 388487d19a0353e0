"""Per-shard kernel time of the C4 search at 8 / 4 / 2 shards against the item-size knob (tuning aid)."""
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
for cnt in (8, 4, 2):
    for cap in (0, 8192, 16384, 24576, 32768, 49152, 65536):
        pol.set_tuning(item_cap=cap)
        out = []
        for s in range(cnt):
            tp.search_resident(pol, ids, [], 8, iters=1, shard_index=s, shard_count=cnt)
            ms, key, info = tp.search_resident(pol, ids, [], 8, iters=5, shard_index=s, shard_count=cnt)
            out.append(ms / 5)
        print("%d shards cap %6d: max %.3f mean %.3f  " % (cnt, cap, max(out), sum(out) / cnt) + " ".join("%.3f" % x for x in out), flush=True)
