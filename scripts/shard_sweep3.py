"""Slowest shard of C4 split 2, 4, 8 ways against the cap on pairs of single values: python scripts/shard_sweep3.py"""
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
total = 28987537150


def slowest(shards, cap, cap1):
    pol.set_tuning(item_cap=cap, item_cap1=cap1)
    worst = 0
    for s in range(shards):
        tp.search_resident(pol, ids, [], 8, iters=2, shard_index=s, shard_count=shards)
        ms, key, info = tp.search_resident(pol, ids, [], 8, iters=10, shard_index=s, shard_count=shards)
        worst = max(worst, ms / 10)
    return worst


for shards in (2, 4, 8):
    line = "c4/%d: default %.4f" % (shards, slowest(shards, 0, 0))
    for div in (165000, 220000, 330000):
        line += "  cap1=T/(%dk s) %.4f" % (div // 1000, slowest(shards, 0, total // (div * shards)))
    print(line, flush=True)
