"""Slowest shard of C4 split 8 (and 2, 4) ways against both item-size knobs: python scripts/shard_sweep2.py"""
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


def slowest(shards, cap, cap1):
    pol.set_tuning(item_cap=cap, item_cap1=cap1)
    worst = 0
    for s in range(shards):
        tp.search_resident(pol, ids, [], 8, iters=2, shard_index=s, shard_count=shards)
        ms, key, info = tp.search_resident(pol, ids, [], 8, iters=10, shard_index=s, shard_count=shards)
        worst = max(worst, ms / 10)
    return worst


for shards in (2, 4, 8):
    print("c4/%d defaults: %.4f ms" % (shards, slowest(shards, 0, 0)), flush=True)
for cap in (12288, 16384, 20480, 24576, 32768):
    line = "c4/8 cap0 %6d:" % cap
    for cap1 in (16384, 24576, 32768, 49152):
        line += "  cap1 %5d: %.4f" % (cap1, slowest(8, cap, cap1))
    print(line, flush=True)
