"""Phase timestamps of block 0 of kernel TRI (needs a -DTP_TRI_TIMING build): TP_ALLOC_LIB=... python scripts/tri_phases.py"""
import os, sys, ctypes as C
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, R); sys.path.insert(0, os.path.join(R, "tests"))
import topo
from __graft_entry__ import package
tp = package()
names = ["entry", "setup in", "copies", "weights", "seeds", "table", "setup out", "items done", "finish"]
CASES = [("cpx", topo.fixture_topology, ("cpx",), 40, 1, 64), ("cpx64", topo.cpx64_topology, (), 64, 1, 64),
         ("c4", topo.c4_topology, (), 80, 1, 80), ("c4 shard 3/8", topo.c4_topology, (), 80, 8, 80)]
for label, mk, args, n, shards, m in CASES:
    devs, pairs = mk(*args)
    pol = tp.BestEffortPolicy(cuda_device=0)
    pol.InitWithWeights([tp.Device(d.Id, d.NodeId, d.NumaNode, d.DevId) for d in devs], pairs)
    ids = [d.Id for d in devs]
    reader = getattr(tp.lib(), "tp_tri_times_m%d_d1" % m)
    si = 3 if shards > 1 else 0
    tp.search_resident(pol, ids[:n], [], 8, iters=3, shard_index=si, shard_count=shards)
    reader((C.c_ulonglong * 16)())
    ms, key, info = tp.search_resident(pol, ids[:n], [], 8, iters=1, shard_index=si, shard_count=shards)
    t = (C.c_ulonglong * 16)()
    reader(t)
    print("%s, 8 of %d: kernel %s, events %.1f us" % (label, n, tp.kernel_name(info["kernel"]), ms * 1e3))
    print("   all blocks: first entry %+.1f, last entry %+.1f, last setup out %+.1f, last items done (thread 0 of a block) %+.1f, last exit %+.1f us relative to block 0's entry" % tuple((t[i] - t[0]) / 1e3 for i in (9, 11, 12, 13, 10)))
    print("   block 0: " + ", ".join("%s +%.1f" % (names[i], (t[i] - t[0]) / 1e3) for i in range(1, 9)), flush=True)
    pol.close()
