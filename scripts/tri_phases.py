"""Phase timestamps of block 0 of kernel TRI (needs a -DTP_TRI_TIMING build): TP_ALLOC_LIB=... python scripts/tri_phases.py"""
import os, sys, ctypes as C
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, R); sys.path.insert(0, os.path.join(R, "tests"))
import topo
from __graft_entry__ import package
tp = package()
devs, pairs = topo.fixture_topology("cpx")
pol = tp.BestEffortPolicy(cuda_device=0)
pol.InitWithWeights([tp.Device(d.Id, d.NodeId, d.NumaNode, d.DevId) for d in devs], pairs)
ids = [d.Id for d in devs]
names = ["entry", "setup in", "copies", "weights", "seeds", "table", "setup out", "items done", "finish"]
tp.lib().tp_tri_times_m64_d1((C.c_ulonglong * 16)())  # reset
for n in (24, 40, 63):
    tp.search_resident(pol, ids[:n], [], 8, iters=3)
    tp.lib().tp_tri_times_m64_d1((C.c_ulonglong * 16)())
    ms, key, info = tp.search_resident(pol, ids[:n], [], 8, iters=1)
    t = (C.c_ulonglong * 16)()
    rc = tp.lib().tp_tri_times_m64_d1(t)
    print("   all blocks: first entry %+.1f, last entry %+.1f, last setup out %+.1f, last items done (thread 0 of a block) %+.1f, last exit %+.1f us relative to block 0's entry" % tuple((t[i] - t[0]) / 1e3 for i in (9, 11, 12, 13, 10)))
    base = t[0]
    print("8 of %d: kernel %s, events %.1f us; block 0: " % (n, tp.kernel_name(info["kernel"]), ms * 1e3) +
          ", ".join("%s +%.1f" % (names[i], (t[i] - base) / 1e3) for i in range(1, 9)))
