"""Kernel time of 8-of-n searches on the C4 topology for a range of n (tuning aid)."""
import os, sys
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, R); sys.path.insert(0, os.path.join(R, "tests"))
import topo
from __graft_entry__ import package
tp = package()
devs, pairs = topo.c4_topology()
ids = [d.Id for d in devs]
for kernel in (0, 2):
    pol = tp.BestEffortPolicy(cuda_device=0, kernel=kernel)
    pol.InitWithWeights([tp.Device(d.Id, d.NodeId, d.NumaNode, d.DevId) for d in devs], pairs)
    for n in [int(a) for a in sys.argv[1:]] or (48, 56, 64, 66, 68, 70, 72, 76, 80):
        tp.search_resident(pol, ids[:n], [], 8, iters=2)
        ms, key, info = tp.search_resident(pol, ids[:n], [], 8, iters=5)
        print("kernel %s 8 of %d: %.4f ms %.4g subsets/s" % (tp.kernel_name(info["kernel"]), n, ms / 5, info["subsets_total"] / (ms / 5 * 1e-3)), flush=True)
    pol.close()
