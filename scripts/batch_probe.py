"""Where does a batch's GPU time go?  Single-request rate by pool size against the rate the
same requests reach inside tp_allocate_batch.   python scripts/batch_probe.py"""
import os, sys, time
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, R); sys.path.insert(0, os.path.join(R, "tests"))
import topo
from __graft_entry__ import package
tp = package()
devs, pairs = topo.fixture_topology("cpx")
ids = [d.Id for d in devs]
pol = tp.BestEffortPolicy(cuda_device=0)
pol.InitWithWeights([tp.Device(d.Id, d.NodeId, d.NumaNode, d.DevId) for d in devs], pairs)
gen = topo.SplitMix64(5)
for k in (8, 4):
    for n in (16, 24, 31, 40, 47, 55, 63):
        pool = sorted(ids, key=lambda _: gen.next())[:n]
        tp.search_resident(pol, pool, [], k, iters=1)
        ms, key, info = tp.search_resident(pol, pool, [], k, iters=5)
        single = info["subsets_total"] / (ms / 5 * 1e-3)
        count = max(4, min(2000, int(3e10 / max(info["subsets_total"], 1))))
        batch = tp.Batch(pol, [(pool, [], k)] * count)
        batch.run()
        t = time.perf_counter(); batch.run(); dt = time.perf_counter() - t
        print("k=%d n=%d kernel %d subsets %.3g: single %.3f ms %.3g/s | batch of %d: %.2f ms %.3g/s" % (
            k, n, info["kernel"], info["subsets_total"], ms / 5, single, count, dt * 1e3,
            count * info["subsets_total"] / dt))
