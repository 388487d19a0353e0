"""Pod storm (BASELINE C5) against the item-size knob: python scripts/storm_sweep.py"""
import os, sys, time
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, R); sys.path.insert(0, os.path.join(R, "tests"))
import topo
from __graft_entry__ import package
tp = package()
devs, pairs = topo.fixture_topology("cpx")
ids = [d.Id for d in devs]
gen = topo.SplitMix64(0x5EED0001)
reqs = []
while len(reqs) < 10000:
    p = (0.25, 0.5, 0.75)[gen.below(3)]
    av = [i for i in ids if gen.chance(p)]
    k = (1, 2, 4, 8)[gen.below(4)]
    if len(av) < k:
        continue
    reqs.append((av, [av[gen.below(len(av))]] if gen.chance(0.2) else [], k))
pol = tp.BestEffortPolicy(cuda_device=0)
pol.InitWithWeights([tp.Device(d.Id, d.NodeId, d.NumaNode, d.DevId) for d in devs], pairs)
batch = tp.Batch(pol, reqs)
for cap in (0, 2048, 4096, 8192, 16384, 32768, 65536):
    pol.set_tuning(item_cap=cap)
    batch.run()
    ts = []
    for _ in range(3):
        t = time.perf_counter(); batch.run(); ts.append(time.perf_counter() - t)
    print("cap %d: batch %.2f ms" % (cap, min(ts) * 1e3), flush=True)
