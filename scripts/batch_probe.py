"""Batch efficiency probe: one 8-of-n request alone (whole grid) against many copies of it in one batch."""
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
gen = topo.SplitMix64(7)
for n in (40, 44, 48, 52, 56):
    av = sorted(ids, key=lambda _: gen.below(1 << 30))[:n]
    av = [i for i in ids if i in set(av)]
    tp.search_resident(pol, av, [], 8, iters=2)
    ms, key, info = tp.search_resident(pol, av, [], 8, iters=5)
    single = info["subsets_total"] / (ms / 5 * 1e-3)
    line = "8 of %d: alone %.4f ms %.3g/s;" % (n, ms / 5, single)
    copies = 592
    for shift in (24, 26, 28, 40):
        pol.set_batch_tuning(share_shift=shift)
        b = tp.Batch(pol, [(av, [], 8)] * copies)
        b.run()
        ts = []
        for _ in range(3):
            t = time.perf_counter(); b.run(); ts.append(time.perf_counter() - t)
        line += " shift %d: %.3g/s" % (shift, copies * info["subsets_total"] / min(ts))
    print(line, flush=True)
