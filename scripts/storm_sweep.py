"""Pod storm (BASELINE C5) against the batch work-plan knobs: python scripts/storm_sweep.py"""
import os, sys, time, threading
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
batch.run()
want, _ = batch.results()


def timed(label):
    batch.run()
    ts = []
    for _ in range(3):
        t = time.perf_counter(); batch.run(); ts.append(time.perf_counter() - t)
    got, _ = batch.results()
    print("%s: batch %.2f ms%s" % (label, min(ts) * 1e3, "" if got == want else "  MISMATCH"), flush=True)


for shift in (24, 25, 26, 27, 28):
    for ipw in (4, 8, 16):
        pol.set_batch_tuning(share_shift=shift, items_per_warp=ipw)
        timed("shift %d items/warp %d" % (shift, ipw))
pol.set_batch_tuning(light_first=True)
timed("defaults, light wave first")
pol.set_batch_tuning()
timed("defaults")
# two batches at once on one policy (separate buffers and streams per call)
b2 = tp.Batch(pol, reqs)
b2.run()
t = time.perf_counter()
th = threading.Thread(target=b2.run)
th.start(); batch.run(); th.join()
dt = time.perf_counter() - t
print("two concurrent batches: %.2f ms for both%s" % (dt * 1e3, "" if b2.results()[0] == want and batch.results()[0] == want else "  MISMATCH"))
