"""Small mixed workload for compute-sanitizer: every kernel family once, tiny sizes."""
import os, sys
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, R); sys.path.insert(0, os.path.join(R, "tests"))
import topo
from __graft_entry__ import package
from oracle.binding import RefPolicy
tp = package()
devs, pairs = topo.fixture_topology("mi308")
ref = RefPolicy(); ref.InitPairs(devs, pairs)
D = [tp.Device(d.Id, d.NodeId, d.NumaNode, d.DevId) for d in devs]
ids = [d.Id for d in devs]
for kern in (0, 1, 3):
    pol = tp.BestEffortPolicy(cuda_device=0, kernel=kern); pol.InitWithWeights(D, pairs)
    for av, rq, k in ((ids[:14], [], 2), (ids[:18], [ids[3]], 5), (ids[:16], [], 6), (ids[:15], [ids[0]], 8), (ids[:14], [], 9), (ids, [], 20)):
        got = pol.Allocate(av, rq, k); want = ref.Allocate(av, rq, k)
        assert got == want, (kern, k, got, want)
        print("kernel opt", kern, "k", k, "ran", pol.last["kernel"], "ok")
    if kern == 0:
        reqs = [(ids[:10 + i % 12], [], 1 + i % 6) for i in range(24)]
        got, _ = tp.allocate_batch(pol, reqs)
        assert got == [ref.Allocate(*r) for r in reqs]
        print("batch ok")
    pol.close()
print("done")
