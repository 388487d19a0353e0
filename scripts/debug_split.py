import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import topo
from __graft_entry__ import package
from oracle.binding import RefPolicy
tp = package()
name = sys.argv[1] if len(sys.argv) > 1 else "c4"
SY = {"c4": topo.c4_topology, "hives": topo.hives_topology, "missing": topo.missing_links_topology, "cpx64": topo.cpx64_topology}
devs, pairs = SY[name]() if name in SY else topo.fixture_topology(name)
ref = RefPolicy(); ref.InitPairs(devs, pairs)
mk = lambda k: tp.BestEffortPolicy(cuda_device=0, kernel=k)
D = [tp.Device(d.Id, d.NodeId, d.NumaNode, d.DevId) for d in devs]
sp = mk(0); sp.InitWithWeights(D, pairs)
lx = mk(1); lx.InitWithWeights(D, pairs)
ids = [d.Id for d in devs]
gen = topo.SplitMix64(int(sys.argv[2], 0) if len(sys.argv) > 2 else 1)
bad = 0
for t in range(int(sys.argv[3]) if len(sys.argv) > 3 else 400):
    av, rq, k = topo.random_request(gen, ids, kmax=7)
    want = ref.Allocate(av, rq, k)
    if not ref.last["searched"]:
        continue
    a = sp.Allocate(av, rq, k); ia = sp.last
    b = lx.Allocate(av, rq, k); ib = lx.last
    if a != want or b != want or ia["score"] != ref.last["score"]:
        bad += 1
        print("MISMATCH n_av=%d n_req=%d k=%d" % (len(av), len(rq), k))
        print(" ref  ", want, ref.last)
        print(" split", a, ia)
        print(" lex  ", b, ib)
        if bad >= 3:
            break
print("done bad=%d" % bad)
