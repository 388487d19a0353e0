"""Kernel TRI against the oracle's BFS on seeded random requests (k' = 6..12, required devices, shards), then
timings of the headline searches with TRI and with SPLIT.   python scripts/tri_check.py [cases]"""
import os, sys, time
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, R); sys.path.insert(0, os.path.join(R, "tests"))
import topo
from __graft_entry__ import package
from oracle.binding import RefPolicy
tp = package()
cases = int(sys.argv[1]) if len(sys.argv) > 1 else 150
SY = {"c4": topo.c4_topology, "hives": topo.hives_topology, "missing": topo.missing_links_topology, "cpx64": topo.cpx64_topology}
def pair(name, kernel=0):
    devs, pairs = SY[name]() if name in SY else topo.fixture_topology(name)
    ref = RefPolicy(); ref.InitPairs(devs, pairs)
    pol = tp.BestEffortPolicy(cuda_device=0, kernel=kernel)
    pol.InitWithWeights([tp.Device(d.Id, d.NodeId, d.NumaNode, d.DevId) for d in devs], pairs)
    return ref, pol, [d.Id for d in devs]
bad = 0
for name in ("cpx", "mi308", "missing", "hives", "c4", "cpx64"):
    ref, pol, ids = pair(name)
    gen = topo.SplitMix64(0x7A1 + len(ids))
    kernels, t0 = {}, time.time()
    for i in range(cases):
        k = 6 + gen.below(7)
        n = min(len(ids), k + 2 + gen.below(30 if k <= 8 else (14 if k <= 10 else 8)))
        pool = sorted(ids, key=lambda _: gen.next())[:n]
        nreq = gen.below(3) if gen.chance(0.4) else 0
        rq = pool[:nreq]
        size = k + nreq
        if len(pool) <= size:
            continue
        want = ref.Allocate(pool, rq, size)
        got = pol.Allocate(pool, rq, size)
        kernels[pol.last["kernel"]] = kernels.get(pol.last["kernel"], 0) + 1
        ok = got == want and pol.last["score"] == ref.last["score"] and pol.last["subsets_scored"] == pol.last["subsets_total"]
        if ok and i % 5 == 0:
            S = 2 + gen.below(4)
            keys = [pol.SearchShard(pool, rq, size, s, S) for s in range(S)]
            ok = pol.Decode(pool, rq, size, min(keys)) == want
        if not ok:
            bad += 1
            print("MISMATCH", name, len(pool), rq, size, got, want, pol.last, ref.last)
            if bad > 5:
                sys.exit(1)
    print(name, "cases", cases, "kernels", kernels, "%.1fs" % (time.time() - t0), flush=True)
    pol.close()
print("mismatches", bad)
for kernel in (0, 2):
    for name, n, size in (("cpx64", 64, 8), ("cpx", 63, 8), ("c4", 80, 8), ("cpx", 47, 8), ("cpx", 40, 8), ("cpx", 24, 8), ("c4", 80, 7), ("cpx", 63, 6), ("cpx", 40, 10)):
        ref, pol, ids = pair(name, kernel)
        tp.search_resident(pol, ids[:n], [], size, iters=2)
        ms, key, info = tp.search_resident(pol, ids[:n], [], size, iters=10)
        want = ref.Allocate(ids[:n], [], size)
        got = pol.Decode(ids[:n], [], size, key)
        print("kernel=%d(%s) %s %d of %d: %.4f ms  %.4g subsets/s  %s" % (kernel, tp.kernel_name(info["kernel"]), name, size, n, ms / 10,
              info["subsets_total"] / (ms / 10 * 1e-3), "ok" if got == want and info["subsets_scored"] == info["subsets_total"] else "WRONG"), flush=True)
        pol.close()
